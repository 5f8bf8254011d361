// mpb_ik.cuh -- batched closed-loop inverse kinematics.
//
// Replaces PinocchioModelTpl::computeIKCLIK (src/pinocchio_model.cpp:377-432) for n independent problems: thread per
// problem, double precision (the reference iterates to |log6| < 1e-5 with damping 1e-12; FP64 keeps that convergence
// test meaningful and B200 has the FP64 rate for it).  Per iteration: FK along the chain of the target joint, the 6-D
// error  err = log6(joint_pose^-1 oMi[joint]),  the joint Jacobian in the LOCAL frame of the target joint
// (pinocchio::computeJointJacobian), masked columns zeroed,  v = -J^T (J J^T + damp I)^-1 err,  q += v dt.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

#include "../../include/mpb.h"

#define MPB_IK_MAX_CHAIN 24

struct IkJoint {  // one joint of the chain root -> target joint, model constants in double
  double P[12];   // placement in the parent joint frame (3x4, row stride 4)
  double ax[3];
  double qconst;  // value when slot < 0
  int type;       // MPB_JOINT_*
  int slot;       // index into the problem's state vector, -1: not part of the state
  int masked;     // 1: the column of J is zeroed (the joint keeps its initial value)
  int pad;
};

struct Xd {
  double r[9], p[3];
};
__device__ __forceinline__ Xd xd_mul(const Xd &a, const Xd &b) {
  Xd o;
#pragma unroll
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int j = 0; j < 3; j++) o.r[3 * i + j] = a.r[3 * i] * b.r[j] + a.r[3 * i + 1] * b.r[3 + j] + a.r[3 * i + 2] * b.r[6 + j];
    o.p[i] = a.r[3 * i] * b.p[0] + a.r[3 * i + 1] * b.p[1] + a.r[3 * i + 2] * b.p[2] + a.p[i];
  }
  return o;
}
__device__ __forceinline__ Xd xd_inv_mul(const Xd &a, const Xd &b) {  // a^-1 b
  Xd o;
  const double d0 = b.p[0] - a.p[0], d1 = b.p[1] - a.p[1], d2 = b.p[2] - a.p[2];
#pragma unroll
  for (int i = 0; i < 3; i++) {
#pragma unroll
    for (int j = 0; j < 3; j++) o.r[3 * i + j] = a.r[i] * b.r[j] + a.r[3 + i] * b.r[3 + j] + a.r[6 + i] * b.r[6 + j];
    o.p[i] = a.r[i] * d0 + a.r[3 + i] * d1 + a.r[6 + i] * d2;
  }
  return o;
}
__device__ __forceinline__ Xd xd_from12(const double *m) {
  Xd o;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    o.r[3 * i] = m[4 * i]; o.r[3 * i + 1] = m[4 * i + 1]; o.r[3 * i + 2] = m[4 * i + 2]; o.p[i] = m[4 * i + 3];
  }
  return o;
}
__device__ __forceinline__ Xd xd_joint_motion(int type, const double *ax, double q) {
  Xd m;
#pragma unroll
  for (int k = 0; k < 9; k++) m.r[k] = (k % 4 == 0) ? 1.0 : 0.0;
  m.p[0] = m.p[1] = m.p[2] = 0.0;
  if (type == MPB_JOINT_PRISMATIC) { m.p[0] = ax[0] * q; m.p[1] = ax[1] * q; m.p[2] = ax[2] * q; return m; }
  double s, c;
  sincos(q, &s, &c);
  const double v = 1.0 - c, x = ax[0], y = ax[1], z = ax[2];
  m.r[0] = x * x * v + c;     m.r[1] = x * y * v - z * s; m.r[2] = x * z * v + y * s;
  m.r[3] = x * y * v + z * s; m.r[4] = y * y * v + c;     m.r[5] = y * z * v - x * s;
  m.r[6] = x * z * v - y * s; m.r[7] = y * z * v + x * s; m.r[8] = z * z * v + c;
  return m;
}

// [upstream Pinocchio 2.6 log3 / log6]: motion (v, w) with exp6(v, w) = M
__device__ __forceinline__ void xd_log6(const Xd &M, double *out6) {
  const double *R = M.r;
  const double PI_D = 3.14159265358979323846;
  const double tr = R[0] + R[4] + R[8];
  double theta = tr >= 3.0 ? 0.0 : (tr <= -1.0 ? PI_D : acos((tr - 1.0) * 0.5));
  double w[3];
  if (theta >= PI_D - 1e-2) {
    const double cphi = -(tr - 1.0) * 0.5, beta = theta * theta / (1.0 + cphi);
    const double t0 = (R[0] + cphi) * beta, t1 = (R[4] + cphi) * beta, t2 = (R[8] + cphi) * beta;
    w[0] = (R[7] > R[5] ? 1.0 : -1.0) * (t0 > 0.0 ? sqrt(t0) : 0.0);
    w[1] = (R[2] > R[6] ? 1.0 : -1.0) * (t1 > 0.0 ? sqrt(t1) : 0.0);
    w[2] = (R[3] > R[1] ? 1.0 : -1.0) * (t2 > 0.0 ? sqrt(t2) : 0.0);
  } else {
    const double t = (theta > 1e-8 ? theta / sin(theta) : 1.0) * 0.5;
    w[0] = t * (R[7] - R[5]); w[1] = t * (R[2] - R[6]); w[2] = t * (R[3] - R[1]);
  }
  const double t2 = theta * theta;
  double alpha, beta;
  if (theta < 1e-4) {
    alpha = 1.0 - t2 / 12.0 - t2 * t2 / 720.0;
    beta = 1.0 / 12.0 + t2 / 720.0;
  } else {
    double st, ct;
    sincos(theta, &st, &ct);
    alpha = theta * st / (2.0 * (1.0 - ct));
    beta = 1.0 / t2 - st / (2.0 * theta * (1.0 - ct));
  }
  const double *p = M.p;
  const double wp = w[0] * p[0] + w[1] * p[1] + w[2] * p[2];
  out6[0] = alpha * p[0] - 0.5 * (w[1] * p[2] - w[2] * p[1]) + beta * wp * w[0];
  out6[1] = alpha * p[1] - 0.5 * (w[2] * p[0] - w[0] * p[2]) + beta * wp * w[1];
  out6[2] = alpha * p[2] - 0.5 * (w[0] * p[1] - w[1] * p[0]) + beta * wp * w[2];
  out6[3] = w[0]; out6[4] = w[1]; out6[5] = w[2];
}

// A y = b for a symmetric positive (semi-)definite 6x6 + damping; Gaussian elimination with partial pivoting
__device__ __forceinline__ void solve6(double A[6][6], double *b) {
  for (int c = 0; c < 6; c++) {
    int piv = c;
    double best = fabs(A[c][c]);
    for (int r = c + 1; r < 6; r++)
      if (fabs(A[r][c]) > best) { best = fabs(A[r][c]); piv = r; }
    if (piv != c) {
      for (int k = 0; k < 6; k++) { const double t = A[c][k]; A[c][k] = A[piv][k]; A[piv][k] = t; }
      const double t = b[c]; b[c] = b[piv]; b[piv] = t;
    }
    const double d = A[c][c];
    if (d == 0.0) continue;
    for (int r = c + 1; r < 6; r++) {
      const double f = A[r][c] / d;
      for (int k = c; k < 6; k++) A[r][k] -= f * A[c][k];
      b[r] -= f * b[c];
    }
  }
  for (int c = 5; c >= 0; c--) {
    double s = b[c];
    for (int k = c + 1; k < 6; k++) s -= A[c][k] * b[k];
    b[c] = A[c][c] != 0.0 ? s / A[c][c] : 0.0;
  }
}

__global__ void __launch_bounds__(64) ik_clik_kernel(const IkJoint *__restrict__ chain, int n_chain, const double *__restrict__ link2joint_inv,
                                                     const double *__restrict__ poses, const double *__restrict__ q_init, int D, long long n,
                                                     double eps, int max_iter, double dt, double damp, double *__restrict__ q_out,
                                                     unsigned char *__restrict__ success, double *__restrict__ err_out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  // joint_pose = link_pose * link2joint^-1   (pose = xyz, wxyz)
  Xd target;
  {
    const double *ps = poses + i * 7;
    const double qw = ps[3], qx = ps[4], qy = ps[5], qz = ps[6];
    Xd lp;
    lp.r[0] = 1 - 2 * (qy * qy + qz * qz); lp.r[1] = 2 * (qx * qy - qz * qw);     lp.r[2] = 2 * (qx * qz + qy * qw);
    lp.r[3] = 2 * (qx * qy + qz * qw);     lp.r[4] = 1 - 2 * (qx * qx + qz * qz); lp.r[5] = 2 * (qy * qz - qx * qw);
    lp.r[6] = 2 * (qx * qz - qy * qw);     lp.r[7] = 2 * (qy * qz + qx * qw);     lp.r[8] = 1 - 2 * (qx * qx + qy * qy);
    lp.p[0] = ps[0]; lp.p[1] = ps[1]; lp.p[2] = ps[2];
    target = xd_mul(lp, xd_from12(link2joint_inv));
  }
  double q[MPB_IK_MAX_CHAIN];
  for (int c = 0; c < n_chain; c++) q[c] = chain[c].slot >= 0 ? q_init[i * D + chain[c].slot] : chain[c].qconst;
  Xd T[MPB_IK_MAX_CHAIN];
  double err[6];
  bool ok = false;
  for (int it = 0;; it++) {
    Xd cur;
    for (int c = 0; c < n_chain; c++) {
      const Xd li = xd_mul(xd_from12(chain[c].P), xd_joint_motion(chain[c].type, chain[c].ax, q[c]));
      cur = c == 0 ? li : xd_mul(cur, li);
      T[c] = cur;
    }
    xd_log6(xd_inv_mul(target, cur), err);
    double e2 = 0;
    for (int k = 0; k < 6; k++) e2 += err[k] * err[k];
    if (sqrt(e2) < eps) { ok = true; break; }
    if (it >= max_iter) break;
    // joint Jacobian in the local frame of the target joint; columns of masked / constant joints are zero
    double J[6][MPB_IK_MAX_CHAIN];
    double A[6][6];
    for (int r = 0; r < 6; r++)
      for (int k = 0; k < 6; k++) A[r][k] = r == k ? damp : 0.0;
    for (int c = 0; c < n_chain; c++) {
      double col[6] = {0, 0, 0, 0, 0, 0};
      if (chain[c].slot >= 0 && !chain[c].masked) {
        const Xd X = xd_inv_mul(cur, T[c]);  // joint c seen from the target joint
        const double *a = chain[c].ax;
        const double ra[3] = {X.r[0] * a[0] + X.r[1] * a[1] + X.r[2] * a[2], X.r[3] * a[0] + X.r[4] * a[1] + X.r[5] * a[2],
                              X.r[6] * a[0] + X.r[7] * a[1] + X.r[8] * a[2]};
        if (chain[c].type == MPB_JOINT_PRISMATIC) { col[0] = ra[0]; col[1] = ra[1]; col[2] = ra[2]; }
        else {
          col[0] = X.p[1] * ra[2] - X.p[2] * ra[1]; col[1] = X.p[2] * ra[0] - X.p[0] * ra[2]; col[2] = X.p[0] * ra[1] - X.p[1] * ra[0];
          col[3] = ra[0]; col[4] = ra[1]; col[5] = ra[2];
        }
        for (int r = 0; r < 6; r++)
          for (int k = 0; k < 6; k++) A[r][k] += col[r] * col[k];
      }
      for (int r = 0; r < 6; r++) J[r][c] = col[r];
    }
    double y[6];
    for (int k = 0; k < 6; k++) y[k] = err[k];
    solve6(A, y);
    for (int c = 0; c < n_chain; c++) {
      double v = 0;
      for (int r = 0; r < 6; r++) v += J[r][c] * y[r];
      q[c] -= v * dt;
    }
  }
  for (int s = 0; s < D; s++) q_out[i * D + s] = q_init[i * D + s];
  for (int c = 0; c < n_chain; c++)
    if (chain[c].slot >= 0) {
      double v = q[c];
      if (chain[c].type == MPB_JOINT_CONTINUOUS) v = atan2(sin(v), cos(v));  // qposPinocchio2User: (cos, sin) -> angle
      q_out[i * D + chain[c].slot] = v;
    }
  success[i] = ok ? 1 : 0;
  for (int k = 0; k < 6; k++) err_out[i * 6 + k] = err[k];
}

// PinocchioModelTpl::computeFullJacobian + getLinkJacobian (src/pinocchio_model.cpp:335-360): the 6 x state_dim Jacobian
// of `link`'s parent joint for n states, rows (linear, angular).  local == 0: Pinocchio's WORLD frame (spatial velocity
// at the world origin; column of a revolute joint = (p x a, a) with p the joint origin and a its axis in the world);
// local != 0: expressed in the link frame (link2world.toActionMatrixInverse() * J).  Columns of joints off the chain
// are zero.  Also writes the link's world placement as a 3x4 row-major matrix.
__global__ void __launch_bounds__(64) link_jacobian_kernel(const IkJoint *__restrict__ chain, int n_chain, const double *__restrict__ link2joint,
                                                           const double *__restrict__ q, int D, long long n, int local,
                                                           double *__restrict__ jac, double *__restrict__ link_T) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Xd T[MPB_IK_MAX_CHAIN];
  Xd cur;
#pragma unroll
  for (int k = 0; k < 9; k++) cur.r[k] = (k % 4 == 0) ? 1.0 : 0.0;
  cur.p[0] = cur.p[1] = cur.p[2] = 0.0;
  for (int c = 0; c < n_chain; c++) {
    const double qc = chain[c].slot >= 0 ? q[i * D + chain[c].slot] : chain[c].qconst;
    const Xd li = xd_mul(xd_from12(chain[c].P), xd_joint_motion(chain[c].type, chain[c].ax, qc));
    cur = c == 0 ? li : xd_mul(cur, li);
    T[c] = cur;
  }
  const Xd lw = xd_mul(cur, xd_from12(link2joint));
  double *J = jac + i * 6 * D;
  for (int k = 0; k < 6 * D; k++) J[k] = 0.0;
  for (int c = 0; c < n_chain; c++) {
    if (chain[c].slot < 0) continue;
    const Xd X = local ? xd_inv_mul(lw, T[c]) : T[c];
    const double *a = chain[c].ax;
    const double ra[3] = {X.r[0] * a[0] + X.r[1] * a[1] + X.r[2] * a[2], X.r[3] * a[0] + X.r[4] * a[1] + X.r[5] * a[2],
                          X.r[6] * a[0] + X.r[7] * a[1] + X.r[8] * a[2]};
    double col[6] = {0, 0, 0, 0, 0, 0};
    if (chain[c].type == MPB_JOINT_PRISMATIC) { col[0] = ra[0]; col[1] = ra[1]; col[2] = ra[2]; }
    else {
      col[0] = X.p[1] * ra[2] - X.p[2] * ra[1]; col[1] = X.p[2] * ra[0] - X.p[0] * ra[2]; col[2] = X.p[0] * ra[1] - X.p[1] * ra[0];
      col[3] = ra[0]; col[4] = ra[1]; col[5] = ra[2];
    }
    for (int r = 0; r < 6; r++) J[r * D + chain[c].slot] = col[r];
  }
  double *o = link_T + i * 12;
  for (int r = 0; r < 3; r++) { o[4 * r] = lw.r[3 * r]; o[4 * r + 1] = lw.r[3 * r + 1]; o[4 * r + 2] = lw.r[3 * r + 2]; o[4 * r + 3] = lw.p[r]; }
}
