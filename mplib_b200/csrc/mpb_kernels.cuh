// mpb_kernels.cuh -- sm_100a kernels of the batched state-validity path.
//
//   fk_cull_kernel<PASS>  thread per configuration: FK over the kinematic forest (replaces
//                         PinocchioModelTpl::computeForwardKinematics + getLinkPose +
//                         FCLModelTpl::updateCollisionObjects + AttachedBodyTpl::updatePose,
//                         src/pinocchio_model.cpp:271-312, src/fcl_model.cpp:138-148,
//                         src/attached_body.h:48-54), then for every admitted pair a bounding-sphere
//                         and an OBB-OBB SAT reject; survivors are ballot-compacted into per-bin work
//                         queues.  PASS 0 counts, PASS 1 fills (exact allocation, no overflow).
//   scan_bins_kernel      prefix sums of the bin counts -> queue offsets and 32-item work units.
//   narrow_kernel         lane per work item, warp per 32 items of ONE bin, so every lane walks the
//                         same hull: the vertex loop is a broadcast LDS.128 + 3 FFMA per vertex for 32
//                         items at once.  Hull pool staged in shared memory by one TMA bulk copy
//                         (cp.async.bulk + mbarrier).  MPR (replaces ::fcl::collide -> ccdMPRIntersect on
//                         the pairs of FCLModelTpl::collideFull / PlanningWorldTpl::selfCollide /
//                         collideWithOthers, src/fcl_model.cpp:181-193, src/planning_world.cpp:276-481)
//                         is written as a state machine with a single support call site so divergent
//                         lanes re-converge at the vertex loop.
//   fk_pose_kernel        FK-only API: link poses as [x,y,z,qw,qx,qy,qz].
#pragma once
#include <float.h>

#include "../../include/mpb.h"
#include "mpb_device.cuh"

struct SceneImg {
  const DevJoint *joints;
  const DevObject *objs;
  const DevPair *pairs;
  const float4 *verts;
  const float *linkL;   // [n_links][12] link placement in its joint frame
  const int *linkJoint; // [n_links]
  int n_joints, n_objects, n_pairs, n_links, n_verts, n_bins, state_dim, pair_words;
};

struct WorkQ {
  float *itf;            // [12][cap] relative transform B in A (SoA)
  int *icfg;             // [cap]
  int *ipair;            // [cap]
  unsigned int *bin_count;   // [n_bins]
  unsigned int *bin_cursor;  // [n_bins]
  unsigned int *bin_start;   // [n_bins+1]
  unsigned int *unit_start;  // [n_bins+1]
  unsigned int *totals;      // [0]=items [1]=units [2]=overflow flag
  unsigned long long *counters;  // [0]=items after cull [1]=support evaluations [2]=capped items
  unsigned int cap;
};

#define MPB_EPS FLT_EPSILON
#define MPB_MPR_TOL 1e-6f
#define MPB_MPR_MAX_ITERS 96

// ---------------------------------------------------------------------------------------------- FK
__device__ __forceinline__ Xf joint_motion(const DevJoint &J, float q) {
  Xf m;
  if (J.type == MPB_JOINT_PRISMATIC) {
    m.r[0] = 1.f; m.r[1] = 0.f; m.r[2] = 0.f; m.r[3] = 0.f; m.r[4] = 1.f; m.r[5] = 0.f; m.r[6] = 0.f; m.r[7] = 0.f; m.r[8] = 1.f;
    m.p[0] = J.ax[0] * q; m.p[1] = J.ax[1] * q; m.p[2] = J.ax[2] * q;
    return m;
  }
  float s, c;
  sincosf(q, &s, &c);
  float v = 1.f - c, x = J.ax[0], y = J.ax[1], z = J.ax[2];
  m.r[0] = fmaf(x * x, v, c);     m.r[1] = fmaf(x * y, v, -z * s); m.r[2] = fmaf(x * z, v, y * s);
  m.r[3] = fmaf(x * y, v, z * s); m.r[4] = fmaf(y * y, v, c);      m.r[5] = fmaf(y * z, v, -x * s);
  m.r[6] = fmaf(x * z, v, -y * s); m.r[7] = fmaf(y * z, v, x * s); m.r[8] = fmaf(z * z, v, c);
  m.p[0] = m.p[1] = m.p[2] = 0.f;
  return m;
}

__device__ __forceinline__ Xf load_xf12(const float *p) {
  Xf o;
  const float4 *p4 = reinterpret_cast<const float4 *>(p);
  float4 a = __ldg(p4), b = __ldg(p4 + 1), c = __ldg(p4 + 2);
  o.r[0] = a.x; o.r[1] = a.y; o.r[2] = a.z; o.p[0] = a.w;
  o.r[3] = b.x; o.r[4] = b.y; o.r[5] = b.z; o.p[1] = b.w;
  o.r[6] = c.x; o.r[7] = c.y; o.r[8] = c.z; o.p[2] = c.w;
  return o;
}

// oMi for all joints of one configuration; T[0] = identity (universe).
__device__ __forceinline__ void fk_joints(const SceneImg &img, const float *qrow, Xf *T) {
  T[0].r[0] = 1.f; T[0].r[1] = 0.f; T[0].r[2] = 0.f; T[0].r[3] = 0.f; T[0].r[4] = 1.f; T[0].r[5] = 0.f;
  T[0].r[6] = 0.f; T[0].r[7] = 0.f; T[0].r[8] = 1.f; T[0].p[0] = T[0].p[1] = T[0].p[2] = 0.f;
  for (int j = 0; j < img.n_joints; j++) {
    const DevJoint &J = img.joints[j];
    float q = J.slot >= 0 ? qrow[J.slot] : J.qconst;
    Xf pl = load_xf12(J.P);
    Xf li = xf_mul(pl, joint_motion(J, q));
    T[j + 1] = xf_mul(T[J.parent], li);
  }
}

// ---------------------------------------------------------------------------------------------- cull
// 15-axis SAT between two OBBs; R,t = box B expressed in box A's frame.  Conservative margin `eps`.
__device__ __forceinline__ bool obb_disjoint(const float *R, const float *t, const float *a, const float *b) {
  const float eps = 1e-5f;
  float Q[9];
#pragma unroll
  for (int i = 0; i < 9; i++) Q[i] = fabsf(R[i]) + 1e-6f;
#pragma unroll
  for (int i = 0; i < 3; i++)
    if (fabsf(t[i]) > a[i] + b[0] * Q[3 * i] + b[1] * Q[3 * i + 1] + b[2] * Q[3 * i + 2] + eps) return true;
#pragma unroll
  for (int j = 0; j < 3; j++) {
    float tj = t[0] * R[j] + t[1] * R[3 + j] + t[2] * R[6 + j];
    if (fabsf(tj) > b[j] + a[0] * Q[j] + a[1] * Q[3 + j] + a[2] * Q[6 + j] + eps) return true;
  }
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3;
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const int j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      float lhs = fabsf(t[i2] * R[3 * i1 + j] - t[i1] * R[3 * i2 + j]);
      float rhs = a[i1] * Q[3 * i2 + j] + a[i2] * Q[3 * i1 + j] + b[j1] * Q[3 * i + j2] + b[j2] * Q[3 * i + j1];
      if (lhs > rhs + eps) return true;
    }
  }
  return false;
}

template <int PASS>
__global__ void __launch_bounds__(128) fk_cull_kernel(SceneImg img, const float *__restrict__ q, long long n, WorkQ wq) {
  const unsigned lane = threadIdx.x & 31;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n_round = (n + 31) / 32 * 32;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride) {
    const bool valid = i < n;
    Xf T[MPB_MAX_JOINTS + 1];
    Xf W[MPB_MAX_DYN];
    if (valid) {
      fk_joints(img, q + i * img.state_dim, T);
      for (int o = 0; o < img.n_objects; o++) {
        const DevObject &O = img.objs[o];
        if (O.dyn >= 0) W[O.dyn] = xf_mul(T[O.joint], load_xf12(O.L));
      }
    }
    for (int p = 0; p < img.n_pairs; p++) {
      const DevPair pr = img.pairs[p];
      bool keep = false;
      Xf rel;
      if (valid) {
        const DevObject &A = img.objs[pr.a];
        const DevObject &B = img.objs[pr.b];
        Xf Wa = A.dyn >= 0 ? W[A.dyn] : load_xf12(A.L);
        Xf Wb = B.dyn >= 0 ? W[B.dyn] : load_xf12(B.L);
        float ca[3], cb[3];
        xf_point(Wa, A.bc, ca);
        xf_point(Wb, B.bc, cb);
        float dx = ca[0] - cb[0], dy = ca[1] - cb[1], dz = ca[2] - cb[2];
        float rr = A.br + B.br + 1e-5f;
        if (fmaf(dx, dx, fmaf(dy, dy, dz * dz)) <= rr * rr) {
          rel = xf_inv_mul(Wa, Wb);
          // box B in box A's frame:  Ra^T (rel.R Rb),  Ra^T (rel*ocB - ocA)
          float M[9], R[9], cbA[3], t[3];
#pragma unroll
          for (int r = 0; r < 3; r++)
#pragma unroll
            for (int c = 0; c < 3; c++)
              M[3 * r + c] = fmaf(rel.r[3 * r], B.oR[c], fmaf(rel.r[3 * r + 1], B.oR[3 + c], rel.r[3 * r + 2] * B.oR[6 + c]));
#pragma unroll
          for (int r = 0; r < 3; r++)
#pragma unroll
            for (int c = 0; c < 3; c++)
              R[3 * r + c] = fmaf(A.oR[r], M[c], fmaf(A.oR[3 + r], M[3 + c], A.oR[6 + r] * M[6 + c]));
          xf_point(rel, B.oc, cbA);
          float d0 = cbA[0] - A.oc[0], d1 = cbA[1] - A.oc[1], d2 = cbA[2] - A.oc[2];
#pragma unroll
          for (int r = 0; r < 3; r++) t[r] = fmaf(A.oR[r], d0, fmaf(A.oR[3 + r], d1, A.oR[6 + r] * d2));
          keep = !obb_disjoint(R, t, A.oh, B.oh);
        }
      }
      const unsigned m = __ballot_sync(0xffffffffu, keep);
      if (m == 0) continue;
      if (PASS == 0) {
        if (lane == 0) atomicAdd(&wq.bin_count[pr.bin], (unsigned)__popc(m));
      } else {
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(&wq.bin_cursor[pr.bin], (unsigned)__popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (keep) {
          const unsigned slot = wq.bin_start[pr.bin] + base + __popc(m & ((1u << lane) - 1u));
          if (slot < wq.cap) {
#pragma unroll
            for (int k = 0; k < 9; k++) wq.itf[(size_t)k * wq.cap + slot] = rel.r[k];
#pragma unroll
            for (int k = 0; k < 3; k++) wq.itf[(size_t)(9 + k) * wq.cap + slot] = rel.p[k];
            wq.icfg[slot] = (int)i;
            wq.ipair[slot] = p;
          }
        }
      }
    }
  }
}

__global__ void scan_bins_kernel(WorkQ wq, int n_bins) {
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    unsigned items = 0, units = 0;
    for (int b = 0; b < n_bins; b++) {
      wq.bin_start[b] = items;
      wq.unit_start[b] = units;
      const unsigned c = wq.bin_count[b];
      items += c;
      units += (c + 31u) >> 5;
      wq.bin_cursor[b] = 0;
    }
    wq.bin_start[n_bins] = items;
    wq.unit_start[n_bins] = units;
    wq.totals[0] = items;
    wq.totals[1] = units;
    if (items > wq.cap) wq.totals[2] = 1u;
    atomicAdd(wq.counters, (unsigned long long)items);
  }
}

// ---------------------------------------------------------------------------------------------- supports
struct DevShape {
  int type, voff, vcnt;
  float p0, p1, p2;
};
__device__ __forceinline__ DevShape shape_of(const DevObject &O) {
  DevShape s;
  s.type = O.type; s.voff = O.voff; s.vcnt = O.vcnt; s.p0 = O.prm[0]; s.p1 = O.prm[1]; s.p2 = O.prm[2];
  return s;
}
__device__ __forceinline__ float sgn_eps(float v) { return fabsf(v) < MPB_EPS ? 0.f : (v < 0.f ? -1.f : 1.f); }

// extreme point of the shape along `d` in the shape's own frame ([upstream FCL support* functions])
template <bool SMEM>
__device__ __forceinline__ V3 support_local(const DevShape &s, V3 d, const float4 *__restrict__ verts) {
  if (s.type == MPB_GEOM_CONVEX) {
    const float4 *v = verts + s.voff;
    float best = -FLT_MAX;
    int bi = 0;
#pragma unroll 4
    for (int i = 0; i < s.vcnt; i++) {
      const float4 p = SMEM ? v[i] : __ldg(v + i);
      const float t = fmaf(p.x, d.x, fmaf(p.y, d.y, p.z * d.z));
      if (t > best) { best = t; bi = i; }
    }
    const float4 p = SMEM ? v[bi] : __ldg(v + bi);
    return v3(p.x, p.y, p.z);
  }
  switch (s.type) {
    case MPB_GEOM_SPHERE: {
      float l2 = dot(d, d);
      return (l2 - MPB_EPS > 0.f) ? d * (s.p0 * rsqrtf(l2)) : v3(0.f, 0.f, 0.f);
    }
    case MPB_GEOM_BOX:
      return v3(sgn_eps(d.x) * s.p0, sgn_eps(d.y) * s.p1, sgn_eps(d.z) * s.p2);
    case MPB_GEOM_CAPSULE: {
      V3 n = d * (s.p0 * rsqrtf(dot(d, d)));
      n.z += (d.z > 0.f) ? s.p1 : -s.p1;
      return n;
    }
    case MPB_GEOM_CYLINDER: {
      float zd = sqrtf(fmaf(d.x, d.x, d.y * d.y));
      if (fabsf(zd) < MPB_EPS) return v3(0.f, 0.f, sgn_eps(d.z) * s.p1);
      float rad = s.p0 / zd;
      return v3(rad * d.x, rad * d.y, sgn_eps(d.z) * s.p1);
    }
    default:
      return v3(0.f, 0.f, 0.f);
  }
}

__device__ __forceinline__ V3 rot(const float *R, V3 v) {
  return v3(fmaf(R[0], v.x, fmaf(R[1], v.y, R[2] * v.z)), fmaf(R[3], v.x, fmaf(R[4], v.y, R[5] * v.z)),
            fmaf(R[6], v.x, fmaf(R[7], v.y, R[8] * v.z)));
}
__device__ __forceinline__ V3 rot_t(const float *R, V3 v) {
  return v3(fmaf(R[0], v.x, fmaf(R[3], v.y, R[6] * v.z)), fmaf(R[1], v.x, fmaf(R[4], v.y, R[7] * v.z)),
            fmaf(R[2], v.x, fmaf(R[5], v.y, R[8] * v.z)));
}
__device__ __forceinline__ bool is_zero(float v) { return fabsf(v) < MPB_EPS; }

// ---------------------------------------------------------------------------------------------- MPR
// ccdMPRIntersect [upstream libccd v2.1 mpr.c] in the frame of shape A; B sits at (R, t).
// Phases: 1 = v1, 2 = v2, 3 = v3 search (discoverPortal), 4 = refinePortal.  One support per trip.
template <bool SMEM>
__device__ __forceinline__ bool mpr_intersect(const DevShape &A, V3 cA, const DevShape &B, V3 cB, const float *R, V3 t,
                                              const float4 *__restrict__ verts, bool active, int &iters, bool &capped) {
  V3 v0 = cA - (rot(R, cB) + t);
  if (is_zero(v0.x) && is_zero(v0.y) && is_zero(v0.z)) v0.x += 10.f * MPB_EPS;
  V3 v1 = v0, v2 = v0, v3p = v0;
  V3 dir = normalized(v0 * -1.f);
  int phase = 1;
  bool done = !active, hit = false;
  iters = 0;
  capped = false;
  for (int it = 0; it < MPB_MPR_MAX_ITERS; it++) {
    if (__all_sync(0xffffffffu, done)) break;
    // Minkowski support: sA(dir) - (R sB(R^T(-dir)) + t)
    V3 sa = support_local<SMEM>(A, dir, verts);
    V3 sb = support_local<SMEM>(B, rot_t(R, dir * -1.f), verts);
    V3 v = sa - (rot(R, sb) + t);
    if (done) continue;
    iters++;
    const float dv = dot(v, dir);
    if (phase == 1) {
      v1 = v;
      if (is_zero(dv) || dv < 0.f) { done = true; continue; }
      dir = cross(v0, v1);
      if (is_zero(dot(dir, dir))) { done = true; hit = true; continue; }  // origin on v1 / on the v0-v1 segment
      dir = normalized(dir);
      phase = 2;
    } else if (phase == 2) {
      v2 = v;
      if (is_zero(dv) || dv < 0.f) { done = true; continue; }
      dir = normalized(cross(v1 - v0, v2 - v0));
      if (dot(dir, v0) > 0.f) { V3 tmp = v1; v1 = v2; v2 = tmp; dir = dir * -1.f; }
      phase = 3;
    } else if (phase == 3) {
      v3p = v;
      if (is_zero(dv) || dv < 0.f) { done = true; continue; }
      bool cont = false;
      float d1 = dot(cross(v1, v3p), v0);
      if (d1 < 0.f && !is_zero(d1)) { v2 = v3p; cont = true; }
      if (!cont) {
        float d2 = dot(cross(v3p, v2), v0);
        if (d2 < 0.f && !is_zero(d2)) { v1 = v3p; cont = true; }
      }
      if (cont) {
        dir = normalized(cross(v1 - v0, v2 - v0));
      } else {
        phase = 4;
        dir = normalized(cross(v2 - v1, v3p - v1));
        float de = dot(dir, v1);
        if (is_zero(de) || de > 0.f) { done = true; hit = true; }
      }
    } else {
      // refinePortal: can v4 expand the portal past the origin, and is there still progress?
      if (!(is_zero(dv) || dv > 0.f)) { done = true; continue; }
      float m = fminf(dv - dot(v1, dir), fminf(dv - dot(v2, dir), dv - dot(v3p, dir)));
      if (m <= MPB_MPR_TOL) { done = true; continue; }
      V3 v4v0 = cross(v, v0);
      if (dot(v1, v4v0) > 0.f) {
        if (dot(v2, v4v0) > 0.f) v1 = v; else v3p = v;
      } else {
        if (dot(v3p, v4v0) > 0.f) v2 = v; else v1 = v;
      }
      dir = normalized(cross(v2 - v1, v3p - v1));
      float de = dot(dir, v1);
      if (is_zero(de) || de > 0.f) { done = true; hit = true; }
    }
  }
  if (!done) capped = true;
  return hit;
}

// ---------------------------------------------------------------------------------------------- TMA staging
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
  unsigned ok = 0;
  while (!ok) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  }
}

// ---------------------------------------------------------------------------------------------- narrowphase
template <bool SMEM>
__global__ void __launch_bounds__(256) narrow_kernel(SceneImg img, WorkQ wq, unsigned int *__restrict__ verdict_bits,
                                                     unsigned int *__restrict__ pair_bits) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) unsigned long long bar;
  const float4 *verts = img.verts;
  if (SMEM) {
    float4 *sv = reinterpret_cast<float4 *>(smem_raw);
    const unsigned total = (unsigned)img.n_verts * 16u;
    if (threadIdx.x == 0) {
      mbar_init(&bar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0 && total > 0) {
      mbar_expect_tx(&bar, total);
      for (unsigned off = 0; off < total; off += 32768u) {
        const unsigned nb = min(32768u, total - off);
        tma_bulk_g2s(smem_raw + off, reinterpret_cast<const unsigned char *>(img.verts) + off, nb, &bar);
      }
    }
    if (total > 0) mbar_wait(&bar, 0);
    verts = sv;
  }
  const unsigned lane = threadIdx.x & 31;
  const unsigned warps_total = (gridDim.x * blockDim.x) >> 5;
  const unsigned warp_id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const unsigned n_units = wq.totals[1];
  unsigned long long it_sum = 0, cap_sum = 0;
  for (unsigned u = warp_id; u < n_units; u += warps_total) {
    // bin of this unit: largest b with unit_start[b] <= u
    int lo = 0, hi = img.n_bins;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (wq.unit_start[mid] <= u) lo = mid; else hi = mid;
    }
    const unsigned first = wq.bin_start[lo] + ((u - wq.unit_start[lo]) << 5);
    const unsigned end = wq.bin_start[lo + 1];
    const unsigned idx = first + lane;
    const bool active = idx < end && idx < wq.cap;
    const unsigned ld = active ? idx : first;  // inactive lanes mirror the first item (same hulls)
    float R[9];
#pragma unroll
    for (int k = 0; k < 9; k++) R[k] = wq.itf[(size_t)k * wq.cap + ld];
    V3 t = v3(wq.itf[(size_t)9 * wq.cap + ld], wq.itf[(size_t)10 * wq.cap + ld], wq.itf[(size_t)11 * wq.cap + ld]);
    const int cfg = wq.icfg[ld];
    const int p = wq.ipair[ld];
    const DevPair pr = img.pairs[p];
    const DevObject &OA = img.objs[pr.a];
    const DevObject &OB = img.objs[pr.b];
    const DevShape A = shape_of(OA), B = shape_of(OB);
    int iters;
    bool capped;
    const bool hit = mpr_intersect<SMEM>(A, v3(OA.ctr[0], OA.ctr[1], OA.ctr[2]), B, v3(OB.ctr[0], OB.ctr[1], OB.ctr[2]), R, t,
                                         verts, active, iters, capped);
    if (active) {
      it_sum += (unsigned)iters;
      cap_sum += capped ? 1u : 0u;
      if (hit) {
        atomicOr(&verdict_bits[cfg >> 5], 1u << (cfg & 31));
        if (pair_bits) atomicOr(&pair_bits[(size_t)cfg * img.pair_words + (p >> 5)], 1u << (p & 31));
      }
    }
  }
  // per-warp reduction of the counters
  for (int o = 16; o > 0; o >>= 1) {
    it_sum += __shfl_xor_sync(0xffffffffu, it_sum, o);
    cap_sum += __shfl_xor_sync(0xffffffffu, cap_sum, o);
  }
  if (lane == 0 && it_sum) atomicAdd(&wq.counters[1], it_sum);
  if (lane == 0 && cap_sum) atomicAdd(&wq.counters[2], cap_sum);
}

// ---------------------------------------------------------------------------------------------- FK-only API
__device__ __forceinline__ void mat_to_quat(const float *R, float *q) {  // Eigen branches, (w,x,y,z)
  float t = R[0] + R[4] + R[8];
  if (t > 0.f) {
    t = sqrtf(t + 1.f);
    q[0] = 0.5f * t;
    t = 0.5f / t;
    q[1] = (R[7] - R[5]) * t; q[2] = (R[2] - R[6]) * t; q[3] = (R[3] - R[1]) * t;
  } else {
    int i = 0;
    if (R[4] > R[0]) i = 1;
    if (R[8] > R[4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = sqrtf(R[4 * i] - R[4 * j] - R[4 * k] + 1.f);
    q[1 + i] = 0.5f * t;
    t = 0.5f / t;
    q[0] = (R[3 * k + j] - R[3 * j + k]) * t;
    q[1 + j] = (R[3 * j + i] + R[3 * i + j]) * t;
    q[1 + k] = (R[3 * k + i] + R[3 * i + k]) * t;
  }
}

__global__ void __launch_bounds__(128) fk_pose_kernel(SceneImg img, const float *__restrict__ q, long long n, float *__restrict__ poses) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    Xf T[MPB_MAX_JOINTS + 1];
    fk_joints(img, q + i * img.state_dim, T);
    for (int l = 0; l < img.n_links; l++) {
      Xf w = xf_mul(T[img.linkJoint[l]], load_xf12(img.linkL + 12 * l));
      float *o = poses + ((size_t)i * img.n_links + l) * 7;
      float qt[4];
      mat_to_quat(w.r, qt);
      o[0] = w.p[0]; o[1] = w.p[1]; o[2] = w.p[2]; o[3] = qt[0]; o[4] = qt[1]; o[5] = qt[2]; o[6] = qt[3];
    }
  }
}
