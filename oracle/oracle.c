/*
 * oracle.c -- CPU restatement (double precision, plain C) of the reference's per-state algorithm
 * for the hot path.  TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs; never by the product path.
 *
 * PARITY UNPINNED: the reference cannot be built or imported in this environment (Pinocchio,
 * FCL, libccd, octomap, OMPL, Eigen, Boost, assimp, urdfdom are all absent) and its own tests
 * hold no golden vectors (tests/test_basic.py:39-50 has no assertions).  MPlib-side logic below is
 * restated from /root/reference and cited file:line.  Third-party arithmetic is restated from the
 * published algorithms of the pinned versions (docker/Dockerfile:17-81): Pinocchio v2.6.21
 * forwardKinematics, FCL 0.7.0 narrowphase dispatch (GJKSolver_libccd), libccd v2.1 mpr.c, octomap
 * v1.9.8 -- marked [upstream] where they appear.  The oracle is pinned by hand-derived FK known
 * answers, the reference's stated example outcomes (examples/detect_collision.py:22-65), analytic
 * primitive cases and an independent exhaustive SAT (orc_sat_*), see tests/.
 *
 * Per-state order follows the reference: scatter state -> FK -> link pose as 7-vector
 * -> matrix -> object transforms -> every admitted pair -> flags.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#include "../include/mpb.h"

#define ORC_EPS DBL_EPSILON /* libccd CCD_EPS for double builds [upstream] */
#define ORC_MPR_TOL 1e-6    /* GJKSolver_libccd::collision_tolerance [upstream] */
#define ORC_MAX_ITERS 500   /* GJKSolver_libccd::max_collision_iterations [upstream] */

typedef struct { double R[9]; double p[3]; } xf_t; /* row-major rotation + translation */
typedef struct { double v[3]; } v3_t;

/* ------------------------------------------------------------------ small algebra */
static inline double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static inline void cross3(double *o, const double *a, const double *b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  o[0] = x; o[1] = y; o[2] = z;
}
static inline void sub3(double *o, const double *a, const double *b) { o[0] = a[0] - b[0]; o[1] = a[1] - b[1]; o[2] = a[2] - b[2]; }
static inline void add3(double *o, const double *a, const double *b) { o[0] = a[0] + b[0]; o[1] = a[1] + b[1]; o[2] = a[2] + b[2]; }
static inline void scale3(double *a, double s) { a[0] *= s; a[1] *= s; a[2] *= s; }
static inline void copy3(double *o, const double *a) { o[0] = a[0]; o[1] = a[1]; o[2] = a[2]; }
static inline double len2_3(const double *a) { return dot3(a, a); }
static inline void normalize3(double *a) { double d = 1.0 / sqrt(len2_3(a)); scale3(a, d); } /* ccdVec3Normalize */
static inline int is_zero(double v) { return fabs(v) < ORC_EPS; }                                /* ccdIsZero */
static inline int eq_real(double a, double b) {                                                  /* ccdEq */
  double ab = fabs(a - b);
  if (ab < ORC_EPS) return 1;
  a = fabs(a); b = fabs(b);
  return (b > a) ? (ab < ORC_EPS * b) : (ab < ORC_EPS * a);
}
static inline int eq3(const double *a, const double *b) { return eq_real(a[0], b[0]) && eq_real(a[1], b[1]) && eq_real(a[2], b[2]); }
static inline double sign_real(double v) { return is_zero(v) ? 0.0 : (v < 0 ? -1.0 : 1.0); }   /* ccdSign */

static void xf_from12(xf_t *o, const double *m) {
  for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) o->R[3 * r + c] = m[4 * r + c]; o->p[r] = m[4 * r + 3]; }
}
static void xf_identity(xf_t *o) { memset(o, 0, sizeof(*o)); o->R[0] = o->R[4] = o->R[8] = 1.0; }
static void xf_mul(xf_t *o, const xf_t *a, const xf_t *b) {
  xf_t t;
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) t.R[3 * r + c] = a->R[3 * r] * b->R[c] + a->R[3 * r + 1] * b->R[3 + c] + a->R[3 * r + 2] * b->R[6 + c];
    t.p[r] = a->R[3 * r] * b->p[0] + a->R[3 * r + 1] * b->p[1] + a->R[3 * r + 2] * b->p[2] + a->p[r];
  }
  *o = t;
}
static inline void xf_rot(double *o, const xf_t *t, const double *v) { /* o = R v */
  double x = t->R[0] * v[0] + t->R[1] * v[1] + t->R[2] * v[2];
  double y = t->R[3] * v[0] + t->R[4] * v[1] + t->R[5] * v[2];
  double z = t->R[6] * v[0] + t->R[7] * v[1] + t->R[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}
static inline void xf_rot_t(double *o, const xf_t *t, const double *v) { /* o = R^T v */
  double x = t->R[0] * v[0] + t->R[3] * v[1] + t->R[6] * v[2];
  double y = t->R[1] * v[0] + t->R[4] * v[1] + t->R[7] * v[2];
  double z = t->R[2] * v[0] + t->R[5] * v[1] + t->R[8] * v[2];
  o[0] = x; o[1] = y; o[2] = z;
}
static inline void xf_apply(double *o, const xf_t *t, const double *v) { xf_rot(o, t, v); add3(o, o, t->p); }
static inline void xf_apply_inv(double *o, const xf_t *t, const double *v) { double d[3]; sub3(d, v, t->p); xf_rot_t(o, t, d); }

/* Eigen Quaternion(Matrix3) -> (w,x,y,z); used by getLinkPose, src/pinocchio_model.cpp:276-312 */
static void mat_to_quat(const double *R, double *q) {
  double t = R[0] + R[4] + R[8];
  if (t > 0) {
    t = sqrt(t + 1.0);
    q[0] = 0.5 * t;
    t = 0.5 / t;
    q[1] = (R[7] - R[5]) * t; q[2] = (R[2] - R[6]) * t; q[3] = (R[3] - R[1]) * t;
  } else {
    int i = 0;
    if (R[4] > R[0]) i = 1;
    if (R[8] > R[4 * i]) i = 2;
    int j = (i + 1) % 3, k = (j + 1) % 3;
    t = sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
    q[1 + i] = 0.5 * t;
    t = 0.5 / t;
    q[0] = (R[3 * k + j] - R[3 * j + k]) * t;
    q[1 + j] = (R[3 * j + i] + R[3 * i + j]) * t;
    q[1 + k] = (R[3 * k + i] + R[3 * i + k]) * t;
  }
}
/* Eigen Quaternion::toRotationMatrix; used by setQpos, src/articulated_model.cpp:118-125 */
static void quat_to_mat(const double *q, double *R) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  double twx = tx * w, twy = ty * w, twz = tz * w, txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}

/* ------------------------------------------------------------------ forward kinematics */
/* Joint motion J(q): [upstream Pinocchio] revolute = rotation about the axis by q (Rodrigues),
 * prismatic = translation q*axis; continuous joints go through (cos q, sin q)
 * (src/pinocchio_model.cpp:511-515) which is the same rotation. */
static void joint_motion(xf_t *o, int type, const double *ax, double q) {
  xf_identity(o);
  if (type == MPB_JOINT_PRISMATIC) { o->p[0] = ax[0] * q; o->p[1] = ax[1] * q; o->p[2] = ax[2] * q; return; }
  double c = cos(q), s = sin(q), v = 1.0 - c;
  double x = ax[0], y = ax[1], z = ax[2];
  o->R[0] = x * x * v + c;     o->R[1] = x * y * v - z * s; o->R[2] = x * z * v + y * s;
  o->R[3] = x * y * v + z * s; o->R[4] = y * y * v + c;     o->R[5] = y * z * v - x * s;
  o->R[6] = x * z * v - y * s; o->R[7] = y * z * v + x * s; o->R[8] = z * z * v + c;
}

/* oMi for every model joint (index 0 = universe).  [upstream] pinocchio::forwardKinematics:
 * liMi = placement * J(q); oMi = oMi[parent] * liMi.   Called from src/pinocchio_model.cpp:271-274. */
static void fk_joints(const mpb_scene_desc *d, const double *qfull, xf_t *oMi) {
  xf_identity(&oMi[0]);
  for (int j = 0; j < d->n_joints; j++) {
    xf_t pl, jm, li;
    xf_from12(&pl, d->joint_placement + 12 * j);
    joint_motion(&jm, d->joint_type[j], d->joint_axis + 3 * j, qfull[d->joint_qidx[j]]);
    xf_mul(&li, &pl, &jm);
    xf_mul(&oMi[j + 1], &oMi[d->joint_parent[j]], &li);
  }
}

/* getLinkPose (src/pinocchio_model.cpp:276-312): oMf = oMi[parent] * placement -> [p, quat] */
static void link_pose7(const mpb_scene_desc *d, const xf_t *oMi, int link, double *pose7) {
  xf_t pl, w;
  xf_from12(&pl, d->link_placement + 12 * link);
  xf_mul(&w, &oMi[d->link_joint[link]], &pl);
  pose7[0] = w.p[0]; pose7[1] = w.p[1]; pose7[2] = w.p[2];
  mat_to_quat(w.R, pose7 + 3);
}

/* scatter a state into q_full: src/articulated_model.cpp:108-113 + src/planning_world.cpp:249-262 */
static void scatter_state(const mpb_scene_desc *d, const double *state, double *qfull) {
  memcpy(qfull, d->q_base, sizeof(double) * (size_t)d->nq_full);
  for (int k = 0; k < d->state_dim; k++) qfull[d->state_map[k]] = state[k];
}

/* Object transforms: link pose -> 7-vector -> matrix (src/articulated_model.cpp:116-125), then
 * T_obj = T_link * origin2link (src/fcl_model.cpp:138-148); attached bodies use the same product
 * with their attach pose (src/attached_body.h:48-54).  World-fixed objects keep obj_local. */
static void object_transforms(const mpb_scene_desc *d, const xf_t *oMi, xf_t *objT, double *link_pose_out /*nullable [nl][7]*/) {
  xf_t *linkT = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_links > 0 ? d->n_links : 1));
  for (int l = 0; l < d->n_links; l++) {
    double p7[7];
    link_pose7(d, oMi, l, p7);
    if (link_pose_out) memcpy(link_pose_out + 7 * l, p7, sizeof(p7));
    quat_to_mat(p7 + 3, linkT[l].R);
    copy3(linkT[l].p, p7);
  }
  for (int o = 0; o < d->n_objects; o++) {
    xf_t loc;
    xf_from12(&loc, d->obj_local + 12 * o);
    if (d->obj_link[o] >= 0) xf_mul(&objT[o], &linkT[d->obj_link[o]], &loc);
    else objT[o] = loc;
  }
  free(linkT);
}

/* ------------------------------------------------------------------ shapes + libccd-style supports */
typedef struct {
  int type;          /* MPB_GEOM_* or ORC_TRI */
  xf_t T;            /* world pose */
  double a, b, c;    /* sphere r | box half sides | capsule/cylinder r, half height */
  const double *verts; int nverts; /* convex */
  const int *nbr;                  /* convex, accelerated leg only: vertex adjacency (CSR: off[nverts + 1], then neighbours) */
  double center_local[3];          /* convex interior point (vertex mean), fcl::Convex ctor [upstream] */
  double tri[9];                   /* triangle vertices in the frame T */
} shape_t;
#define ORC_TRI 100

static void shape_from_object(const mpb_scene_desc *d, int o, const xf_t *T, shape_t *s) {
  memset(s, 0, sizeof(*s));
  s->type = d->obj_type[o];
  s->T = *T;
  const double *p = d->obj_param + 4 * o;
  switch (s->type) {
    case MPB_GEOM_SPHERE: s->a = p[0]; break;
    case MPB_GEOM_BOX: s->a = 0.5 * p[0]; s->b = 0.5 * p[1]; s->c = 0.5 * p[2]; break; /* ccd_box_t dim = side/2 [upstream] */
    case MPB_GEOM_CAPSULE:
    case MPB_GEOM_CYLINDER: s->a = p[0]; s->b = 0.5 * p[1]; break;                     /* height = lz/2 [upstream] */
    case MPB_GEOM_CONVEX: {
      int h = d->obj_geom[o];
      int v0 = d->hull_vert_off[h], v1 = d->hull_vert_off[h + 1];
      s->verts = d->hull_verts + 3 * v0; s->nverts = v1 - v0;
      double c[3] = {0, 0, 0};
      for (int i = 0; i < s->nverts; i++) add3(c, c, s->verts + 3 * i);
      scale3(c, 1.0 / (double)s->nverts);
      copy3(s->center_local, c);
    } break;
    default: break;
  }
}

/* [upstream FCL 0.7.0 gjk_libccd-inl.h] support{Box,Sphere,Cap,Cyl,Convex,Triangle}: direction into
 * the shape frame, extreme point there, back to the world.  Convex: argmax over the vertices (first
 * maximum wins; FCL hill-climbs on hulls with >32 vertices, same point except on exact ties). */
static void shape_support(const shape_t *s, const double *dir_w, double *out_w) {
  double dir[3], v[3];
  xf_rot_t(dir, &s->T, dir_w);
  switch (s->type) {
    case MPB_GEOM_BOX:
      v[0] = sign_real(dir[0]) * s->a; v[1] = sign_real(dir[1]) * s->b; v[2] = sign_real(dir[2]) * s->c;
      break;
    case MPB_GEOM_SPHERE: {
      double len = len2_3(dir);
      if (len - ORC_EPS > 0) { copy3(v, dir); scale3(v, s->a / sqrt(len)); } else { v[0] = v[1] = v[2] = 0; }
    } break;
    case MPB_GEOM_CAPSULE: {
      double n[3]; copy3(n, dir); normalize3(n); scale3(n, s->a);
      v[0] = n[0]; v[1] = n[1]; v[2] = n[2] + (dir[2] > 0 ? s->b : -s->b);
    } break;
    case MPB_GEOM_CYLINDER: {
      double zdist = sqrt(dir[0] * dir[0] + dir[1] * dir[1]);
      if (is_zero(zdist)) { v[0] = 0; v[1] = 0; v[2] = sign_real(dir[2]) * s->b; }
      else { double rad = s->a / zdist; v[0] = rad * dir[0]; v[1] = rad * dir[1]; v[2] = sign_real(dir[2]) * s->b; }
    } break;
    case MPB_GEOM_CONVEX: {
      int best = 0; double bd = dot3(s->verts, dir);
      if (s->nbr && s->nverts > 32) {
        /* [upstream FCL 0.7.0 geometry/shape/convex-inl.h] Convex::findExtremeVertex on hulls with more than 32 vertices:
         * start at vertex 0, walk to any not-yet-visited neighbour that is at least as extreme, until none is.  On a
         * convex hull this ends on a global maximum (same vertex as the scan above except on exact ties). */
        static __thread unsigned *stamp = NULL; static __thread int stamp_n = 0; static __thread unsigned cur = 0;
        if (stamp_n < s->nverts) { free(stamp); stamp_n = s->nverts; stamp = (unsigned *)calloc((size_t)stamp_n, sizeof(unsigned)); cur = 0; }
        if (++cur == 0) { memset(stamp, 0, sizeof(unsigned) * (size_t)stamp_n); cur = 1; }
        stamp[0] = cur;
        int keep = 1;
        while (keep) {
          keep = 0;
          const int n0 = s->nbr[best], n1 = s->nbr[best + 1];
          for (int k = n0; k < n1; k++) {
            const int j = s->nbr[s->nverts + 1 + k];
            if (stamp[j] == cur) continue;
            stamp[j] = cur;
            const double t = dot3(s->verts + 3 * j, dir);
            if (t >= bd) { keep = 1; best = j; bd = t; }
          }
        }
      } else
        for (int i = 1; i < s->nverts; i++) { double t = dot3(s->verts + 3 * i, dir); if (t > bd) { bd = t; best = i; } }
      copy3(v, s->verts + 3 * best);
    } break;
    case ORC_TRI: {
      int best = 0; double bd = dot3(s->tri, dir);
      for (int i = 1; i < 3; i++) { double t = dot3(s->tri + 3 * i, dir); if (t > bd) { bd = t; best = i; } }
      copy3(v, s->tri + 3 * best);
    } break;
    default: v[0] = v[1] = v[2] = 0; break;
  }
  xf_apply(out_w, &s->T, v);
}
/* [upstream] centerShape / centerConvex / centerTriangle */
static void shape_center(const shape_t *s, double *c) {
  if (s->type == MPB_GEOM_CONVEX) xf_apply(c, &s->T, s->center_local);
  else if (s->type == ORC_TRI) {
    double m[3] = {(s->tri[0] + s->tri[3] + s->tri[6]) / 3, (s->tri[1] + s->tri[4] + s->tri[7]) / 3, (s->tri[2] + s->tri[5] + s->tri[8]) / 3};
    xf_apply(c, &s->T, m);
  } else copy3(c, s->T.p);
}
/* [upstream libccd support.c] __ccdSupport: v = support1(dir) - support2(-dir) */
static void mink_support(const shape_t *a, const shape_t *b, const double *dir, double *v) {
  double p1[3], p2[3], nd[3] = {-dir[0], -dir[1], -dir[2]};
  shape_support(a, dir, p1);
  shape_support(b, nd, p2);
  sub3(v, p1, p2);
}

/* ------------------------------------------------------------------ MPR  [upstream libccd v2.1 src/mpr.c] */
typedef struct { double v[4][3]; int size; } portal_t;
typedef struct { long iters; } orc_counter_t;
static __thread orc_counter_t g_cnt;

static void portal_dir(const portal_t *p, double *dir) {
  double v2v1[3], v3v1[3];
  sub3(v2v1, p->v[2], p->v[1]);
  sub3(v3v1, p->v[3], p->v[1]);
  cross3(dir, v2v1, v3v1);
  normalize3(dir);
}
static int portal_encapsules_origin(const portal_t *p, const double *dir) { double d = dot3(dir, p->v[1]); return is_zero(d) || d > 0; }
static int portal_can_encapsule_origin(const double *v4, const double *dir) { double d = dot3(v4, dir); return is_zero(d) || d > 0; }
static int portal_reach_tolerance(const portal_t *p, const double *v4, const double *dir, double tol) {
  double dv1 = dot3(p->v[1], dir), dv2 = dot3(p->v[2], dir), dv3 = dot3(p->v[3], dir), dv4 = dot3(v4, dir);
  double d1 = dv4 - dv1, d2 = dv4 - dv2, d3 = dv4 - dv3;
  double m = d1 < d2 ? d1 : d2; m = m < d3 ? m : d3;
  return eq_real(m, tol) || m < tol;
}
static void expand_portal(portal_t *p, const double *v4) {
  double v4v0[3];
  cross3(v4v0, v4, p->v[0]);
  double d = dot3(p->v[1], v4v0);
  if (d > 0) {
    d = dot3(p->v[2], v4v0);
    if (d > 0) copy3(p->v[1], v4); else copy3(p->v[3], v4);
  } else {
    d = dot3(p->v[3], v4v0);
    if (d > 0) copy3(p->v[2], v4); else copy3(p->v[1], v4);
  }
}
/* discoverPortal: -1 no intersection, 0 portal found, 1 origin on v1, 2 origin on v0-v1 segment */
static int discover_portal(const shape_t *A, const shape_t *B, portal_t *p) {
  double dir[3], va[3], vb[3], c1[3], c2[3], dot;
  static const double origin[3] = {0, 0, 0};
  shape_center(A, c1); shape_center(B, c2);
  sub3(p->v[0], c1, c2); /* findOrigin */
  p->size = 1;
  if (eq3(p->v[0], origin)) p->v[0][0] += ORC_EPS * 10.0;
  copy3(dir, p->v[0]); scale3(dir, -1.0); normalize3(dir);
  mink_support(A, B, dir, p->v[1]); g_cnt.iters++;
  p->size = 2;
  dot = dot3(p->v[1], dir);
  if (is_zero(dot) || dot < 0) return -1;
  cross3(dir, p->v[0], p->v[1]);
  if (is_zero(len2_3(dir))) return eq3(p->v[1], origin) ? 1 : 2;
  normalize3(dir);
  mink_support(A, B, dir, p->v[2]); g_cnt.iters++;
  dot = dot3(p->v[2], dir);
  if (is_zero(dot) || dot < 0) return -1;
  p->size = 3;
  sub3(va, p->v[1], p->v[0]); sub3(vb, p->v[2], p->v[0]);
  cross3(dir, va, vb); normalize3(dir);
  dot = dot3(dir, p->v[0]);
  if (dot > 0) { double t[3]; copy3(t, p->v[1]); copy3(p->v[1], p->v[2]); copy3(p->v[2], t); scale3(dir, -1.0); }
  while (p->size < 4) {
    mink_support(A, B, dir, p->v[3]); g_cnt.iters++;
    dot = dot3(p->v[3], dir);
    if (is_zero(dot) || dot < 0) return -1;
    int cont = 0;
    cross3(va, p->v[1], p->v[3]);
    dot = dot3(va, p->v[0]);
    if (dot < 0 && !is_zero(dot)) { copy3(p->v[2], p->v[3]); cont = 1; }
    if (!cont) {
      cross3(va, p->v[3], p->v[2]);
      dot = dot3(va, p->v[0]);
      if (dot < 0 && !is_zero(dot)) { copy3(p->v[1], p->v[3]); cont = 1; }
    }
    if (cont) {
      sub3(va, p->v[1], p->v[0]); sub3(vb, p->v[2], p->v[0]);
      cross3(dir, va, vb); normalize3(dir);
    } else p->size = 4;
  }
  return 0;
}
static int refine_portal(const shape_t *A, const shape_t *B, portal_t *p) {
  double dir[3], v4[3];
  for (;;) {
    portal_dir(p, dir);
    if (portal_encapsules_origin(p, dir)) return 0;
    mink_support(A, B, dir, v4); g_cnt.iters++;
    if (!portal_can_encapsule_origin(v4, dir) || portal_reach_tolerance(p, v4, dir, ORC_MPR_TOL)) return -1;
    expand_portal(p, v4);
  }
}
/* ccdMPRIntersect */
static int mpr_intersect(const shape_t *A, const shape_t *B) {
  portal_t p;
  int res = discover_portal(A, B, &p);
  if (res < 0) return 0;
  if (res > 0) return 1;
  return refine_portal(A, B, &p) == 0 ? 1 : 0;
}

/* closest point on triangle to origin -> squared distance, witness (libccd ccdVec3PointTriDist2 role) */
static double point_seg_dist2(const double *P, const double *a, const double *b, double *w) {
  double d[3], ap[3]; sub3(d, b, a); sub3(ap, P, a);
  double t = dot3(ap, d), dd = len2_3(d);
  if (dd <= 0) t = 0; else { t /= dd; if (t < 0) t = 0; if (t > 1) t = 1; }
  double q[3] = {a[0] + t * d[0], a[1] + t * d[1], a[2] + t * d[2]};
  if (w) copy3(w, q);
  double r[3]; sub3(r, P, q); return len2_3(r);
}
static double point_tri_dist2(const double *P, const double *a, const double *b, const double *c, double *w) {
  /* Ericson, Real-Time Collision Detection 5.1.5 */
  double ab[3], ac[3], ap[3]; sub3(ab, b, a); sub3(ac, c, a); sub3(ap, P, a);
  double d1 = dot3(ab, ap), d2 = dot3(ac, ap), q[3];
  if (d1 <= 0 && d2 <= 0) { copy3(q, a); goto done; }
  { double bp[3]; sub3(bp, P, b);
    double d3 = dot3(ab, bp), d4 = dot3(ac, bp);
    if (d3 >= 0 && d4 <= d3) { copy3(q, b); goto done; }
    double vc = d1 * d4 - d3 * d2;
    if (vc <= 0 && d1 >= 0 && d3 <= 0) { double v = d1 / (d1 - d3); q[0] = a[0] + v * ab[0]; q[1] = a[1] + v * ab[1]; q[2] = a[2] + v * ab[2]; goto done; }
    double cp[3]; sub3(cp, P, c);
    double d5 = dot3(ab, cp), d6 = dot3(ac, cp);
    if (d6 >= 0 && d5 <= d6) { copy3(q, c); goto done; }
    double vb = d5 * d2 - d1 * d6;
    if (vb <= 0 && d2 >= 0 && d6 <= 0) { double w2 = d2 / (d2 - d6); q[0] = a[0] + w2 * ac[0]; q[1] = a[1] + w2 * ac[1]; q[2] = a[2] + w2 * ac[2]; goto done; }
    double va = d3 * d6 - d5 * d4;
    if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) {
      double w2 = (d4 - d3) / ((d4 - d3) + (d5 - d6));
      q[0] = b[0] + w2 * (c[0] - b[0]); q[1] = b[1] + w2 * (c[1] - b[1]); q[2] = b[2] + w2 * (c[2] - b[2]); goto done;
    }
    double den = 1.0 / (va + vb + vc), v = vb * den, w2 = vc * den;
    q[0] = a[0] + ab[0] * v + ac[0] * w2; q[1] = a[1] + ab[1] * v + ac[1] * w2; q[2] = a[2] + ab[2] * v + ac[2] * w2;
  }
done:
  if (w) copy3(w, q);
  { double r[3]; sub3(r, P, q); return len2_3(r); }
}

/* ccdMPRPenetration depth (findPenetr): distance from the origin to the final portal triangle.
 * An UPPER bound of the true penetration depth; used only to classify the 1e-4 m contact band. */
static double mpr_penetration_depth(const shape_t *A, const shape_t *B) {
  portal_t p;
  static const double origin[3] = {0, 0, 0};
  int res = discover_portal(A, B, &p);
  if (res < 0) return -1.0; /* separated */
  if (res == 1) return 0.0; /* touching */
  if (res == 2) { /* findPenetrSegment: depth = |v1| */ return sqrt(len2_3(p.v[1])); }
  if (refine_portal(A, B, &p) < 0) return -1.0;
  double dir[3], v4[3];
  for (long it = 0;; it++) {
    portal_dir(&p, dir);
    mink_support(A, B, dir, v4);
    if (portal_reach_tolerance(&p, v4, dir, ORC_MPR_TOL) || it > ORC_MAX_ITERS)
      return sqrt(point_tri_dist2(origin, p.v[1], p.v[2], p.v[3], NULL));
    expand_portal(&p, v4);
  }
}

/* ------------------------------------------------------------------ GJK distance
 * Role of FCL GJKDistance / libccd-style GJK (tolerance 1e-6): returns the separation distance, or -1
 * when the shapes touch or intersect (enable_signed_distance = false, python/pybind_fcl.hpp:311-315).
 * Implemented as textbook GJK with closest-point-on-simplex; equal to the upstream result within its
 * distance tolerance.  Witness points are not produced. */
static double closest_on_simplex(double S[4][3], int *n, double *cp) {
  static const double O[3] = {0, 0, 0};
  if (*n == 1) { copy3(cp, S[0]); return len2_3(cp); }
  if (*n == 2) {
    double d2 = point_seg_dist2(O, S[0], S[1], cp);
    /* reduce */
    if (eq3(cp, S[0])) { *n = 1; } else if (eq3(cp, S[1])) { copy3(S[0], S[1]); *n = 1; }
    return d2;
  }
  if (*n == 3) {
    double best = point_tri_dist2(O, S[0], S[1], S[2], cp);
    /* reduce to the supporting sub-simplex by testing edges/vertices */
    double w[3]; int keep[3] = {1, 1, 1};
    for (int e = 0; e < 3; e++) {
      int a = e, b = (e + 1) % 3, c = (e + 2) % 3;
      double d2 = point_seg_dist2(O, S[a], S[b], w);
      if (d2 <= best * (1 + 1e-14) + 1e-300) { keep[c] = 0; (void)c; /* closest lies on edge ab */
        double T[2][3]; copy3(T[0], S[a]); copy3(T[1], S[b]); copy3(S[0], T[0]); copy3(S[1], T[1]); *n = 2; copy3(cp, w);
        return closest_on_simplex(S, n, cp);
      }
    }
    (void)keep;
    return best;
  }
  /* tetrahedron: inside test, else closest over the 4 faces */
  {
    double best = DBL_MAX, bcp[3] = {0, 0, 0}; int bf = -1;
    static const int F[4][3] = {{0, 1, 2}, {0, 1, 3}, {0, 2, 3}, {1, 2, 3}};
    static const int opp[4] = {3, 2, 1, 0};
    int inside = 1;
    for (int f = 0; f < 4; f++) {
      double ab[3], ac[3], nrm[3], ao[3], ad[3];
      sub3(ab, S[F[f][1]], S[F[f][0]]); sub3(ac, S[F[f][2]], S[F[f][0]]); cross3(nrm, ab, ac);
      sub3(ad, S[opp[f]], S[F[f][0]]);
      ao[0] = -S[F[f][0]][0]; ao[1] = -S[F[f][0]][1]; ao[2] = -S[F[f][0]][2];
      double so = dot3(nrm, ao), sd = dot3(nrm, ad);
      if (so * sd < 0 || sd == 0) { /* origin on the outer side of this face (or degenerate tetra) */
        inside = 0;
        double w[3]; double d2 = point_tri_dist2(O, S[F[f][0]], S[F[f][1]], S[F[f][2]], w);
        if (d2 < best) { best = d2; copy3(bcp, w); bf = f; }
      }
    }
    if (inside) { cp[0] = cp[1] = cp[2] = 0; return 0.0; }
    double T[3][3]; copy3(T[0], S[F[bf][0]]); copy3(T[1], S[F[bf][1]]); copy3(T[2], S[F[bf][2]]);
    copy3(S[0], T[0]); copy3(S[1], T[1]); copy3(S[2], T[2]); *n = 3;
    copy3(cp, bcp);
    return closest_on_simplex(S, n, cp);
  }
}
static double gjk_distance(const shape_t *A, const shape_t *B) {
  double S[4][3]; int n = 0;
  double c1[3], c2[3], dir[3], cp[3], w[3];
  shape_center(A, c1); shape_center(B, c2); sub3(dir, c2, c1); /* search towards B */
  if (len2_3(dir) < 1e-24) { dir[0] = 1; dir[1] = 0; dir[2] = 0; }
  normalize3(dir);
  mink_support(A, B, dir, S[0]); n = 1;
  copy3(cp, S[0]);
  double d2 = len2_3(cp);
  for (int it = 0; it < ORC_MAX_ITERS; it++) {
    if (d2 < 1e-24) return -1.0;
    double dist = sqrt(d2);
    dir[0] = -cp[0] / dist; dir[1] = -cp[1] / dist; dir[2] = -cp[2] / dist;
    mink_support(A, B, dir, w); g_cnt.iters++;
    /* lower bound from the supporting plane: -w.dir ; converged when gap small */
    double lb = -dot3(w, dir);
    if (dist - lb <= 1e-10 * (dist > 1 ? dist : 1.0)) return dist;
    int dup = 0;
    for (int i = 0; i < n; i++) if (eq3(S[i], w)) dup = 1;
    if (dup) return dist;
    copy3(S[n], w); n++;
    d2 = closest_on_simplex(S, &n, cp);
    if (n == 4 && d2 == 0.0) return -1.0;
  }
  return sqrt(d2);
}

/* ------------------------------------------------------------------ closed-form primitives [upstream FCL 0.7.0
 * narrowphase/detail/primitive_shape_algorithm/{sphere_sphere,sphere_box,sphere_capsule,sphere_cylinder,box_box}-inl.h] */
static int sphere_sphere(const shape_t *s1, const shape_t *s2) {
  double d[3]; sub3(d, s2->T.p, s1->T.p);
  return !(sqrt(len2_3(d)) > s1->a + s2->a);
}
static int sphere_box(const shape_t *sp, const shape_t *bx) {
  double c[3], n[3]; xf_apply_inv(c, &bx->T, sp->T.p);
  const double h[3] = {bx->a, bx->b, bx->c};
  int clamped = 0;
  for (int i = 0; i < 3; i++) { n[i] = c[i]; if (c[i] < -h[i]) { n[i] = -h[i]; clamped = 1; } if (c[i] > h[i]) { n[i] = h[i]; clamped = 1; } }
  if (!clamped) return 1;
  double d[3]; sub3(d, c, n);
  return !(len2_3(d) > sp->a * sp->a);
}
static int sphere_capsule(const shape_t *sp, const shape_t *cap) {
  double c[3]; xf_apply_inv(c, &cap->T, sp->T.p);
  double a[3] = {0, 0, cap->b}, b[3] = {0, 0, -cap->b}, w[3];
  double d2 = point_seg_dist2(c, a, b, w);
  double dist = sqrt(d2) - sp->a - cap->a;
  return !(dist > 0);
}
static int sphere_cylinder(const shape_t *sp, const shape_t *cy) {
  double c[3], n[3]; xf_apply_inv(c, &cy->T, sp->T.p);
  int clamped = 0;
  copy3(n, c);
  double r2 = c[0] * c[0] + c[1] * c[1];
  if (r2 > cy->a * cy->a) { double s = cy->a / sqrt(r2); n[0] = c[0] * s; n[1] = c[1] * s; clamped = 1; }
  if (c[2] > cy->b) { n[2] = cy->b; clamped = 1; } else if (c[2] < -cy->b) { n[2] = -cy->b; clamped = 1; }
  if (!clamped) return 1;
  double d[3]; sub3(d, c, n);
  return !(len2_3(d) > sp->a * sp->a);
}
/* boxBox2 as a boolean: 15-axis SAT, separated iff some axis has s > 0 (tested before normalisation) */
static int box_box(const shape_t *b1, const shape_t *b2) {
  double R[9], Q[9], p[3], pp[3];
  sub3(p, b2->T.p, b1->T.p);
  xf_rot_t(pp, &b1->T, p);
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
    double s = 0; for (int k = 0; k < 3; k++) s += b1->T.R[3 * k + i] * b2->T.R[3 * k + j];
    R[3 * i + j] = s; Q[3 * i + j] = fabs(s);
  }
  const double A[3] = {b1->a, b1->b, b1->c}, B[3] = {b2->a, b2->b, b2->c};
  for (int i = 0; i < 3; i++) { double s2 = fabs(pp[i]) - (A[i] + B[0] * Q[3 * i] + B[1] * Q[3 * i + 1] + B[2] * Q[3 * i + 2]); if (s2 > 0) return 0; }
  for (int j = 0; j < 3; j++) {
    double t = R[j] * pp[0] + R[3 + j] * pp[1] + R[6 + j] * pp[2];
    double s2 = fabs(t) - (B[j] + A[0] * Q[j] + A[1] * Q[3 + j] + A[2] * Q[6 + j]);
    if (s2 > 0) return 0;
  }
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
    int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    double expr1 = pp[i2] * R[3 * i1 + j] - pp[i1] * R[3 * i2 + j];
    double expr2 = A[i1] * Q[3 * i2 + j] + A[i2] * Q[3 * i1 + j] + B[j1] * Q[3 * i + j2] + B[j2] * Q[3 * i + j1];
    if (fabs(expr1) - expr2 > 0) return 0;
  }
  return 1;
}

/* GJKSolver_libccd::shapeIntersect dispatch [upstream]: specialised pairs, everything else -> MPR */
static int shape_intersect(const shape_t *A, const shape_t *B) {
  int ta = A->type, tb = B->type;
  if (ta == MPB_GEOM_SPHERE && tb == MPB_GEOM_SPHERE) return sphere_sphere(A, B);
  if (ta == MPB_GEOM_SPHERE && tb == MPB_GEOM_BOX) return sphere_box(A, B);
  if (ta == MPB_GEOM_BOX && tb == MPB_GEOM_SPHERE) return sphere_box(B, A);
  if (ta == MPB_GEOM_SPHERE && tb == MPB_GEOM_CAPSULE) return sphere_capsule(A, B);
  if (ta == MPB_GEOM_CAPSULE && tb == MPB_GEOM_SPHERE) return sphere_capsule(B, A);
  if (ta == MPB_GEOM_SPHERE && tb == MPB_GEOM_CYLINDER) return sphere_cylinder(A, B);
  if (ta == MPB_GEOM_CYLINDER && tb == MPB_GEOM_SPHERE) return sphere_cylinder(B, A);
  if (ta == MPB_GEOM_BOX && tb == MPB_GEOM_BOX) return box_box(A, B);
  return mpr_intersect(A, B);
}

/* ------------------------------------------------------------------ mesh / octree (exhaustive leaves)
 * FCL walks an OBBRSS hierarchy / the octree and tests leaves.  The hierarchy only prunes,
 * so the verdict equals "some leaf intersects"; the oracle tests every leaf behind a conservative
 * bounding-sphere reject. */
static double shape_bound_radius(const mpb_scene_desc *d, int o, double *center_local) {
  center_local[0] = center_local[1] = center_local[2] = 0;
  const double *p = d->obj_param + 4 * o;
  switch (d->obj_type[o]) {
    case MPB_GEOM_SPHERE: return p[0];
    case MPB_GEOM_BOX: return 0.5 * sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]);
    case MPB_GEOM_CAPSULE: return p[0] + 0.5 * p[1];
    case MPB_GEOM_CYLINDER: return sqrt(p[0] * p[0] + 0.25 * p[1] * p[1]);
    case MPB_GEOM_CONVEX: {
      int h = d->obj_geom[o]; int v0 = d->hull_vert_off[h], v1 = d->hull_vert_off[h + 1];
      double r = 0; for (int i = v0; i < v1; i++) { double l = sqrt(len2_3(d->hull_verts + 3 * i)); if (l > r) r = l; }
      return r;
    }
    default: return 1e30;
  }
}
/* [upstream] shapeTriangleIntersect: sphere-triangle closed form, else MPR with a triangle support */
static int shape_triangle(const shape_t *S, const double *P0, const double *P1, const double *P2, const xf_t *Ttri) {
  if (S->type == MPB_GEOM_SPHERE) {
    double a[3], b[3], c[3]; xf_apply(a, Ttri, P0); xf_apply(b, Ttri, P1); xf_apply(c, Ttri, P2);
    return !(point_tri_dist2(S->T.p, a, b, c, NULL) > S->a * S->a);
  }
  shape_t t; memset(&t, 0, sizeof(t)); t.type = ORC_TRI; t.T = *Ttri;
  copy3(t.tri, P0); copy3(t.tri + 3, P1); copy3(t.tri + 6, P2);
  return mpr_intersect(S, &t);
}
static int mesh_shape(const mpb_scene_desc *d, int om, const xf_t *Tm, int os, const xf_t *Ts, int mesh_first) {
  int m = d->obj_geom[om];
  const double *V = d->mesh_verts + 3 * d->mesh_vert_off[m];
  const int32_t *Tr = d->mesh_tris + 3 * d->mesh_tri_off[m];
  int nt = d->mesh_tri_off[m + 1] - d->mesh_tri_off[m];
  shape_t S; shape_from_object(d, os, Ts, &S);
  double cl[3]; double rad = shape_bound_radius(d, os, cl);
  double cw[3]; copy3(cw, Ts->p);
  (void)mesh_first;
  for (int t = 0; t < nt; t++) {
    const double *P0 = V + 3 * Tr[3 * t], *P1 = V + 3 * Tr[3 * t + 1], *P2 = V + 3 * Tr[3 * t + 2];
    double a[3], b[3], c[3]; xf_apply(a, Tm, P0); xf_apply(b, Tm, P1); xf_apply(c, Tm, P2);
    if (point_tri_dist2(cw, a, b, c, NULL) > rad * rad) continue;
    if (shape_triangle(&S, P0, P1, P2, Tm)) return 1;
  }
  return 0;
}
/* triangle-triangle intersection [upstream FCL 0.7.0 Intersect::intersect_Triangle = PQP TriContact]: separating-axis
 * test over 17 axes -- the two normals, the nine edge x edge products and the six in-plane edge normals (edge x own
 * normal), which are the ones that decide COPLANAR triangles.  Closed sets: touching counts as intersecting
 * (project6 separates only on a strict gap). */
static int project6(const double *ax, const double *p1, const double *p2, const double *p3, const double *q1, const double *q2, const double *q3) {
  double P1 = dot3(ax, p1), P2 = dot3(ax, p2), P3 = dot3(ax, p3), Q1 = dot3(ax, q1), Q2 = dot3(ax, q2), Q3 = dot3(ax, q3);
  double mx1 = fmax(P1, fmax(P2, P3)), mn1 = fmin(P1, fmin(P2, P3)), mx2 = fmax(Q1, fmax(Q2, Q3)), mn2 = fmin(Q1, fmin(Q2, Q3));
  if (mn1 > mx2) return 0;
  if (mn2 > mx1) return 0;
  return 1;
}
static int tri_tri(const double *A, const double *B) {
  double p1[3] = {0, 0, 0}, p2[3], p3[3], q1[3], q2[3], q3[3];
  sub3(p2, A + 3, A); sub3(p3, A + 6, A); sub3(q1, B, A); sub3(q2, B + 3, A); sub3(q3, B + 6, A);
  double e[3][3], f[3][3], n1[3], m1[3], ax[3];
  sub3(e[0], p2, p1); sub3(e[1], p3, p2); sub3(e[2], p1, p3);
  sub3(f[0], q2, q1); sub3(f[1], q3, q2); sub3(f[2], q1, q3);
  cross3(n1, e[0], e[1]); cross3(m1, f[0], f[1]);
  if (!project6(n1, p1, p2, p3, q1, q2, q3)) return 0;
  if (!project6(m1, p1, p2, p3, q1, q2, q3)) return 0;
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { cross3(ax, e[i], f[j]); if (!project6(ax, p1, p2, p3, q1, q2, q3)) return 0; }
  for (int i = 0; i < 3; i++) { cross3(ax, e[i], n1); if (!project6(ax, p1, p2, p3, q1, q2, q3)) return 0; }
  for (int j = 0; j < 3; j++) { cross3(ax, f[j], m1); if (!project6(ax, p1, p2, p3, q1, q2, q3)) return 0; }
  return 1;
}
static int mesh_mesh(const mpb_scene_desc *d, int oa, const xf_t *Ta, int ob, const xf_t *Tb) {
  int ma = d->obj_geom[oa], mb = d->obj_geom[ob];
  const double *Va = d->mesh_verts + 3 * d->mesh_vert_off[ma], *Vb = d->mesh_verts + 3 * d->mesh_vert_off[mb];
  const int32_t *Ia = d->mesh_tris + 3 * d->mesh_tri_off[ma], *Ib = d->mesh_tris + 3 * d->mesh_tri_off[mb];
  int na = d->mesh_tri_off[ma + 1] - d->mesh_tri_off[ma], nb = d->mesh_tri_off[mb + 1] - d->mesh_tri_off[mb];
  double *WB = (double *)malloc(sizeof(double) * 9 * (size_t)nb);
  for (int t = 0; t < nb; t++) for (int k = 0; k < 3; k++) xf_apply(WB + 9 * t + 3 * k, Tb, Vb + 3 * Ib[3 * t + k]);
  int hit = 0;
  for (int t = 0; t < na && !hit; t++) {
    double A[9]; for (int k = 0; k < 3; k++) xf_apply(A + 3 * k, Ta, Va + 3 * Ia[3 * t + k]);
    for (int u = 0; u < nb; u++) if (tri_tri(A, WB + 9 * u)) { hit = 1; break; }
  }
  free(WB);
  return hit;
}
/* OcTreeShapeIntersectRecurse leaves [upstream]: Box(res) centred on the voxel vs the shape through
 * shape_intersect(box, shape); octomap voxel centre = (cell + 0.5) * res (key - 32768 + 0.5). */
static int octree_shape(const mpb_scene_desc *d, int oo, const xf_t *To, int os, const xf_t *Ts) {
  int t = d->obj_geom[oo];
  double res = d->oct_res[t];
  const int32_t *C = d->oct_cells + 3 * d->oct_cell_off[t];
  int nc = d->oct_cell_off[t + 1] - d->oct_cell_off[t];
  shape_t S; shape_from_object(d, os, Ts, &S);
  double cl[3]; double rad = shape_bound_radius(d, os, cl);
  double cs[3]; xf_apply_inv(cs, To, Ts->p); /* shape origin in the octree frame */
  double reach = rad + 0.5 * res * sqrt(3.0);
  shape_t bx; memset(&bx, 0, sizeof(bx)); bx.type = MPB_GEOM_BOX; bx.a = bx.b = bx.c = 0.5 * res;
  for (int i = 0; i < nc; i++) {
    double c[3] = {(C[3 * i] + 0.5) * res, (C[3 * i + 1] + 0.5) * res, (C[3 * i + 2] + 0.5) * res};
    double dd[3]; sub3(dd, c, cs);
    if (len2_3(dd) > reach * reach) continue;
    bx.T = *To; xf_apply(bx.T.p, To, c);
    if (shape_intersect(&bx, &S)) return 1;
  }
  return 0;
}
static int octree_mesh(const mpb_scene_desc *d, int oo, const xf_t *To, int om, const xf_t *Tm) {
  int t = d->obj_geom[oo]; double res = d->oct_res[t];
  const int32_t *C = d->oct_cells + 3 * d->oct_cell_off[t];
  int nc = d->oct_cell_off[t + 1] - d->oct_cell_off[t];
  int m = d->obj_geom[om];
  const double *V = d->mesh_verts + 3 * d->mesh_vert_off[m];
  const int32_t *Tr = d->mesh_tris + 3 * d->mesh_tri_off[m];
  int nt = d->mesh_tri_off[m + 1] - d->mesh_tri_off[m];
  shape_t bx; memset(&bx, 0, sizeof(bx)); bx.type = MPB_GEOM_BOX; bx.a = bx.b = bx.c = 0.5 * res;
  double reach = 0.5 * res * sqrt(3.0);
  for (int u = 0; u < nt; u++) {
    const double *P0 = V + 3 * Tr[3 * u], *P1 = V + 3 * Tr[3 * u + 1], *P2 = V + 3 * Tr[3 * u + 2];
    double a[3], b[3], c[3]; xf_apply(a, Tm, P0); xf_apply(b, Tm, P1); xf_apply(c, Tm, P2);
    double al[3], bl[3], cl2[3]; xf_apply_inv(al, To, a); xf_apply_inv(bl, To, b); xf_apply_inv(cl2, To, c);
    for (int i = 0; i < nc; i++) {
      double cc[3] = {(C[3 * i] + 0.5) * res, (C[3 * i + 1] + 0.5) * res, (C[3 * i + 2] + 0.5) * res};
      if (point_tri_dist2(cc, al, bl, cl2, NULL) > reach * reach) continue;
      bx.T = *To; xf_apply(bx.T.p, To, cc);
      if (shape_triangle(&bx, P0, P1, P2, Tm)) return 1;
    }
  }
  return 0;
}

/* ------------------------------------------------------------------ hierarchical leg (the TIMED CPU baseline)
 * FCL does not test every leaf: ::fcl::collide descends a BVHModel<OBBRSS> (load_mesh_as_BVH, src/urdf_utils.cpp:135-154)
 * or the octree (addPointCloud, src/planning_world.cpp:101-110) and stops at the first hit (num_max_contacts = 1), and
 * fcl::Convex finds its support vertex by walking the hull's edges.  The functions below do that work: an AABB tree per
 * mesh / per occupied-voxel set (median split, one primitive per leaf), node-vs-shape-OBB overlap for the descent
 * (role of OBBRSS::overlap / OcTreeSolver's obb1.overlap(obb2)), the SAME leaf tests as the exhaustive functions above,
 * early exit.  A pruned leaf cannot intersect (the bounds are conservative, padded by ORC_PRUNE_PAD), so the verdict of
 * every pair equals the exhaustive leg's -- tests/test_oracle_accel.py checks that bit for bit.  The exhaustive leg
 * stays the CHECKER; this one is what bench.py times. */
#define ORC_PRUNE_PAD 1e-6
typedef struct { double c[3], h[3]; int left, right; } orc_node; /* inner: children; leaf: left = ~primitive, right = 1 */
typedef struct { orc_node *nodes; int n_nodes, cap; } orc_bvh;
typedef struct orc_accel {
  int n_meshes, n_octrees, n_hulls;
  orc_bvh *mesh, *oct;
  int **hull_nbr;        /* per hull: CSR adjacency, NULL for hulls of <= 32 vertices */
  double *hull_box;      /* per hull: centre[3], half[3] of the vertex AABB (hull frame) */
} orc_accel;

typedef struct { double lo[3], hi[3]; int prim; } orc_prim_box;
static int g_sort_axis;
static int cmp_prim_box(const void *a, const void *b) {
  const orc_prim_box *x = (const orc_prim_box *)a, *y = (const orc_prim_box *)b;
  double cx = x->lo[g_sort_axis] + x->hi[g_sort_axis], cy = y->lo[g_sort_axis] + y->hi[g_sort_axis];
  return cx < cy ? -1 : (cx > cy ? 1 : (x->prim - y->prim));
}
static int bvh_build(orc_bvh *t, orc_prim_box *pb, int first, int count) {
  if (t->n_nodes == t->cap) { t->cap = t->cap ? 2 * t->cap : 1024; t->nodes = (orc_node *)realloc(t->nodes, sizeof(orc_node) * (size_t)t->cap); }
  int me = t->n_nodes++;
  double lo[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, hi[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX}, clo[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, chi[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
  for (int i = first; i < first + count; i++) for (int a = 0; a < 3; a++) {
    if (pb[i].lo[a] < lo[a]) lo[a] = pb[i].lo[a];
    if (pb[i].hi[a] > hi[a]) hi[a] = pb[i].hi[a];
    double cc = 0.5 * (pb[i].lo[a] + pb[i].hi[a]); if (cc < clo[a]) clo[a] = cc; if (cc > chi[a]) chi[a] = cc;
  }
  for (int a = 0; a < 3; a++) { t->nodes[me].c[a] = 0.5 * (lo[a] + hi[a]); t->nodes[me].h[a] = 0.5 * (hi[a] - lo[a]) + ORC_PRUNE_PAD; }
  if (count == 1) { t->nodes[me].left = ~pb[first].prim; t->nodes[me].right = 1; return me; }
  int ax = 0; if (chi[1] - clo[1] > chi[ax] - clo[ax]) ax = 1; if (chi[2] - clo[2] > chi[ax] - clo[ax]) ax = 2;
  g_sort_axis = ax;
  qsort(pb + first, (size_t)count, sizeof(orc_prim_box), cmp_prim_box);
  int mid = count / 2;
  int l = bvh_build(t, pb, first, mid);
  int r = bvh_build(t, pb, first + mid, count - mid);
  t->nodes[me].left = l; t->nodes[me].right = r;
  return me;
}

void orc_accel_destroy(orc_accel *ac) {
  if (!ac) return;
  for (int m = 0; m < ac->n_meshes; m++) free(ac->mesh[m].nodes);
  for (int t = 0; t < ac->n_octrees; t++) free(ac->oct[t].nodes);
  for (int h = 0; h < ac->n_hulls; h++) free(ac->hull_nbr[h]);
  free(ac->mesh); free(ac->oct); free(ac->hull_nbr); free(ac->hull_box); free(ac);
}

orc_accel *orc_accel_create(const mpb_scene_desc *d) {
  orc_accel *ac = (orc_accel *)calloc(1, sizeof(orc_accel));
  ac->n_meshes = d->n_meshes; ac->n_octrees = d->n_octrees; ac->n_hulls = d->n_hulls;
  ac->mesh = (orc_bvh *)calloc((size_t)(d->n_meshes > 0 ? d->n_meshes : 1), sizeof(orc_bvh));
  ac->oct = (orc_bvh *)calloc((size_t)(d->n_octrees > 0 ? d->n_octrees : 1), sizeof(orc_bvh));
  ac->hull_nbr = (int **)calloc((size_t)(d->n_hulls > 0 ? d->n_hulls : 1), sizeof(int *));
  ac->hull_box = (double *)calloc((size_t)(d->n_hulls > 0 ? d->n_hulls : 1) * 6, sizeof(double));
  for (int m = 0; m < d->n_meshes; m++) {
    const double *V = d->mesh_verts + 3 * d->mesh_vert_off[m];
    const int32_t *Tr = d->mesh_tris + 3 * d->mesh_tri_off[m];
    int nt = d->mesh_tri_off[m + 1] - d->mesh_tri_off[m];
    if (nt <= 0) continue;
    orc_prim_box *pb = (orc_prim_box *)malloc(sizeof(orc_prim_box) * (size_t)nt);
    for (int t = 0; t < nt; t++) {
      pb[t].prim = t;
      for (int a = 0; a < 3; a++) {
        double x0 = V[3 * Tr[3 * t] + a], x1 = V[3 * Tr[3 * t + 1] + a], x2 = V[3 * Tr[3 * t + 2] + a];
        pb[t].lo[a] = fmin(x0, fmin(x1, x2)); pb[t].hi[a] = fmax(x0, fmax(x1, x2));
      }
    }
    bvh_build(&ac->mesh[m], pb, 0, nt);
    free(pb);
  }
  for (int t = 0; t < d->n_octrees; t++) {
    double res = d->oct_res[t];
    const int32_t *C = d->oct_cells + 3 * d->oct_cell_off[t];
    int nc = d->oct_cell_off[t + 1] - d->oct_cell_off[t];
    if (nc <= 0) continue;
    orc_prim_box *pb = (orc_prim_box *)malloc(sizeof(orc_prim_box) * (size_t)nc);
    for (int i = 0; i < nc; i++) { pb[i].prim = i; for (int a = 0; a < 3; a++) { pb[i].lo[a] = C[3 * i + a] * res; pb[i].hi[a] = (C[3 * i + a] + 1) * res; } }
    bvh_build(&ac->oct[t], pb, 0, nc);
    free(pb);
  }
  for (int h = 0; h < d->n_hulls; h++) {
    int v0 = d->hull_vert_off[h], nv = d->hull_vert_off[h + 1] - v0;
    const double *V = d->hull_verts + 3 * v0;
    double lo[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, hi[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
    for (int i = 0; i < nv; i++) for (int a = 0; a < 3; a++) { if (V[3 * i + a] < lo[a]) lo[a] = V[3 * i + a]; if (V[3 * i + a] > hi[a]) hi[a] = V[3 * i + a]; }
    for (int a = 0; a < 3; a++) { ac->hull_box[6 * h + a] = 0.5 * (lo[a] + hi[a]); ac->hull_box[6 * h + 3 + a] = 0.5 * (hi[a] - lo[a]); }
    if (nv <= 32) continue;
    /* adjacency from the triangle faces (each edge listed once per endpoint) */
    const int32_t *F = d->hull_faces + 3 * d->hull_face_off[h];
    int nf = d->hull_face_off[h + 1] - d->hull_face_off[h];
    unsigned char *adj = (unsigned char *)calloc((size_t)nv * (size_t)nv, 1);
    for (int f = 0; f < nf; f++) for (int e = 0; e < 3; e++) { int i = F[3 * f + e], j = F[3 * f + (e + 1) % 3]; if (i >= 0 && i < nv && j >= 0 && j < nv && i != j) { adj[(size_t)i * nv + j] = 1; adj[(size_t)j * nv + i] = 1; } }
    int total = 0; for (size_t k = 0; k < (size_t)nv * nv; k++) total += adj[k];
    int *nbr = (int *)malloc(sizeof(int) * (size_t)(nv + 1 + total));
    int off = 0, connected = 1;
    for (int i = 0; i < nv; i++) {
      nbr[i] = off;
      for (int j = 0; j < nv; j++) if (adj[(size_t)i * nv + j]) nbr[nv + 1 + off++] = j;
      if (off == nbr[i]) connected = 0; /* a vertex no face references: the walk could miss it -> keep the scan */
    }
    nbr[nv] = off;
    free(adj);
    if (connected) ac->hull_nbr[h] = nbr; else free(nbr);
  }
  return ac;
}

/* the shape's OBB in its own frame: centre + half extents (primitives: FCL computeBV<OBB>; convex: vertex AABB) */
static void shape_local_box(const mpb_scene_desc *d, const orc_accel *ac, int o, double *c, double *h) {
  const double *p = d->obj_param + 4 * o;
  c[0] = c[1] = c[2] = 0;
  switch (d->obj_type[o]) {
    case MPB_GEOM_SPHERE: h[0] = h[1] = h[2] = p[0]; break;
    case MPB_GEOM_BOX: h[0] = 0.5 * p[0]; h[1] = 0.5 * p[1]; h[2] = 0.5 * p[2]; break;
    case MPB_GEOM_CAPSULE: h[0] = h[1] = p[0]; h[2] = 0.5 * p[1] + p[0]; break;
    case MPB_GEOM_CYLINDER: h[0] = h[1] = p[0]; h[2] = 0.5 * p[1]; break;
    case MPB_GEOM_CONVEX: { const double *b = ac->hull_box + 6 * d->obj_geom[o]; copy3(c, b); copy3(h, b + 3); } break;
    default: h[0] = h[1] = h[2] = 1e30; break;
  }
}
/* box A (axis-aligned in its frame, centre ca, half ha) vs box B (axes = columns of R, centre cb in A's frame, half hb) */
static int boxes_disjoint(const double *ca, const double *ha, const double *R, const double *cb, const double *hb) {
  double t[3] = {cb[0] - ca[0], cb[1] - ca[1], cb[2] - ca[2]}, Q[9];
  for (int i = 0; i < 9; i++) Q[i] = fabs(R[i]) + 1e-12;
  for (int i = 0; i < 3; i++) if (fabs(t[i]) > ha[i] + hb[0] * Q[3 * i] + hb[1] * Q[3 * i + 1] + hb[2] * Q[3 * i + 2]) return 1;
  for (int j = 0; j < 3; j++) {
    double tj = t[0] * R[j] + t[1] * R[3 + j] + t[2] * R[6 + j];
    if (fabs(tj) > hb[j] + ha[0] * Q[j] + ha[1] * Q[3 + j] + ha[2] * Q[6 + j]) return 1;
  }
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
    int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    double lhs = fabs(t[i2] * R[3 * i1 + j] - t[i1] * R[3 * i2 + j]);
    double rhs = ha[i1] * Q[3 * i2 + j] + ha[i2] * Q[3 * i1 + j] + hb[j1] * Q[3 * i + j2] + hb[j2] * Q[3 * i + j1];
    if (lhs > rhs) return 1;
  }
  return 0;
}
static void xf_inv_mul_(xf_t *o, const xf_t *a, const xf_t *b) { /* a^-1 b */
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { double v = 0; for (int k = 0; k < 3; k++) v += a->R[3 * k + i] * b->R[3 * k + j]; o->R[3 * i + j] = v; }
  double dd[3]; sub3(dd, b->p, a->p); xf_rot_t(o->p, a, dd);
}
#define ORC_STACK 256
/* shape (object os at Ts) against the tree of object ob (mesh or octree) at Tb */
static int tree_shape_h(const mpb_scene_desc *d, const orc_accel *ac, int ob, const xf_t *Tb, int os, const xf_t *Ts) {
  const int is_oct = d->obj_type[ob] == MPB_GEOM_OCTREE;
  const int g = d->obj_geom[ob];
  const orc_bvh *t = is_oct ? &ac->oct[g] : &ac->mesh[g];
  if (t->n_nodes == 0) return 0;
  shape_t S; shape_from_object(d, os, Ts, &S);
  if (S.type == MPB_GEOM_CONVEX) S.nbr = ac->hull_nbr[d->obj_geom[os]];
  xf_t rel; xf_inv_mul_(&rel, Tb, Ts); /* shape in the tree's frame */
  double cl[3], hb[3], cb[3]; shape_local_box(d, ac, os, cl, hb); xf_apply(cb, &rel, cl);
  for (int a = 0; a < 3; a++) hb[a] += ORC_PRUNE_PAD;
  double res = is_oct ? d->oct_res[g] : 0;
  const int32_t *C = is_oct ? d->oct_cells + 3 * d->oct_cell_off[g] : NULL;
  const double *V = is_oct ? NULL : d->mesh_verts + 3 * d->mesh_vert_off[g];
  const int32_t *Tr = is_oct ? NULL : d->mesh_tris + 3 * d->mesh_tri_off[g];
  shape_t bx; memset(&bx, 0, sizeof(bx)); bx.type = MPB_GEOM_BOX; bx.a = bx.b = bx.c = 0.5 * res;
  int stack[ORC_STACK], sp = 0; stack[sp++] = 0;
  while (sp > 0) {
    const orc_node *nd = &t->nodes[stack[--sp]];
    if (boxes_disjoint(nd->c, nd->h, rel.R, cb, hb)) continue;
    if (nd->left >= 0) { if (sp + 2 > ORC_STACK) return -1; stack[sp++] = nd->right; stack[sp++] = nd->left; continue; }
    const int prim = ~nd->left;
    if (is_oct) {
      double c[3] = {(C[3 * prim] + 0.5) * res, (C[3 * prim + 1] + 0.5) * res, (C[3 * prim + 2] + 0.5) * res};
      bx.T = *Tb; xf_apply(bx.T.p, Tb, c);
      if (shape_intersect(&bx, &S)) return 1;
    } else if (shape_triangle(&S, V + 3 * Tr[3 * prim], V + 3 * Tr[3 * prim + 1], V + 3 * Tr[3 * prim + 2], Tb)) return 1;
  }
  return 0;
}
/* mesh oa at Ta against mesh / octree ob at Tb: simultaneous descent of both trees */
static int tree_tree_h(const mpb_scene_desc *d, const orc_accel *ac, int oa, const xf_t *Ta, int ob, const xf_t *Tb) {
  const int ma = d->obj_geom[oa], gb = d->obj_geom[ob];
  const int b_oct = d->obj_type[ob] == MPB_GEOM_OCTREE;
  const orc_bvh *ta = &ac->mesh[ma], *tb = b_oct ? &ac->oct[gb] : &ac->mesh[gb];
  if (ta->n_nodes == 0 || tb->n_nodes == 0) return 0;
  xf_t rel; xf_inv_mul_(&rel, Ta, Tb); /* B in A */
  const double *Va = d->mesh_verts + 3 * d->mesh_vert_off[ma];
  const int32_t *Ia = d->mesh_tris + 3 * d->mesh_tri_off[ma];
  const double *Vb = b_oct ? NULL : d->mesh_verts + 3 * d->mesh_vert_off[gb];
  const int32_t *Ib = b_oct ? NULL : d->mesh_tris + 3 * d->mesh_tri_off[gb];
  double res = b_oct ? d->oct_res[gb] : 0;
  const int32_t *C = b_oct ? d->oct_cells + 3 * d->oct_cell_off[gb] : NULL;
  shape_t bx; memset(&bx, 0, sizeof(bx)); bx.type = MPB_GEOM_BOX; bx.a = bx.b = bx.c = 0.5 * res;
  int stack[2 * ORC_STACK], sp = 0; stack[sp++] = 0; stack[sp++] = 0;
  while (sp > 0) {
    const int ib = stack[--sp], ia = stack[--sp];
    const orc_node *na = &ta->nodes[ia], *nb = &tb->nodes[ib];
    double cb[3]; xf_apply(cb, &rel, nb->c);
    if (boxes_disjoint(na->c, na->h, rel.R, cb, nb->h)) continue;
    if (na->left < 0 && nb->left < 0) {
      const int pa = ~na->left, pb = ~nb->left;
      if (b_oct) {
        double c[3] = {(C[3 * pb] + 0.5) * res, (C[3 * pb + 1] + 0.5) * res, (C[3 * pb + 2] + 0.5) * res};
        bx.T = *Tb; xf_apply(bx.T.p, Tb, c);
        if (shape_triangle(&bx, Va + 3 * Ia[3 * pa], Va + 3 * Ia[3 * pa + 1], Va + 3 * Ia[3 * pa + 2], Ta)) return 1;
      } else {
        double A[9], B[9];
        for (int k = 0; k < 3; k++) { xf_apply(A + 3 * k, Ta, Va + 3 * Ia[3 * pa + k]); xf_apply(B + 3 * k, Tb, Vb + 3 * Ib[3 * pb + k]); }
        if (tri_tri(A, B)) return 1;
      }
      continue;
    }
    if (sp + 4 > 2 * ORC_STACK) return -1;
    const double sa = na->h[0] + na->h[1] + na->h[2], sb = nb->h[0] + nb->h[1] + nb->h[2];
    if (nb->left < 0 || (na->left >= 0 && sa >= sb)) { stack[sp++] = na->right; stack[sp++] = ib; stack[sp++] = na->left; stack[sp++] = ib; }
    else { stack[sp++] = ia; stack[sp++] = nb->right; stack[sp++] = ia; stack[sp++] = nb->left; }
  }
  return 0;
}

/* ::fcl::collide(o1, o2) as a boolean for the geometry types on the hot path */
static int object_collide_ac(const mpb_scene_desc *d, const orc_accel *ac, int a, const xf_t *Ta, int b, const xf_t *Tb);
static int object_collide(const mpb_scene_desc *d, int a, const xf_t *Ta, int b, const xf_t *Tb) {
  int ta = d->obj_type[a], tb = d->obj_type[b];
  if (ta == MPB_GEOM_OCTREE && tb == MPB_GEOM_OCTREE) return 0; /* not produced by the path */
  if (ta == MPB_GEOM_OCTREE) return tb == MPB_GEOM_MESH ? octree_mesh(d, a, Ta, b, Tb) : octree_shape(d, a, Ta, b, Tb);
  if (tb == MPB_GEOM_OCTREE) return ta == MPB_GEOM_MESH ? octree_mesh(d, b, Tb, a, Ta) : octree_shape(d, b, Tb, a, Ta);
  if (ta == MPB_GEOM_MESH && tb == MPB_GEOM_MESH) return mesh_mesh(d, a, Ta, b, Tb);
  if (ta == MPB_GEOM_MESH) return mesh_shape(d, a, Ta, b, Tb, 1);
  if (tb == MPB_GEOM_MESH) return mesh_shape(d, b, Tb, a, Ta, 0);
  shape_t A, B; shape_from_object(d, a, Ta, &A); shape_from_object(d, b, Tb, &B);
  return shape_intersect(&A, &B);
}

/* the same dispatch through the hierarchies (ac != NULL); a traversal that outgrows its stack falls back to the exhaustive test */
static int object_collide_ac(const mpb_scene_desc *d, const orc_accel *ac, int a, const xf_t *Ta, int b, const xf_t *Tb) {
  if (!ac) return object_collide(d, a, Ta, b, Tb);
  int ta = d->obj_type[a], tb = d->obj_type[b], r;
  if (ta == MPB_GEOM_OCTREE && tb == MPB_GEOM_OCTREE) return 0;
  if (ta >= MPB_GEOM_MESH && tb >= MPB_GEOM_MESH) r = ta == MPB_GEOM_MESH ? tree_tree_h(d, ac, a, Ta, b, Tb) : tree_tree_h(d, ac, b, Tb, a, Ta);
  else if (ta >= MPB_GEOM_MESH) r = tree_shape_h(d, ac, a, Ta, b, Tb);
  else if (tb >= MPB_GEOM_MESH) r = tree_shape_h(d, ac, b, Tb, a, Ta);
  else {
    shape_t A, B; shape_from_object(d, a, Ta, &A); shape_from_object(d, b, Tb, &B);
    if (A.type == MPB_GEOM_CONVEX) A.nbr = ac->hull_nbr[d->obj_geom[a]];
    if (B.type == MPB_GEOM_CONVEX) B.nbr = ac->hull_nbr[d->obj_geom[b]];
    return shape_intersect(&A, &B);
  }
  return r < 0 ? object_collide(d, a, Ta, b, Tb) : r;
}

/* ------------------------------------------------------------------ exported API */
/* FK of a batch: link poses [n][n_links][7] (getLinkPose for every link). */
int orc_fk_batch(const mpb_scene_desc *d, const double *states, int64_t n, double *poses) {
  xf_t *oMi = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_joints + 1));
  double *qf = (double *)malloc(sizeof(double) * (size_t)(d->nq_full > 0 ? d->nq_full : 1));
  for (int64_t i = 0; i < n; i++) {
    scatter_state(d, states + i * d->state_dim, qf);
    fk_joints(d, qf, oMi);
    for (int l = 0; l < d->n_links; l++) link_pose7(d, oMi, l, poses + (i * d->n_links + l) * 7);
  }
  free(oMi); free(qf);
  return 0;
}

/* Object world transforms [n][n_objects][12] after setQposAll (for tests). */
int orc_object_poses(const mpb_scene_desc *d, const double *states, int64_t n, double *out12) {
  xf_t *oMi = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_joints + 1));
  xf_t *objT = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_objects > 0 ? d->n_objects : 1));
  double *qf = (double *)malloc(sizeof(double) * (size_t)(d->nq_full > 0 ? d->nq_full : 1));
  for (int64_t i = 0; i < n; i++) {
    scatter_state(d, states + i * d->state_dim, qf);
    fk_joints(d, qf, oMi);
    object_transforms(d, oMi, objT, NULL);
    for (int o = 0; o < d->n_objects; o++) {
      double *m = out12 + (i * d->n_objects + o) * 12;
      for (int r = 0; r < 3; r++) { m[4 * r] = objT[o].R[3 * r]; m[4 * r + 1] = objT[o].R[3 * r + 1]; m[4 * r + 2] = objT[o].R[3 * r + 2]; m[4 * r + 3] = objT[o].p[r]; }
    }
  }
  free(oMi); free(objT); free(qf);
  return 0;
}

/* ValidityCheckerTpl::isValid for a batch (src/ompl_planner.h:59-62): per state, every admitted pair
 * is tested (no early exit, src/planning_world.h:248-250).  pair_hit: [n][n_pairs] bytes (nullable),
 * verdict: [n] bytes (nullable).  Returns total support evaluations through *iters (nullable). */
static int64_t collide_range(const mpb_scene_desc *d, const orc_accel *ac, const double *states, int64_t i0, int64_t i1, uint8_t *pair_hit, uint8_t *verdict) {
  xf_t *oMi = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_joints + 1));
  xf_t *objT = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_objects > 0 ? d->n_objects : 1));
  double *qf = (double *)malloc(sizeof(double) * (size_t)(d->nq_full > 0 ? d->nq_full : 1));
  g_cnt.iters = 0;
  for (int64_t i = i0; i < i1; i++) {
    scatter_state(d, states + i * d->state_dim, qf);
    fk_joints(d, qf, oMi);
    object_transforms(d, oMi, objT, NULL);
    int any = 0;
    for (int p = 0; p < d->n_pairs; p++) {
      int a = d->pair_a[p], b = d->pair_b[p];
      int hit = object_collide_ac(d, ac, a, &objT[a], b, &objT[b]);
      if (pair_hit) pair_hit[i * d->n_pairs + p] = (uint8_t)hit;
      any |= hit;
    }
    if (verdict) verdict[i] = (uint8_t)any;
  }
  free(oMi); free(objT); free(qf);
  return g_cnt.iters;
}
int orc_collide_batch(const mpb_scene_desc *d, const double *states, int64_t n, uint8_t *pair_hit, uint8_t *verdict, int64_t *iters) {
  int64_t it = collide_range(d, NULL, states, 0, n, pair_hit, verdict);
  if (iters) *iters = it;
  return 0;
}
/* All-host-core variant: the stateful reference can only use more than one core as independent world
 * clones, one per thread; each thread owns its scratch and a contiguous slice. */
static int collide_batch_mt(const mpb_scene_desc *d, const orc_accel *ac, const double *states, int64_t n, uint8_t *pair_hit, uint8_t *verdict, int64_t *iters, int nthreads) {
  int64_t total = 0;
  if (nthreads < 1) nthreads = 1;
  /* slices of 256 states handed out dynamically: states near obstacles cost far more than free ones */
  const int64_t slice = 256, n_slices = (n + slice - 1) / slice;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1) reduction(+ : total)
  for (int64_t k = 0; k < n_slices; k++) {
    int64_t i0 = k * slice, i1 = i0 + slice < n ? i0 + slice : n;
    total += collide_range(d, ac, states, i0, i1, pair_hit, verdict);
  }
  if (iters) *iters = total;
  return 0;
}
int orc_collide_batch_mt(const mpb_scene_desc *d, const double *states, int64_t n, uint8_t *pair_hit, uint8_t *verdict, int64_t *iters, int nthreads) {
  return collide_batch_mt(d, NULL, states, n, pair_hit, verdict, iters, nthreads);
}
/* the timed baseline: same loop through the hierarchies of `ac` (orc_accel_create) */
int orc_collide_batch_accel(const mpb_scene_desc *d, const orc_accel *ac, const double *states, int64_t n, uint8_t *pair_hit, uint8_t *verdict, int64_t *iters, int nthreads) {
  return collide_batch_mt(d, ac, states, n, pair_hit, verdict, iters, nthreads);
}

/* Band classifier for one (state, pair): signed distance estimate.  > 0: GJK separation distance;
 * < 0: -(MPR penetration depth) (an upper bound of the true depth => |value| >= true depth);
 * mesh/octree pairs return NaN (callers treat those with the exhaustive-leaf margin test). */
static double leafset_signed_distance(const mpb_scene_desc *d, int ob, const xf_t *Tb, int os, const xf_t *Ts) {
  /* ob = mesh or octree object, os = shape.  Separated: min leaf distance; intersecting: -(max leaf MPR depth). */
  shape_t S; shape_from_object(d, os, Ts, &S);
  double cl[3]; double rad = shape_bound_radius(d, os, cl);
  double mind = DBL_MAX, maxdepth = -1.0;
  if (d->obj_type[ob] == MPB_GEOM_OCTREE) {
    int t = d->obj_geom[ob]; double res = d->oct_res[t];
    const int32_t *C = d->oct_cells + 3 * d->oct_cell_off[t];
    int nc = d->oct_cell_off[t + 1] - d->oct_cell_off[t];
    double cs[3]; xf_apply_inv(cs, Tb, Ts->p);
    double reach = rad + 0.5 * res * sqrt(3.0) + 0.01;
    shape_t bx; memset(&bx, 0, sizeof(bx)); bx.type = MPB_GEOM_BOX; bx.a = bx.b = bx.c = 0.5 * res;
    for (int i = 0; i < nc; i++) {
      double c[3] = {(C[3 * i] + 0.5) * res, (C[3 * i + 1] + 0.5) * res, (C[3 * i + 2] + 0.5) * res};
      double dd[3]; sub3(dd, c, cs);
      if (len2_3(dd) > reach * reach) continue;
      bx.T = *Tb; xf_apply(bx.T.p, Tb, c);
      double g = gjk_distance(&bx, &S);
      if (g > 0) { if (g < mind) mind = g; }
      else { double pd = mpr_penetration_depth(&bx, &S); if (pd < 0) pd = 0; if (pd > maxdepth) maxdepth = pd; }
    }
  } else {
    int m = d->obj_geom[ob];
    const double *V = d->mesh_verts + 3 * d->mesh_vert_off[m];
    const int32_t *Tr = d->mesh_tris + 3 * d->mesh_tri_off[m];
    int nt = d->mesh_tri_off[m + 1] - d->mesh_tri_off[m];
    for (int t = 0; t < nt; t++) {
      shape_t tri; memset(&tri, 0, sizeof(tri)); tri.type = ORC_TRI; tri.T = *Tb;
      copy3(tri.tri, V + 3 * Tr[3 * t]); copy3(tri.tri + 3, V + 3 * Tr[3 * t + 1]); copy3(tri.tri + 6, V + 3 * Tr[3 * t + 2]);
      double g = gjk_distance(&S, &tri);
      if (g > 0) { if (g < mind) mind = g; }
      else { double pd = mpr_penetration_depth(&S, &tri); if (pd < 0) pd = 0; if (pd > maxdepth) maxdepth = pd; }
    }
  }
  if (maxdepth >= 0) return -maxdepth;
  return mind;
}

int orc_pair_signed_distance(const mpb_scene_desc *d, const double *state, int pair, double *out) {
  xf_t *oMi = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_joints + 1));
  xf_t *objT = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_objects > 0 ? d->n_objects : 1));
  double *qf = (double *)malloc(sizeof(double) * (size_t)(d->nq_full > 0 ? d->nq_full : 1));
  scatter_state(d, state, qf); fk_joints(d, qf, oMi); object_transforms(d, oMi, objT, NULL);
  int a = d->pair_a[pair], b = d->pair_b[pair];
  int ta = d->obj_type[a], tb = d->obj_type[b];
  if (ta >= MPB_GEOM_MESH && tb >= MPB_GEOM_MESH) { *out = NAN; }
  else if (ta >= MPB_GEOM_MESH) *out = leafset_signed_distance(d, a, &objT[a], b, &objT[b]);
  else if (tb >= MPB_GEOM_MESH) *out = leafset_signed_distance(d, b, &objT[b], a, &objT[a]);
  else {
    shape_t A, B; shape_from_object(d, a, &objT[a], &A); shape_from_object(d, b, &objT[b], &B);
    double g = gjk_distance(&A, &B);
    if (g > 0) *out = g;
    else { double pd = mpr_penetration_depth(&A, &B); *out = pd < 0 ? 0.0 : -pd; }
  }
  free(oMi); free(objT); free(qf);
  return 0;
}

/* ---- ::fcl::distance for pairs that involve a BVHModel or an OcTree (distance column): the minimum over
 * the leaf primitives of the leaf distance -- what MeshShapeDistanceTraversalNode / OcTreeSolver::OcTree*DistanceRecurse /
 * MeshDistanceTraversalNode compute with their bounding-volume pruning [upstream FCL 0.7.0].  Leaf distances:
 *   shape x triangle, box(voxel) x shape, box(voxel) x triangle : GJK (shapeTriangleDistance / shapeDistance), -1 when
 *                                                                touching or intersecting (signed distance disabled)
 *   triangle x triangle (mesh x mesh)                          : TriangleDistance::triDistance, which returns 0 for
 *                                                                overlapping triangles (not -1)
 * Exhaustive over the leaves; a leaf is skipped only when a bounding-sphere lower bound already exceeds the running
 * minimum (cannot change the result). */
static void tri_shape(shape_t *t, const xf_t *T, const double *P0, const double *P1, const double *P2) {
  memset(t, 0, sizeof(*t)); t->type = ORC_TRI; t->T = *T;
  copy3(t->tri, P0); copy3(t->tri + 3, P1); copy3(t->tri + 6, P2);
}
static double tri_bound(const double *a, const double *b, const double *c, double *ctr) {
  for (int k = 0; k < 3; k++) ctr[k] = (a[k] + b[k] + c[k]) / 3.0;
  double d[3], r = 0, l;
  sub3(d, a, ctr); l = len2_3(d); if (l > r) r = l;
  sub3(d, b, ctr); l = len2_3(d); if (l > r) r = l;
  sub3(d, c, ctr); l = len2_3(d); if (l > r) r = l;
  return sqrt(r);
}
static double bvh_object_distance(const mpb_scene_desc *d, int a, const xf_t *Ta, int b, const xf_t *Tb) {
  int ta = d->obj_type[a], tb = d->obj_type[b];
  double best = DBL_MAX;
  if (ta >= MPB_GEOM_MESH && tb >= MPB_GEOM_MESH) {
    if (ta == MPB_GEOM_OCTREE && tb == MPB_GEOM_OCTREE) return DBL_MAX; /* not produced by the path */
    if (ta == MPB_GEOM_OCTREE) { int t = a; a = b; b = t; const xf_t *T = Ta; Ta = Tb; Tb = T; tb = MPB_GEOM_OCTREE; }
    /* a = mesh; b = mesh or octree */
    int ma = d->obj_geom[a];
    const double *Va = d->mesh_verts + 3 * d->mesh_vert_off[ma];
    const int32_t *Ia = d->mesh_tris + 3 * d->mesh_tri_off[ma];
    int na = d->mesh_tri_off[ma + 1] - d->mesh_tri_off[ma];
    for (int i = 0; i < na; i++) {
      double A[9], ca[3];
      for (int k = 0; k < 3; k++) xf_apply(A + 3 * k, Ta, Va + 3 * Ia[3 * i + k]);
      double ra = tri_bound(A, A + 3, A + 6, ca);
      shape_t tA; xf_t I; xf_identity(&I); tri_shape(&tA, &I, A, A + 3, A + 6);
      if (tb == MPB_GEOM_MESH) {
        int mb = d->obj_geom[b];
        const double *Vb = d->mesh_verts + 3 * d->mesh_vert_off[mb];
        const int32_t *Ib = d->mesh_tris + 3 * d->mesh_tri_off[mb];
        int nb = d->mesh_tri_off[mb + 1] - d->mesh_tri_off[mb];
        for (int j = 0; j < nb; j++) {
          double B[9], cb[3], dd[3];
          for (int k = 0; k < 3; k++) xf_apply(B + 3 * k, Tb, Vb + 3 * Ib[3 * j + k]);
          double rb = tri_bound(B, B + 3, B + 6, cb);
          sub3(dd, ca, cb);
          if (sqrt(len2_3(dd)) - ra - rb >= best) continue;
          shape_t tB; tri_shape(&tB, &I, B, B + 3, B + 6);
          double g = gjk_distance(&tA, &tB);
          if (g < 0) g = 0.0; /* triDistance: overlapping triangles -> 0 */
          if (g < best) best = g;
        }
      } else {
        int t = d->obj_geom[b]; double res = d->oct_res[t];
        const int32_t *C = d->oct_cells + 3 * d->oct_cell_off[t];
        int nc = d->oct_cell_off[t + 1] - d->oct_cell_off[t];
        shape_t bx; memset(&bx, 0, sizeof(bx)); bx.type = MPB_GEOM_BOX; bx.a = bx.b = bx.c = 0.5 * res;
        double rb = 0.5 * res * sqrt(3.0);
        for (int j = 0; j < nc; j++) {
          double c[3] = {(C[3 * j] + 0.5) * res, (C[3 * j + 1] + 0.5) * res, (C[3 * j + 2] + 0.5) * res}, cw[3], dd[3];
          xf_apply(cw, Tb, c); sub3(dd, ca, cw);
          if (sqrt(len2_3(dd)) - ra - rb >= best && best >= 0) continue;
          bx.T = *Tb; copy3(bx.T.p, cw);
          double g = gjk_distance(&bx, &tA);
          if (g < best) best = g;
          if (best < 0) return -1.0;
        }
      }
    }
    return best;
  }
  /* shape x mesh / octree */
  int os = ta >= MPB_GEOM_MESH ? b : a, ob = ta >= MPB_GEOM_MESH ? a : b;
  const xf_t *Ts = ta >= MPB_GEOM_MESH ? Tb : Ta, *To = ta >= MPB_GEOM_MESH ? Ta : Tb;
  shape_t S; shape_from_object(d, os, Ts, &S);
  double cl[3]; double rad = shape_bound_radius(d, os, cl);
  if (d->obj_type[ob] == MPB_GEOM_OCTREE) {
    int t = d->obj_geom[ob]; double res = d->oct_res[t];
    const int32_t *C = d->oct_cells + 3 * d->oct_cell_off[t];
    int nc = d->oct_cell_off[t + 1] - d->oct_cell_off[t];
    shape_t bx; memset(&bx, 0, sizeof(bx)); bx.type = MPB_GEOM_BOX; bx.a = bx.b = bx.c = 0.5 * res;
    double rb = 0.5 * res * sqrt(3.0);
    for (int j = 0; j < nc; j++) {
      double c[3] = {(C[3 * j] + 0.5) * res, (C[3 * j + 1] + 0.5) * res, (C[3 * j + 2] + 0.5) * res}, cw[3], dd[3];
      xf_apply(cw, To, c); sub3(dd, Ts->p, cw);
      if (sqrt(len2_3(dd)) - rad - rb >= best) continue;
      bx.T = *To; copy3(bx.T.p, cw);
      double g = gjk_distance(&bx, &S);
      if (g < best) best = g;
      if (best < 0) return -1.0;
    }
  } else {
    int m = d->obj_geom[ob];
    const double *V = d->mesh_verts + 3 * d->mesh_vert_off[m];
    const int32_t *Tr = d->mesh_tris + 3 * d->mesh_tri_off[m];
    int nt = d->mesh_tri_off[m + 1] - d->mesh_tri_off[m];
    for (int t = 0; t < nt; t++) {
      double A[9], ca[3], dd[3];
      for (int k = 0; k < 3; k++) xf_apply(A + 3 * k, To, V + 3 * Tr[3 * t + k]);
      double ra = tri_bound(A, A + 3, A + 6, ca);
      sub3(dd, Ts->p, ca);
      if (sqrt(len2_3(dd)) - rad - ra >= best) continue;
      shape_t tri; tri_shape(&tri, To, V + 3 * Tr[3 * t], V + 3 * Tr[3 * t + 1], V + 3 * Tr[3 * t + 2]);
      double g = gjk_distance(&S, &tri);
      if (g < best) best = g;
      if (best < 0) return -1.0;
    }
  }
  return best;
}

/* closest distance between two segments (Ericson 5.1.9) and the closed-form triangle-triangle distance built from it
 * (6 vertex-face + 9 edge-edge candidates; valid when the triangles are disjoint): an INDEPENDENT check of the GJK leaf
 * distance above (tests/test_oracle_distance.py). */
static double seg_seg_dist2(const double *p1, const double *q1, const double *p2, const double *q2) {
  double d1[3], d2[3], r[3]; sub3(d1, q1, p1); sub3(d2, q2, p2); sub3(r, p1, p2);
  double a = len2_3(d1), e = len2_3(d2), f = dot3(d2, r), s, t;
  if (a <= 1e-300 && e <= 1e-300) return len2_3(r);
  if (a <= 1e-300) { s = 0; t = f / e; t = t < 0 ? 0 : (t > 1 ? 1 : t); }
  else {
    double c = dot3(d1, r);
    if (e <= 1e-300) { t = 0; s = -c / a; s = s < 0 ? 0 : (s > 1 ? 1 : s); }
    else {
      double b = dot3(d1, d2), den = a * e - b * b;
      s = den != 0 ? (b * f - c * e) / den : 0; s = s < 0 ? 0 : (s > 1 ? 1 : s);
      t = (b * s + f) / e;
      if (t < 0) { t = 0; s = -c / a; s = s < 0 ? 0 : (s > 1 ? 1 : s); }
      else if (t > 1) { t = 1; s = (b - c) / a; s = s < 0 ? 0 : (s > 1 ? 1 : s); }
    }
  }
  double c1[3] = {p1[0] + d1[0] * s, p1[1] + d1[1] * s, p1[2] + d1[2] * s}, c2[3] = {p2[0] + d2[0] * t, p2[1] + d2[1] * t, p2[2] + d2[2] * t}, dd[3];
  sub3(dd, c1, c2); return len2_3(dd);
}
double orc_tri_tri_distance_closed(const double *A, const double *B) {
  double best = DBL_MAX;
  for (int i = 0; i < 3; i++) {
    double v = point_tri_dist2(A + 3 * i, B, B + 3, B + 6, NULL); if (v < best) best = v;
    v = point_tri_dist2(B + 3 * i, A, A + 3, A + 6, NULL); if (v < best) best = v;
    for (int j = 0; j < 3; j++) { v = seg_seg_dist2(A + 3 * i, A + 3 * ((i + 1) % 3), B + 3 * j, B + 3 * ((j + 1) % 3)); if (v < best) best = v; }
  }
  return sqrt(best);
}
double orc_tri_tri_distance_gjk(const double *A, const double *B) {
  shape_t tA, tB; xf_t I; xf_identity(&I);
  tri_shape(&tA, &I, A, A + 3, A + 6); tri_shape(&tB, &I, B, B + 3, B + 6);
  return gjk_distance(&tA, &tB);
}
int orc_tri_tri_intersect(const double *A, const double *B) { return tri_tri(A, B); }
/* INDEPENDENT intersection test (shares nothing with the SAT above): an edge of one triangle pierces the other
 * (segment / plane intersection + barycentric containment), or -- coplanar and touching configurations -- one of the
 * vertex-face / edge-edge distance candidates is zero. */
static int seg_pierces_tri(const double *p, const double *q, const double *a, const double *b, const double *c) {
  double ab[3], ac[3], n[3], pq[3], ap[3];
  sub3(ab, b, a); sub3(ac, c, a); cross3(n, ab, ac); sub3(pq, q, p); sub3(ap, p, a);
  double dn = dot3(n, pq), t = -dot3(n, ap);
  if (dn == 0) return 0;
  t /= dn; if (t < 0 || t > 1) return 0;
  double x[3] = {p[0] + t * pq[0] - a[0], p[1] + t * pq[1] - a[1], p[2] + t * pq[2] - a[2]};
  double d00 = dot3(ab, ab), d01 = dot3(ab, ac), d11 = dot3(ac, ac), d20 = dot3(x, ab), d21 = dot3(x, ac);
  double den = d00 * d11 - d01 * d01; if (den == 0) return 0;
  double v = (d11 * d20 - d01 * d21) / den, w = (d00 * d21 - d01 * d20) / den;
  return v >= 0 && w >= 0 && v + w <= 1;
}
int orc_tri_tri_intersect_independent(const double *A, const double *B, double zero_tol) {
  for (int e = 0; e < 3; e++) {
    if (seg_pierces_tri(A + 3 * e, A + 3 * ((e + 1) % 3), B, B + 3, B + 6)) return 1;
    if (seg_pierces_tri(B + 3 * e, B + 3 * ((e + 1) % 3), A, A + 3, A + 6)) return 1;
  }
  return orc_tri_tri_distance_closed(A, B) <= zero_tol;
}

/* ::fcl::distance of one admitted pair in world poses */
static double object_distance(const mpb_scene_desc *d, int a, const xf_t *Ta, int b, const xf_t *Tb) {
  if (d->obj_type[a] >= MPB_GEOM_MESH || d->obj_type[b] >= MPB_GEOM_MESH) return bvh_object_distance(d, a, Ta, b, Tb);
  shape_t A, B; shape_from_object(d, a, Ta, &A); shape_from_object(d, b, Tb, &B);
  return gjk_distance(&A, &B);
}

/* min over admitted pairs of ::fcl::distance (distanceFull, src/planning_world.cpp:714-720); -1 for touching /
 * penetrating pairs as FCL returns with signed distance disabled (0 for overlapping mesh x mesh, see above).
 * pair_dist: nullable [n][n_pairs], the distance of every pair. */
int orc_distance_batch_ex(const mpb_scene_desc *d, const double *states, int64_t n, double *min_dist, int32_t *argmin_pair, double *pair_dist, int nthreads) {
  if (nthreads < 1) nthreads = 1;
#pragma omp parallel num_threads(nthreads)
  {
    xf_t *oMi = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_joints + 1));
    xf_t *objT = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_objects > 0 ? d->n_objects : 1));
    double *qf = (double *)malloc(sizeof(double) * (size_t)(d->nq_full > 0 ? d->nq_full : 1));
#pragma omp for schedule(dynamic, 1)
    for (int64_t i = 0; i < n; i++) {
      scatter_state(d, states + i * d->state_dim, qf); fk_joints(d, qf, oMi); object_transforms(d, oMi, objT, NULL);
      double best = DBL_MAX; int bi = -1; /* WorldDistanceResult::min_distance starts at max (src/planning_world.h:30-41) */
      for (int p = 0; p < d->n_pairs; p++) {
        int a = d->pair_a[p], b = d->pair_b[p];
        double g = object_distance(d, a, &objT[a], b, &objT[b]);
        if (pair_dist) pair_dist[i * d->n_pairs + p] = g;
        if (g < best) { best = g; bi = p; }
      }
      min_dist[i] = best; if (argmin_pair) argmin_pair[i] = bi;
    }
    free(oMi); free(objT); free(qf);
  }
  return 0;
}
int orc_distance_batch(const mpb_scene_desc *d, const double *states, int64_t n, double *min_dist, int32_t *argmin_pair) {
  return orc_distance_batch_ex(d, states, n, min_dist, argmin_pair, NULL, 1);
}

/* ------------------------------------------------------------------ independent exact check: SAT on polytopes
 * Signed separation of two convex polytopes (vertices + triangle faces) in world poses: the maximum over
 * all candidate axes (face normals of A, of B, edge x edge) of the projection gap.  > 0 => separated by at
 * least that much (a lower bound of the distance), < 0 => exact penetration depth.  Shares no code path
 * with MPR/GJK above. */
static void project_hull(const double *W, int nv, const double *ax, double *lo, double *hi) {
  double l = DBL_MAX, h = -DBL_MAX;
  for (int i = 0; i < nv; i++) { double t = dot3(W + 3 * i, ax); if (t < l) l = t; if (t > h) h = t; }
  *lo = l; *hi = h;
}
static double sat_axis_gap(const double *WA, int na, const double *WB, int nb, double *ax) {
  double l2 = len2_3(ax);
  if (l2 < 1e-24) return -DBL_MAX;
  scale3(ax, 1.0 / sqrt(l2));
  double la, ha, lb, hb; project_hull(WA, na, ax, &la, &ha); project_hull(WB, nb, ax, &lb, &hb);
  double g1 = lb - ha, g2 = la - hb;
  return g1 > g2 ? g1 : g2;
}
int orc_sat_pair(const mpb_scene_desc *d, const double *state, int pair, double *gap_out) {
  int a = d->pair_a[pair], b = d->pair_b[pair];
  if (d->obj_type[a] != MPB_GEOM_CONVEX || d->obj_type[b] != MPB_GEOM_CONVEX) return MPB_ERR_UNSUPPORTED;
  xf_t *oMi = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_joints + 1));
  xf_t *objT = (xf_t *)malloc(sizeof(xf_t) * (size_t)d->n_objects);
  double *qf = (double *)malloc(sizeof(double) * (size_t)d->nq_full);
  scatter_state(d, state, qf); fk_joints(d, qf, oMi); object_transforms(d, oMi, objT, NULL);
  int ha = d->obj_geom[a], hb = d->obj_geom[b];
  int na = d->hull_vert_off[ha + 1] - d->hull_vert_off[ha], nb = d->hull_vert_off[hb + 1] - d->hull_vert_off[hb];
  int fa = d->hull_face_off[ha + 1] - d->hull_face_off[ha], fb = d->hull_face_off[hb + 1] - d->hull_face_off[hb];
  const double *VA = d->hull_verts + 3 * d->hull_vert_off[ha], *VB = d->hull_verts + 3 * d->hull_vert_off[hb];
  const int32_t *FA = d->hull_faces + 3 * d->hull_face_off[ha], *FB = d->hull_faces + 3 * d->hull_face_off[hb];
  double *WA = (double *)malloc(sizeof(double) * 3 * (size_t)na), *WB = (double *)malloc(sizeof(double) * 3 * (size_t)nb);
  for (int i = 0; i < na; i++) xf_apply(WA + 3 * i, &objT[a], VA + 3 * i);
  for (int i = 0; i < nb; i++) xf_apply(WB + 3 * i, &objT[b], VB + 3 * i);
  double best = -DBL_MAX;
  for (int f = 0; f < fa; f++) {
    double e1[3], e2[3], ax[3]; sub3(e1, WA + 3 * FA[3 * f + 1], WA + 3 * FA[3 * f]); sub3(e2, WA + 3 * FA[3 * f + 2], WA + 3 * FA[3 * f]);
    cross3(ax, e1, e2); double g = sat_axis_gap(WA, na, WB, nb, ax); if (g > best) best = g;
  }
  for (int f = 0; f < fb; f++) {
    double e1[3], e2[3], ax[3]; sub3(e1, WB + 3 * FB[3 * f + 1], WB + 3 * FB[3 * f]); sub3(e2, WB + 3 * FB[3 * f + 2], WB + 3 * FB[3 * f]);
    cross3(ax, e1, e2); double g = sat_axis_gap(WA, na, WB, nb, ax); if (g > best) best = g;
  }
  for (int f = 0; f < fa; f++) for (int e = 0; e < 3; e++) {
    int i0 = FA[3 * f + e], i1 = FA[3 * f + (e + 1) % 3];
    if (i0 > i1) continue; /* each undirected edge once */
    double ea[3]; sub3(ea, WA + 3 * i1, WA + 3 * i0);
    for (int g2 = 0; g2 < fb; g2++) for (int e2 = 0; e2 < 3; e2++) {
      int j0 = FB[3 * g2 + e2], j1 = FB[3 * g2 + (e2 + 1) % 3];
      if (j0 > j1) continue;
      double eb[3], ax[3]; sub3(eb, WB + 3 * j1, WB + 3 * j0); cross3(ax, ea, eb);
      double g = sat_axis_gap(WA, na, WB, nb, ax); if (g > best) best = g;
    }
  }
  *gap_out = best;
  free(WA); free(WB); free(oMi); free(objT); free(qf);
  return 0;
}

/* Standalone narrowphase on two primitive/convex objects given world poses [12] each (tests). */
int orc_object_collide(const mpb_scene_desc *d, int a, const double *Ta12, int b, const double *Tb12) {
  xf_t Ta, Tb; xf_from12(&Ta, Ta12); xf_from12(&Tb, Tb12);
  return object_collide(d, a, &Ta, b, &Tb);
}
double orc_object_distance(const mpb_scene_desc *d, int a, const double *Ta12, int b, const double *Tb12) {
  xf_t Ta, Tb; xf_from12(&Ta, Ta12); xf_from12(&Tb, Tb12);
  shape_t A, B; shape_from_object(d, a, &Ta, &A); shape_from_object(d, b, &Tb, &B);
  return gjk_distance(&A, &B);
}

/* Batched orc_sat_pair over (state index, pair) items on `nthreads` OpenMP threads (tests). */
int orc_sat_batch_mt(const mpb_scene_desc *d, const double *states, int64_t n_items, const int64_t *item_state,
                     const int32_t *item_pair, double *gap_out, int nthreads) {
  int rc = 0;
  if (nthreads < 1) nthreads = 1;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 8)
  for (int64_t k = 0; k < n_items; k++) {
    int r = orc_sat_pair(d, states + item_state[k] * d->state_dim, item_pair[k], gap_out + k);
    if (r != 0) {
#pragma omp atomic write
      rc = r;
    }
  }
  return rc;
}

/* ------------------------------------------------------------------ planning layer (CPU restatement)
 * State space of OMPLPlannerTpl::build_state_space (src/ompl_planner.cpp:245-296): 1-D RealVector subspaces
 * (and SO(2) for continuous joints), weight 1 each => distance = sum_j |delta_j|  [upstream OMPL 1.6.0
 * CompoundStateSpace::distance / interpolate / validSegmentCount].  seg[j] = 0.01 * extent_j. */
#include <time.h>
#define ORC_PI 3.14159265358979323846
static double ss_delta(double a, double b, int so2) {
  double d = b - a;
  if (so2) { if (d > ORC_PI) d -= 2 * ORC_PI; else if (d < -ORC_PI) d += 2 * ORC_PI; }
  return d;
}
static double ss_interp1(double a, double b, double t, int so2) {
  double v = a + ss_delta(a, b, so2) * t;
  if (so2) { if (v > ORC_PI) v -= 2 * ORC_PI; else if (v < -ORC_PI) v += 2 * ORC_PI; }
  return v;
}
static double ss_distance(int D, const uint8_t *so2, const double *a, const double *b) {
  double d = 0; for (int j = 0; j < D; j++) d += fabs(ss_delta(a[j], b[j], so2 ? so2[j] : 0)); return d;
}
static int ss_segment_count(int D, const double *seg, const uint8_t *so2, const double *a, const double *b) {
  int nd = 0;
  for (int j = 0; j < D; j++) { int c = (int)ceil(fabs(ss_delta(a[j], b[j], so2 ? so2[j] : 0)) / seg[j]); if (c > nd) nd = c; }
  return nd;
}

typedef struct { const mpb_scene_desc *d; const orc_accel *ac; xf_t *oMi; xf_t *objT; double *qf; int64_t checks; } orc_checker;
static void checker_init(orc_checker *c, const mpb_scene_desc *d) {
  c->d = d; c->checks = 0; c->ac = NULL;
  c->oMi = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_joints + 1));
  c->objT = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_objects > 0 ? d->n_objects : 1));
  c->qf = (double *)malloc(sizeof(double) * (size_t)(d->nq_full > 0 ? d->nq_full : 1));
}
static void checker_free(orc_checker *c) { free(c->oMi); free(c->objT); free(c->qf); }
/* ValidityCheckerTpl::isValid (src/ompl_planner.h:59-62): setQposAll + !collide(); collide() has no early exit
 * (src/planning_world.h:248-250), so every admitted pair is tested. */
static int checker_is_valid(orc_checker *c, const double *state) {
  const mpb_scene_desc *d = c->d;
  scatter_state(d, state, c->qf); fk_joints(d, c->qf, c->oMi); object_transforms(d, c->oMi, c->objT, NULL);
  int any = 0;
  for (int p = 0; p < d->n_pairs; p++) { int a = d->pair_a[p], b = d->pair_b[p]; any |= object_collide_ac(d, c->ac, a, &c->objT[a], b, &c->objT[b]); }
  c->checks++;
  return !any;
}
/* [upstream OMPL 1.6.0] DiscreteMotionValidator::checkMotion(s1, s2): s2 first, then the interior points in
 * bisection order with early exit. */
static int checker_motion(orc_checker *c, int D, const double *seg, const uint8_t *so2, const double *s1, const double *s2) {
  if (!checker_is_valid(c, s2)) return 0;
  int nd = ss_segment_count(D, seg, so2, s1, s2);
  int result = 1;
  if (nd >= 2) {
    int *queue = (int *)malloc(sizeof(int) * 2 * (size_t)(nd + 2));
    double *test = (double *)malloc(sizeof(double) * (size_t)D);
    int head = 0, tail = 0;
    queue[2 * tail] = 1; queue[2 * tail + 1] = nd - 1; tail++;
    while (head < tail) {
      int lo = queue[2 * head], hi = queue[2 * head + 1];
      int mid = (lo + hi) / 2;
      for (int j = 0; j < D; j++) test[j] = ss_interp1(s1[j], s2[j], (double)mid / (double)nd, so2 ? so2[j] : 0);
      if (!checker_is_valid(c, test)) { result = 0; break; }
      head++;
      if (lo < mid) { queue[2 * tail] = lo; queue[2 * tail + 1] = mid - 1; tail++; }
      if (hi > mid) { queue[2 * tail] = mid + 1; queue[2 * tail + 1] = hi; tail++; }
    }
    free(queue); free(test);
  }
  return result;
}

int orc_check_motion_batch(const mpb_scene_desc *d, int D, const double *seg, const uint8_t *so2, const double *a, const double *b,
                           int64_t n_edges, uint8_t *valid, int64_t *n_states, int nthreads) {
  if (D != d->state_dim) return MPB_ERR_INVALID;
  int64_t total = 0;
  if (nthreads < 1) nthreads = 1;
#pragma omp parallel for num_threads(nthreads) schedule(static) reduction(+ : total)
  for (int t = 0; t < nthreads; t++) {
    orc_checker c; checker_init(&c, d);
    for (int64_t e = n_edges * t / nthreads; e < n_edges * (t + 1) / nthreads; e++)
      valid[e] = (uint8_t)checker_motion(&c, D, seg, so2, a + e * D, b + e * D);
    total += c.checks;
    checker_free(&c);
  }
  if (n_states) *n_states = total;
  return 0;
}

/* ---- RRTConnect [upstream OMPL 1.6.0 geometric/planners/rrt/src/RRTConnect.cpp], as OMPLPlannerTpl::plan runs it
 * (src/ompl_planner.cpp:160-165, 217-221): no path simplification, GoalStates goal, brute-force nearest. */
typedef struct { double *st; int *parent; int n, cap; } orc_tree;
static int tree_add(orc_tree *t, int D, const double *s, int parent) {
  if (t->n == t->cap) { t->cap = t->cap ? 2 * t->cap : 256; t->st = (double *)realloc(t->st, sizeof(double) * (size_t)t->cap * D); t->parent = (int *)realloc(t->parent, sizeof(int) * (size_t)t->cap); }
  memcpy(t->st + (size_t)t->n * D, s, sizeof(double) * (size_t)D); t->parent[t->n] = parent; return t->n++;
}
static uint64_t orc_rng_next(uint64_t *s) { uint64_t z = (*s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }
static double orc_rng_uniform(uint64_t *s) { return (double)(orc_rng_next(s) >> 11) * (1.0 / 9007199254740992.0); }

typedef struct { int D; const double *lo, *hi, *seg; const uint8_t *so2; double range; orc_checker *chk; double *xstate; int xmotion; } orc_rrt_ctx;
enum { ORC_TRAPPED = 0, ORC_ADVANCED = 1, ORC_REACHED = 2 };
static int grow_tree(orc_rrt_ctx *c, orc_tree *tree, int is_start, const double *rstate) {
  int D = c->D, nm = 0; double best = DBL_MAX;
  for (int i = 0; i < tree->n; i++) { double dd = ss_distance(D, c->so2, tree->st + (size_t)i * D, rstate); if (dd < best) { best = dd; nm = i; } }
  const double *ns = tree->st + (size_t)nm * D;
  const double *dstate = rstate; int reach = 1;
  if (best > c->range) {
    for (int j = 0; j < D; j++) c->xstate[j] = ss_interp1(ns[j], rstate[j], c->range / best, c->so2 ? c->so2[j] : 0);
    int same = 1; for (int j = 0; j < D; j++) if (c->xstate[j] != ns[j]) same = 0;
    if (same) return ORC_TRAPPED;
    dstate = c->xstate; reach = 0;
  }
  int valid = is_start ? checker_motion(c->chk, D, c->seg, c->so2, ns, dstate)
                       : (checker_is_valid(c->chk, dstate) && checker_motion(c->chk, D, c->seg, c->so2, dstate, ns));
  if (!valid) return ORC_TRAPPED;
  double *tmp = (double *)malloc(sizeof(double) * (size_t)D); memcpy(tmp, dstate, sizeof(double) * (size_t)D);
  c->xmotion = tree_add(tree, D, tmp, nm); free(tmp);
  return reach ? ORC_REACHED : ORC_ADVANCED;
}

/* Returns 1 = "Exact solution", 0 = timeout / iteration budget exhausted, 3 = no valid goal.
 * path_out: [max_path][D]; *path_len = 0 unless solved and it fits. */
static int rrtconnect_ac(const mpb_scene_desc *d, const orc_accel *ac, int D, const double *lo, const double *hi, const double *seg, const uint8_t *so2,
                   const double *start, const double *goals, int n_goals, double range, int64_t max_grow_calls, double time_limit_s,
                   uint64_t seed, double *path_out, int max_path, int *path_len, int64_t *grow_calls, int64_t *state_checks) {
  if (D != d->state_dim) return MPB_ERR_INVALID;
  orc_checker chk; checker_init(&chk, d); chk.ac = ac;
  orc_rrt_ctx c; c.D = D; c.lo = lo; c.hi = hi; c.seg = seg; c.so2 = so2; c.range = range; c.chk = &chk;
  c.xstate = (double *)malloc(sizeof(double) * (size_t)D); c.xmotion = -1;
  if (!(range > 1e-6)) { double ext = 0; for (int j = 0; j < D; j++) ext += (so2 && so2[j]) ? ORC_PI : hi[j] - lo[j]; c.range = 0.2 * ext; }
  orc_tree T[2]; memset(T, 0, sizeof(T));
  tree_add(&T[0], D, start, -1);
  double *rstate = (double *)malloc(sizeof(double) * (size_t)D);
  int goals_used = 0, start_tree = 1, solved = 0; int64_t calls = 0;
  struct timespec t0; clock_gettime(CLOCK_MONOTONIC, &t0);
  *path_len = 0;
  int status = 0;
  while (calls < max_grow_calls) {
    if (time_limit_s > 0) { struct timespec t1; clock_gettime(CLOCK_MONOTONIC, &t1); if ((t1.tv_sec - t0.tv_sec) + 1e-9 * (t1.tv_nsec - t0.tv_nsec) > time_limit_s) break; }
    orc_tree *tree = start_tree ? &T[0] : &T[1]; int tgi_start = start_tree; start_tree = !start_tree;
    orc_tree *other = start_tree ? &T[0] : &T[1];
    if (T[1].n == 0 || goals_used < T[1].n / 2) {
      while (goals_used < n_goals) { /* pis_.nextGoal(): invalid goal states are skipped */
        const double *g = goals + (size_t)goals_used * D; goals_used++;
        if (checker_is_valid(&chk, g)) { tree_add(&T[1], D, g, -1); break; }
      }
      if (T[1].n == 0) { status = 3; break; }
    }
    for (int j = 0; j < D; j++) rstate[j] = lo[j] + (hi[j] - lo[j]) * orc_rng_uniform(&seed);
    int gs = grow_tree(&c, tree, tgi_start, rstate); calls++;
    if (gs != ORC_TRAPPED) {
      int added = c.xmotion;
      if (gs != ORC_REACHED) memcpy(rstate, tree->st + (size_t)added * D, sizeof(double) * (size_t)D);
      int other_is_start = start_tree;
      int gsc = ORC_ADVANCED;
      while (gsc == ORC_ADVANCED && calls < max_grow_calls) { gsc = grow_tree(&c, other, other_is_start, rstate); calls++; }
      if (gsc == ORC_REACHED) {
        int is = other_is_start ? c.xmotion : added, ig = other_is_start ? added : c.xmotion;
        if (T[0].parent[is] >= 0) is = T[0].parent[is]; else ig = T[1].parent[ig];
        int ls = 0, lg = 0;
        for (int i = is; i >= 0; i = T[0].parent[i]) ls++;
        for (int i = ig; i >= 0; i = T[1].parent[i]) lg++;
        if (ls + lg <= max_path) {
          int k = ls - 1;
          for (int i = is; i >= 0; i = T[0].parent[i], k--) memcpy(path_out + (size_t)k * D, T[0].st + (size_t)i * D, sizeof(double) * (size_t)D);
          k = ls;
          for (int i = ig; i >= 0; i = T[1].parent[i], k++) memcpy(path_out + (size_t)k * D, T[1].st + (size_t)i * D, sizeof(double) * (size_t)D);
          *path_len = ls + lg;
        }
        solved = 1; status = 1; break;
      }
    }
  }
  (void)solved;
  if (grow_calls) *grow_calls = calls;
  if (state_checks) *state_checks = chk.checks;
  free(T[0].st); free(T[0].parent); free(T[1].st); free(T[1].parent); free(rstate); free(c.xstate); checker_free(&chk);
  return status;
}

int orc_rrtconnect(const mpb_scene_desc *d, int D, const double *lo, const double *hi, const double *seg, const uint8_t *so2,
                   const double *start, const double *goals, int n_goals, double range, int64_t max_grow_calls, double time_limit_s,
                   uint64_t seed, double *path_out, int max_path, int *path_len, int64_t *grow_calls, int64_t *state_checks) {
  return rrtconnect_ac(d, NULL, D, lo, hi, seg, so2, start, goals, n_goals, range, max_grow_calls, time_limit_s, seed, path_out, max_path, path_len,
                       grow_calls, state_checks);
}

/* Q independent queries, one per OpenMP thread at a time (the all-host-core CPU baseline of BASELINE.md 3).
 * ac: NULL = exhaustive leaf tests, else the hierarchies of orc_accel_create (the timed baseline). */
int orc_rrtconnect_batch_accel(const mpb_scene_desc *d, const orc_accel *ac, int D, const double *lo, const double *hi, const double *seg, const uint8_t *so2,
                         const double *starts, const double *goals, const int32_t *n_goals, int max_goals, int64_t n_queries, double range,
                         int64_t max_grow_calls, double time_limit_s, uint64_t seed, int32_t *status, int32_t *path_len, double *paths,
                         int max_path, int64_t *grow_calls, int64_t *state_checks, int nthreads) {
  int64_t gc = 0, sc = 0;
  if (nthreads < 1) nthreads = 1;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1) reduction(+ : gc, sc)
  for (int64_t q = 0; q < n_queries; q++) {
    int pl = 0; int64_t g1 = 0, s1 = 0;
    status[q] = rrtconnect_ac(d, ac, D, lo, hi, seg, so2, starts + q * D, goals + (size_t)q * max_goals * D, n_goals[q], range, max_grow_calls,
                              time_limit_s, seed + 7919ull * (uint64_t)q, paths + (size_t)q * max_path * D, max_path, &pl, &g1, &s1);
    path_len[q] = pl; gc += g1; sc += s1;
  }
  if (grow_calls) *grow_calls = gc;
  if (state_checks) *state_checks = sc;
  return 0;
}
int orc_rrtconnect_batch(const mpb_scene_desc *d, int D, const double *lo, const double *hi, const double *seg, const uint8_t *so2,
                         const double *starts, const double *goals, const int32_t *n_goals, int max_goals, int64_t n_queries, double range,
                         int64_t max_grow_calls, double time_limit_s, uint64_t seed, int32_t *status, int32_t *path_len, double *paths,
                         int max_path, int64_t *grow_calls, int64_t *state_checks, int nthreads) {
  int64_t gc = 0, sc = 0;
  if (nthreads < 1) nthreads = 1;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1) reduction(+ : gc, sc)
  for (int64_t q = 0; q < n_queries; q++) {
    int pl = 0; int64_t g1 = 0, s1 = 0;
    status[q] = orc_rrtconnect(d, D, lo, hi, seg, so2, starts + q * D, goals + (size_t)q * max_goals * D, n_goals[q], range, max_grow_calls,
                               time_limit_s, seed + 7919ull * (uint64_t)q, paths + (size_t)q * max_path * D, max_path, &pl, &g1, &s1);
    path_len[q] = pl; gc += g1; sc += s1;
  }
  if (grow_calls) *grow_calls = gc;
  if (state_checks) *state_checks = sc;
  return 0;
}

/* ------------------------------------------------------------------ inverse kinematics */
/* [upstream Pinocchio 2.6] log6 of an SE3: motion (v, w) */
static void log6_xf(const xf_t *M, double *o6) {
  const double *R = M->R;
  double tr = R[0] + R[4] + R[8];
  double theta = tr >= 3.0 ? 0.0 : (tr <= -1.0 ? M_PI : acos((tr - 1.0) * 0.5));
  double w[3];
  if (theta >= M_PI - 1e-2) {
    double cphi = -(tr - 1.0) * 0.5, beta = theta * theta / (1.0 + cphi);
    double t0 = (R[0] + cphi) * beta, t1 = (R[4] + cphi) * beta, t2 = (R[8] + cphi) * beta;
    w[0] = (R[7] > R[5] ? 1.0 : -1.0) * (t0 > 0 ? sqrt(t0) : 0.0);
    w[1] = (R[2] > R[6] ? 1.0 : -1.0) * (t1 > 0 ? sqrt(t1) : 0.0);
    w[2] = (R[3] > R[1] ? 1.0 : -1.0) * (t2 > 0 ? sqrt(t2) : 0.0);
  } else {
    double t = (theta > 1e-8 ? theta / sin(theta) : 1.0) * 0.5;
    w[0] = t * (R[7] - R[5]); w[1] = t * (R[2] - R[6]); w[2] = t * (R[3] - R[1]);
  }
  double t2 = theta * theta, alpha, beta;
  if (theta < 1e-4) { alpha = 1.0 - t2 / 12.0 - t2 * t2 / 720.0; beta = 1.0 / 12.0 + t2 / 720.0; }
  else { double st = sin(theta), ct = cos(theta); alpha = theta * st / (2.0 * (1.0 - ct)); beta = 1.0 / t2 - st / (2.0 * theta * (1.0 - ct)); }
  const double *p = M->p;
  double wp = dot3(w, p), c[3]; cross3(c, w, p);
  for (int k = 0; k < 3; k++) { o6[k] = alpha * p[k] - 0.5 * c[k] + beta * wp * w[k]; o6[3 + k] = w[k]; }
}
static void xf_inv_mul(xf_t *o, const xf_t *a, const xf_t *b) { /* a^-1 b */
  double d[3]; sub3(d, b->p, a->p);
  for (int i = 0; i < 3; i++) {
    for (int j = 0; j < 3; j++) o->R[3 * i + j] = a->R[i] * b->R[j] + a->R[3 + i] * b->R[3 + j] + a->R[6 + i] * b->R[6 + j];
    o->p[i] = a->R[i] * d[0] + a->R[3 + i] * d[1] + a->R[6 + i] * d[2];
  }
}
static void solve6(double A[6][6], double *b) {
  for (int c = 0; c < 6; c++) {
    int piv = c; double best = fabs(A[c][c]);
    for (int r = c + 1; r < 6; r++) if (fabs(A[r][c]) > best) { best = fabs(A[r][c]); piv = r; }
    if (piv != c) { for (int k = 0; k < 6; k++) { double t = A[c][k]; A[c][k] = A[piv][k]; A[piv][k] = t; } double t = b[c]; b[c] = b[piv]; b[piv] = t; }
    double d = A[c][c]; if (d == 0.0) continue;
    for (int r = c + 1; r < 6; r++) { double f = A[r][c] / d; for (int k = c; k < 6; k++) A[r][k] -= f * A[c][k]; b[r] -= f * b[c]; }
  }
  for (int c = 5; c >= 0; c--) { double s = b[c]; for (int k = c + 1; k < 6; k++) s -= A[c][k] * b[k]; b[c] = A[c][c] != 0.0 ? s / A[c][c] : 0.0; }
}
/* PinocchioModelTpl::computeIKCLIK (src/pinocchio_model.cpp:377-432) over the scene's batch state */
int orc_ik_clik_batch(const mpb_scene_desc *d, int link, const double *poses, const double *q_init, const uint8_t *mask, int64_t n,
                      double eps, int max_iter, double dt, double damp, double *q_out, uint8_t *success, double *err_out, int nthreads) {
  if (link < 0 || link >= d->n_links) return -1;
  int chain[64], nc = 0;
  for (int j = d->link_joint[link]; j != 0 && nc < 64; j = d->joint_parent[j - 1]) chain[nc++] = j;
  if (nc == 0) return -1;
  for (int a = 0, b = nc - 1; a < b; a++, b--) { int t = chain[a]; chain[a] = chain[b]; chain[b] = t; }
  int *slot_of_q = (int *)malloc(sizeof(int) * (size_t)d->nq_full);
  for (int k = 0; k < d->nq_full; k++) slot_of_q[k] = -1;
  for (int k = 0; k < d->state_dim; k++) slot_of_q[d->state_map[k]] = k;
  xf_t L, Li; xf_from12(&L, d->link_placement + 12 * link);
  xf_t I; xf_identity(&I); xf_inv_mul(&Li, &L, &I);
  const int D = d->state_dim;
  if (nthreads < 1) nthreads = 1;
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 16)
  for (int64_t i = 0; i < n; i++) {
    const double *ps = poses + 7 * i;
    double quat[4] = {ps[3], ps[4], ps[5], ps[6]};
    xf_t lp, target; quat_to_mat(quat, lp.R); copy3(lp.p, ps);
    xf_mul(&target, &lp, &Li);
    double q[64]; xf_t T[64];
    for (int c = 0; c < nc; c++) { int j = chain[c] - 1, sl = slot_of_q[d->joint_qidx[j]]; q[c] = sl >= 0 ? q_init[i * D + sl] : d->q_base[d->joint_qidx[j]]; }
    double err[6]; int ok = 0;
    for (int it = 0;; it++) {
      xf_t cur; xf_identity(&cur);
      for (int c = 0; c < nc; c++) {
        int j = chain[c] - 1; xf_t pl, jm, li, nx;
        xf_from12(&pl, d->joint_placement + 12 * j); joint_motion(&jm, d->joint_type[j], d->joint_axis + 3 * j, q[c]);
        xf_mul(&li, &pl, &jm);
        if (c == 0) cur = li; else { xf_mul(&nx, &cur, &li); cur = nx; }
        T[c] = cur;
      }
      xf_t dMi; xf_inv_mul(&dMi, &target, &cur); log6_xf(&dMi, err);
      double e2 = 0; for (int k = 0; k < 6; k++) e2 += err[k] * err[k];
      if (sqrt(e2) < eps) { ok = 1; break; }
      if (it >= max_iter) break;
      double J[6][64], A[6][6];
      for (int r = 0; r < 6; r++) for (int k = 0; k < 6; k++) A[r][k] = r == k ? damp : 0.0;
      for (int c = 0; c < nc; c++) {
        int j = chain[c] - 1, sl = slot_of_q[d->joint_qidx[j]]; double col[6] = {0, 0, 0, 0, 0, 0};
        if (sl >= 0 && !(mask && mask[sl])) {
          xf_t X; xf_inv_mul(&X, &cur, &T[c]);
          double ra[3]; xf_rot(ra, &X, d->joint_axis + 3 * j);
          if (d->joint_type[j] == MPB_JOINT_PRISMATIC) copy3(col, ra);
          else { cross3(col, X.p, ra); copy3(col + 3, ra); }
          for (int r = 0; r < 6; r++) for (int k = 0; k < 6; k++) A[r][k] += col[r] * col[k];
        }
        for (int r = 0; r < 6; r++) J[r][c] = col[r];
      }
      double y[6]; memcpy(y, err, sizeof(y)); solve6(A, y);
      for (int c = 0; c < nc; c++) { double v = 0; for (int r = 0; r < 6; r++) v += J[r][c] * y[r]; q[c] -= v * dt; }
    }
    for (int s = 0; s < D; s++) q_out[i * D + s] = q_init[i * D + s];
    for (int c = 0; c < nc; c++) {
      int j = chain[c] - 1, sl = slot_of_q[d->joint_qidx[j]];
      if (sl >= 0) q_out[i * D + sl] = d->joint_type[j] == MPB_JOINT_CONTINUOUS ? atan2(sin(q[c]), cos(q[c])) : q[c];
    }
    success[i] = (uint8_t)ok;
    memcpy(err_out + 6 * i, err, sizeof(err));
  }
  free(slot_of_q);
  return 0;
}

/* PinocchioModelTpl::computeFullJacobian + getLinkJacobian (src/pinocchio_model.cpp:335-360): Pinocchio's
 * computeJointJacobians fills data.J with column j = oMi[j].act(S_j) (the joint's motion subspace seen from the world
 * origin: (p x a, a) for a revolute joint with origin p and world axis a, (a, 0) for a prismatic one); getJointJacobian
 * (WORLD) keeps the columns of the joints supporting `link`'s parent joint; local: link2world.toActionMatrixInverse() * J.
 * jac [n][6][state_dim], link_T [n][12] (3x4 row-major world placement of the link). */
int orc_link_jacobian_batch(const mpb_scene_desc *d, int link, const double *q, int64_t n, int local, double *jac, double *link_T) {
  if (link < 0 || link >= d->n_links) return -1;
  const int D = d->state_dim;
  double *qfull = (double *)malloc(sizeof(double) * (size_t)(d->nq_full > 0 ? d->nq_full : 1));
  xf_t *oMi = (xf_t *)malloc(sizeof(xf_t) * (size_t)(d->n_joints + 1));
  int *slot_of_q = (int *)malloc(sizeof(int) * (size_t)(d->nq_full > 0 ? d->nq_full : 1));
  for (int k = 0; k < d->nq_full; k++) slot_of_q[k] = -1;
  for (int k = 0; k < D; k++) slot_of_q[d->state_map[k]] = k;
  for (int64_t i = 0; i < n; i++) {
    memcpy(qfull, d->q_base, sizeof(double) * (size_t)d->nq_full);
    for (int k = 0; k < D; k++) qfull[d->state_map[k]] = q[i * D + k];
    fk_joints(d, qfull, oMi);
    xf_t L, lw; xf_from12(&L, d->link_placement + 12 * link);
    const int pj = d->link_joint[link];
    if (pj == 0) lw = L; else xf_mul(&lw, &oMi[pj], &L);
    double *J = jac + i * 6 * D;
    for (int k = 0; k < 6 * D; k++) J[k] = 0.0;
    for (int j = pj; j != 0; j = d->joint_parent[j - 1]) {
      const int sl = slot_of_q[d->joint_qidx[j - 1]];
      if (sl < 0) continue;
      xf_t X;
      if (local) xf_inv_mul(&X, &lw, &oMi[j]); else X = oMi[j];
      double a[3], col[6] = {0, 0, 0, 0, 0, 0};
      xf_rot(a, &X, d->joint_axis + 3 * (j - 1));
      if (d->joint_type[j - 1] == MPB_JOINT_PRISMATIC) copy3(col, a);
      else { cross3(col, X.p, a); copy3(col + 3, a); }
      for (int r = 0; r < 6; r++) J[r * D + sl] = col[r];
    }
    if (link_T) for (int r = 0; r < 3; r++) { for (int c = 0; c < 3; c++) link_T[12 * i + 4 * r + c] = lw.R[3 * r + c]; link_T[12 * i + 4 * r + 3] = lw.p[r]; }
  }
  free(qfull); free(oMi); free(slot_of_q);
  return 0;
}
