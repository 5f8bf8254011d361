// mpb_kernels.cuh -- sm_100a kernels of the batched state-validity path.
//
//   fk_table_kernel    thread per configuration.  FK streams down the kinematic forest with the running joint
//                      transform in registers; joint / moving-object records come from the parameter space (FkTab).
//                      Writes the pose tables in HBM: OBB rows (quaternion + translation, 32 B per moving object) and
//                      the word-major sphere table.  (Replaces PinocchioModelTpl::computeForwardKinematics +
//                      getLinkPose + FCLModelTpl::updateCollisionObjects + AttachedBodyTpl::updatePose,
//                      src/pinocchio_model.cpp:271-312, src/fcl_model.cpp:138-148, src/attached_body.h:48-54.)
//   sphere_kernel      thread per configuration over the cull table (admitted pairs in runs that share their moving
//   sphere_blk_kernel  object): bounding-sphere test, 32 entries per mask word, survivors pushed into warp-private
//   sphere_grid_kernel chunks of the survivor queue.  Small tables ride in the parameter space with the tile's
//   sphere_pairs_kernel sphere centres staged in shared memory; large ones with world-fixed partners go through a uniform
//                      grid per moving object (a handful of candidates instead of the whole run); small passes use a
//                      thread per (configuration, entry).
//   dense2_kernel      thread per survivor, two converged phases: 15-axis OBB SAT from the two OBB rows + inscribed-ball
//   (dense_kernel)     test; then the relative pose of what is left, the sufficient-hit test with the inner balls, and one
//                      64-byte ITEM RECORD + 2-byte bin key per item (warp-private chunks of 256 slots).
//   scan_bins_kernel   prefix sums of the per-bin counts -> offsets of the binned order.
//   scatter_kernel     ranks the bin keys and writes order[]: the items are sorted by INDEX, the records never move
//                      (bin = (geometry A, geometry B) class).
//   narrow2_kernel     lane per work item, a warp owns a slice of ONE bin (so all lanes walk the same hulls): a converged
//   (narrow_kernel)    prologue over 32 fresh items (fetch through order[], first two supports and their exits), survivors
//                      buffered in shared memory, then the portal loop over full warps of survivors.  Hull pool + support
//                      cell tables staged in shared memory by TMA bulk copies (cp.async.bulk + mbarrier).  MPR (replaces
//                      ::fcl::collide -> ccdMPRIntersect on the pairs of FCLModelTpl::collideFull /
//                      PlanningWorldTpl::selfCollide / collideWithOthers, src/fcl_model.cpp:181-193,
//                      src/planning_world.cpp:276-481).
//   expand_kernel      shape x BVH object (mesh / octree): AABB-tree descent against the shape's box, leaf items for a
//   expand_lane_kernel second narrowphase round; expand2_kernel: mesh x mesh / mesh x octree with exact leaf tests in place.
//   hits_compact_kernel dense pair mask of a pass -> (configuration, pair) list.
//   fk_pose_kernel     FK-only API: link poses as [x,y,z,qw,qx,qy,qz].
#pragma once
#include <float.h>

#include "../../include/mpb.h"
#include "mpb_device.cuh"

struct SceneImg {
  const DevJoint *joints;
  const DevObject *objs;
  const DevPair *pairs;
  const float4 *verts;      // hull vertex pool, followed by the support cell tables (pool_bytes in all)
  const int *dynlist;   // moving objects grouped by owner joint
  const DevBvhNode *nodes;  // BVH node pool (meshes, octrees); DevObject::voff = root, ::vcnt = leaf type
  const float4 *prims;      // leaf primitives, 3 float4 each
  const float4 *cobb;       // [n_objects][MPB_COBB_WORDS] cull records (see dense_kernel)
  const float4 *cullc;      // [n_pairs] cull table: world-fixed partner's sphere centre (xyz), (R1 + R2 + margin)^2 (w)
  const float4 *cullq;      // [n_pairs] world-fixed partners in expanded form: (-2c, |c|^2 - (R1 + R2)^2 - slack), see sphere_kernel
  const int4 *cullp;        // [n_pairs] cull table: {pair index, object a, object b, (slot a + 1) | (slot b + 1) << 16}
  const int2 *cullm;        // [n_pairs] cull table: {pair index, (moving slot 1 + 1) | (moving slot 2 + 1) << 16}
  const int4 *runs;         // cull-table runs: {moving slot or -1, first entry, count, partner kind (0 fixed, 1 moving)}
  // broadphase grid over the world-fixed partners (sphere_grid_kernel): one cell table per run of fixed partners
  const int *grid_start;    // [n_grid_runs][gx * gy * gz + 1] offsets into grid_list (relative to the run's block)
  const unsigned short *grid_list;  // cull-table entries (relative to the run's first entry) per cell
  const int2 *grid_run;     // [n_runs] {offset of the run's cell table in grid_start or -1, offset of its lists in grid_list}
  float grid_o[3], grid_inv_h; // grid origin, 1 / cell size
  int grid_n[3];            // cells per axis
  const int *pair_bin2;     // [n_pairs] bin of the leaf-level items a BVH pair expands into (-1: not a BVH pair)
  int expand_bin;           // first-round bin that collects the shape x BVH pairs (-1: none)
  int expand2_bin;          // first-round bin that collects the BVH x BVH pairs (-1: none)
  int n_bins2;
  const float *linkL;   // [n_links][12] link placement in its joint frame
  const int *linkJoint; // [n_links]
  int n_joints, n_objects, n_pairs, n_links, n_verts, n_bins, state_dim, pair_words, n_dyn, n_jslots;
  int pool_bytes;         // vertex pool + cell tables, staged as one block in shared memory
  int w_sphere, w_joint;  // sphere-table word offsets: [n_dyn][3] sphere centres, then [n_jslots][12] branch joints
};

#define MPB_KEY_HOLE 0xffffu
struct WorkQ {
  float4 *rec[2];        // item records, 64 bytes each, in the order they were produced: [0] first round (dense cull),
                         // [1] leaf items of the second round (BVH expansion).  R (9), t (3): B in A; cfg, pair, aux
                         // (leaf primitive index in the second round, else 0)
  unsigned short *skey;  // [cap] bin of the record in slot s of the round being binned; MPB_KEY_HOLE = unused chunk tail
  unsigned int *order;   // [cap] binned order: order[k] = slot of the k-th item (bins are ranges of k)
  unsigned int *bin_count;   // [n_bins]
  unsigned int *bin_cursor;  // [n_bins]
  unsigned int *bin_start;   // [n_bins+1]
  unsigned int *unit_start;  // [n_bins+1]   units of MPB_UNIT_ITEMS items
  unsigned int *totals;      // [0]=record slots handed out (holes included) [1]=units [2]=overflow flag [3]=unit ticket [4]=items per unit
  unsigned long long *counters;  // [0]=items after cull [1]=support evaluations [2]=capped items
  unsigned int cap;
};

// Items are SORTED BY INDEX, not moved: the producer writes one 64-byte record per item (full sectors) and a 2-byte bin
// key; scatter_kernel ranks the keys and writes order[]; the consumers read order[k] and gather the record (two full
// sectors).  The first layout (structure of arrays, staged, then copied to its binned position by scatter_kernel) moved
// 112 bytes per item through the scatter and made it 11 % of a C3 pass.
__device__ __forceinline__ void item_store(const WorkQ &wq, int round, unsigned slot, const float *R, const float *t, int cfg, int pair, int aux, int bin) {
  float4 *r = wq.rec[round] + 4 * (size_t)slot;
  r[0] = make_float4(R[0], R[1], R[2], R[3]);
  r[1] = make_float4(R[4], R[5], R[6], R[7]);
  r[2] = make_float4(R[8], t[0], t[1], t[2]);
  r[3] = make_float4(__int_as_float(cfg), __int_as_float(pair), __int_as_float(aux), 0.f);
  wq.skey[slot] = (unsigned short)bin;
}
// record of the k-th binned item
__device__ __forceinline__ const float4 *item_rec(const WorkQ &wq, int round, unsigned k) { return wq.rec[round] + 4 * (size_t)wq.order[k]; }
__device__ __forceinline__ void item_ids(const float4 *r, int &cfg, int &pair, int &aux) {
  const float4 d = __ldg(r + 3);
  cfg = __float_as_int(d.x); pair = __float_as_int(d.y); aux = __float_as_int(d.z);
}
__device__ __forceinline__ void item_pose(const float4 *r, float *R, float *t) {
  const float4 a = __ldg(r), b = __ldg(r + 1), c = __ldg(r + 2);
  R[0] = a.x; R[1] = a.y; R[2] = a.z; R[3] = a.w; R[4] = b.x; R[5] = b.y; R[6] = b.z; R[7] = b.w; R[8] = c.x;
  t[0] = c.y; t[1] = c.z; t[2] = c.w;
}

// Where the kernels of one pass report a colliding (configuration, pair): the verdict word and, optionally, the dense
// pair mask of the pass.  (The sparse hit list of mpb_collide_batch_hits is compacted out of the mask after the pass by
// hits_compact_kernel: a list appended to from inside the narrowphase costs one same-address atomic per hit, which
// at ~10 hits per configuration in the C3 scene slowed the whole pass by 40 %.)
struct HitOut {
  unsigned int *verdict;    // [ceil(m / 32)] words of this pass
  unsigned int *pair_bits;  // [m][pair_words] of this pass, or nullptr
  int pair_words;
};
__device__ __forceinline__ void report_hit(const HitOut &o, unsigned cfg, int p) {
  atomicOr(&o.verdict[cfg >> 5], 1u << (cfg & 31));
  if (o.pair_bits) atomicOr(&o.pair_bits[(size_t)cfg * o.pair_words + (p >> 5)], 1u << (p & 31));
}
// this (configuration, pair) -- or, without a pair mask, this configuration -- is already known to collide
__device__ __forceinline__ bool hit_known(const HitOut &o, unsigned cfg, int p) {
  return o.pair_bits ? ((o.pair_bits[(size_t)cfg * o.pair_words + (p >> 5)] >> (p & 31)) & 1u) : ((o.verdict[cfg >> 5] >> (cfg & 31)) & 1u);
}

// Dense pair mask of a pass -> (configuration, pair) entries appended to the call-wide list.  A block compacts tiles of
// 1024 mask words (one 16-byte load per thread): popcounts, warp + block prefix sums, ONE atomic per tile for the
// slots; the threads expand their set bits into a shared-memory stage at their rank within the tile, and the block
// copies the stage out with consecutive 8-byte stores (a tile with more hits than the stage holds writes directly).
// Entries are ordered by configuration within a tile; the reservation order of the tiles is the only nondeterminism.
#define MPB_HC_STAGE 2048
__global__ void __launch_bounds__(256) hits_compact_kernel(const unsigned int *__restrict__ pair_bits, long long n_words, int pair_words,
                                                           unsigned cfg0, uint2 *__restrict__ hits, unsigned int *__restrict__ n_hits,
                                                           unsigned hit_cap) {
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ unsigned s_warp[8], s_base, s_tot;
  __shared__ uint2 s_stage[MPB_HC_STAGE];
  const long long n_quads = (n_words + 3) / 4, n_tiles = (n_quads + 255) / 256;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const long long qd = tile * 256 + threadIdx.x;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (qd * 4 + 3 < n_words) v = __ldg(reinterpret_cast<const uint4 *>(pair_bits) + qd);
    else if (qd * 4 < n_words) {
      v.x = pair_bits[qd * 4];
      if (qd * 4 + 1 < n_words) v.y = pair_bits[qd * 4 + 1];
      if (qd * 4 + 2 < n_words) v.z = pair_bits[qd * 4 + 2];
    }
    const unsigned c = __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w);
    unsigned incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned t = __shfl_up_sync(FULL, incl, o);
      if ((int)lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned tot = 0;
#pragma unroll
      for (int k = 0; k < 8; k++) { const unsigned t = s_warp[k]; s_warp[k] = tot; tot += t; }
      s_tot = tot;
      s_base = tot ? atomicAdd(n_hits, tot) : 0u;
    }
    __syncthreads();
    const unsigned tot = s_tot, base = s_base;
    const bool staged = tot <= MPB_HC_STAGE;
    unsigned pos = s_warp[warp] + incl - c;  // rank within the tile
    if (c) {
      // (configuration, first pair) of the thread's first word: one division per thread, then counted forward
      const long long w0 = qd * 4;
      unsigned cfg = cfg0 + (unsigned)(w0 / pair_words), pw = (unsigned)(w0 % pair_words);
      const unsigned words[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; j++) {
        unsigned bits = words[j];
        const unsigned p0 = pw * 32u;
        while (bits) {
          const int k = __ffs(bits) - 1;
          bits &= bits - 1;
          const uint2 ent = make_uint2(cfg, p0 + (unsigned)k);
          if (staged) s_stage[pos] = ent;
          else if (base + pos < hit_cap) hits[base + pos] = ent;
          pos++;
        }
        if (++pw == (unsigned)pair_words) { pw = 0; cfg++; }
      }
    }
    __syncthreads();  // the stage is complete; s_warp / s_base / s_tot may be rewritten after this point
    if (staged)
      for (unsigned k = threadIdx.x; k < tot; k += blockDim.x)
        if (base + k < hit_cap) hits[base + k] = s_stage[k];
    // (the next tile writes the stage only after its own two barriers, which every thread reaches after this copy)
  }
}

__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
__device__ __forceinline__ V3 rot(const float *R, V3 v) {
  return v3(fmaf(R[0], v.x, fmaf(R[1], v.y, R[2] * v.z)), fmaf(R[3], v.x, fmaf(R[4], v.y, R[5] * v.z)),
            fmaf(R[6], v.x, fmaf(R[7], v.y, R[8] * v.z)));
}
__device__ __forceinline__ V3 rot_t(const float *R, V3 v) {
  return v3(fmaf(R[0], v.x, fmaf(R[3], v.y, R[6] * v.z)), fmaf(R[1], v.x, fmaf(R[4], v.y, R[7] * v.z)),
            fmaf(R[2], v.x, fmaf(R[5], v.y, R[8] * v.z)));
}

#define MPB_EPS FLT_EPSILON
#define MPB_MPR_TOL 1e-6f
#define MPB_MPR_MAX_ITERS 500  // libccd: GJKSolver_libccd::max_collision_iterations [upstream]
#define MPB_UNIT_ITEMS 128   // most items a warp owns at a time in the narrowphase (scan_bins_kernel picks 2..128)
#define MPB_COOP_MAX 10      // busy lanes at or below which a warp evaluates hull supports cooperatively
#ifndef MPB_STAGE_CHUNK
#define MPB_STAGE_CHUNK 256  // staging slots a warp reserves per atomic (>= 32)
#endif
#define MPB_SMEM_BINS 1024    // bins a block can count in shared memory (more: global atomics)

// ---------------------------------------------------------------------------------------------- FK
__device__ __forceinline__ Xf xf_identity() {
  Xf m;
  m.r[0] = 1.f; m.r[1] = 0.f; m.r[2] = 0.f; m.r[3] = 0.f; m.r[4] = 1.f; m.r[5] = 0.f; m.r[6] = 0.f; m.r[7] = 0.f; m.r[8] = 1.f;
  m.p[0] = m.p[1] = m.p[2] = 0.f;
  return m;
}
__device__ __forceinline__ Xf joint_motion(const DevJoint &J, float q) {
  Xf m;
  if (J.type == MPB_JOINT_PRISMATIC) {
    m = xf_identity();
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

__device__ __forceinline__ Xf xf_rows(float4 a, float4 b, float4 c) {
  Xf o;
  o.r[0] = a.x; o.r[1] = a.y; o.r[2] = a.z; o.p[0] = a.w;
  o.r[3] = b.x; o.r[4] = b.y; o.r[5] = b.z; o.p[1] = b.w;
  o.r[6] = c.x; o.r[7] = c.y; o.r[8] = c.z; o.p[2] = c.w;
  return o;
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

// oMi for all joints of one configuration into a per-thread table; T[0] = identity (FK-only API).
__device__ __forceinline__ void fk_joints(const SceneImg &img, const float *qrow, Xf *T) {
  T[0] = xf_identity();
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

// world transform of joint J given its parent's: cur = parent * P * motion(q).  Axis-aligned revolute joints
// (Pinocchio JointModelRX/RY/RZ) only mix two columns of parent * P.
__device__ __forceinline__ Xf joint_step(const DevJoint &J, const Xf *parent, float q) {
  Xf a = load_xf12(J.P);
  if (parent) a = xf_mul(*parent, a);
  if (J.kind == MPB_JK_PRISMATIC) {
#pragma unroll
    for (int r = 0; r < 3; r++)
      a.p[r] = fmaf(a.r[3 * r], J.ax[0] * q, fmaf(a.r[3 * r + 1], J.ax[1] * q, fmaf(a.r[3 * r + 2], J.ax[2] * q, a.p[r])));
    return a;
  }
  if (J.kind == MPB_JK_GENERIC) {
    const Xf m = joint_motion(J, q);
    Xf o;
#pragma unroll
    for (int i = 0; i < 3; i++) {
#pragma unroll
      for (int j = 0; j < 3; j++)
        o.r[3 * i + j] = fmaf(a.r[3 * i], m.r[j], fmaf(a.r[3 * i + 1], m.r[3 + j], a.r[3 * i + 2] * m.r[6 + j]));
      o.p[i] = a.p[i];
    }
    return o;
  }
  // MUFU sine / cosine: absolute error < 1e-6 for |q| <= 2 pi (joint limits), i.e. < 1e-6 m per joint at 1 m reach; the
  // pose table only feeds the culls and the narrowphase (contact band 1e-4 m).  The FK API (fk_pose_kernel) keeps sincosf.
  float s, c;
  __sincosf(q, &s, &c);
  // rotation about +-axis k mixes the two other columns (u, v) of parent * P; -axis: opposite angle
#define MPB_MIX(u, v)                                    \
  _Pragma("unroll") for (int r = 0; r < 3; r++) {        \
    const float cu = a.r[3 * r + u], cv = a.r[3 * r + v]; \
    a.r[3 * r + u] = fmaf(c, cu, s * cv);                \
    a.r[3 * r + v] = fmaf(c, cv, -s * cu);               \
  }
  if (J.kind == MPB_JK_RX) { s *= J.ax[0]; MPB_MIX(1, 2) }
  else if (J.kind == MPB_JK_RY) { s *= J.ax[1]; MPB_MIX(2, 0) }
  else { s *= J.ax[2]; MPB_MIX(0, 1) }
#undef MPB_MIX
  return a;
}

// ---------------------------------------------------------------------------------------------- FK + cull, three kernels
// The pose tables live in HBM.  OBB table: [moving object][configuration][3 x float4], the world pose of each moving
// object's OBB frame (fk_table_kernel transposes a warp's 32 frames through shared memory and stores 1536 contiguous
// bytes; dense_kernel reads three 16-byte words per object).  Sphere table, word-major [word][configuration]: 3 words per moving object's bounding-sphere
// centre (coalesced for the thread-per-configuration sphere phase) and 12 words per branch-point joint (a parent that a
// later, non-adjacent child needs).  ~700 B per Panda configuration is cheap at HBM3e rates (a million configurations per
// 0.2 ms, read + write) and it frees the kernels from a shared-memory pose table that capped them at ten warps per
// SM: all three kernels below are register-only and run at full occupancy.
//   fk_table_kernel   thread per configuration: FK down the forest; OBB frames + sphere centres -> pose table
//   sphere_kernel     thread per configuration over the CULL TABLE: the admitted pairs re-ordered into runs that share
//                     their moving object (its centre stays in registers) and whose partner is either world-fixed
//                     (centre and squared radius sum in one float4, a uniform 16-byte load) or another moving object;
//                     survivors (configuration, entry) go to the survivor queue through warp-private chunks
//   dense_kernel      thread per survivor: 15-axis OBB SAT straight from the two OBB frames (pose table for moving
//                     objects, cull record for fixed ones), inscribed-ball sufficient test, relative pose of what is
//                     left -> staging queue
// MODE 0 (collide): as above.  MODE 1 (distance): a pair is a candidate for the minimum iff its bounding-sphere lower
// bound |c| - R does not exceed the smallest upper bound |c| + R over all admitted pairs of the configuration;
// candidates skip the SAT (overlap is not the question) and go to the GJK distance kernel.
#define MPB_SQ_CHUNK 128  // survivor-queue slots a warp reserves per atomic
#define POSE(w, i) pose[(size_t)(w) * ld + (i)]

#define MPB_COBB_WORDS 11  // float4 per object in the cull records
// 12 consecutive words of the pose table, starting at word w0, of configuration i
__device__ __forceinline__ Xf pose_joint(const float *__restrict__ pose, size_t ld, int w0, long long i) {
  Xf w;
#pragma unroll
  for (int k = 0; k < 9; k++) w.r[k] = POSE(w0 + k, i);
#pragma unroll
  for (int k = 0; k < 3; k++) w.p[k] = POSE(w0 + 9 + k, i);
  return w;
}

// OBB table rows are (unit quaternion xyzw, translation): 32 B per object and configuration instead of a 48 B matrix.
__device__ __forceinline__ float4 rot_to_quat(const float *R) {  // Shepperd's branches, normalised
  float x, y, z, w;
  const float tr = R[0] + R[4] + R[8];
  if (tr > 0.f) { w = tr + 1.f; x = R[7] - R[5]; y = R[2] - R[6]; z = R[3] - R[1]; }
  else if (R[0] >= R[4] && R[0] >= R[8]) { x = 1.f + R[0] - R[4] - R[8]; y = R[1] + R[3]; z = R[2] + R[6]; w = R[7] - R[5]; }
  else if (R[4] >= R[8]) { y = 1.f + R[4] - R[0] - R[8]; x = R[1] + R[3]; z = R[5] + R[7]; w = R[2] - R[6]; }
  else { z = 1.f + R[8] - R[0] - R[4]; x = R[2] + R[6]; y = R[5] + R[7]; w = R[3] - R[1]; }
  const float s = rsqrtf(fmaf(x, x, fmaf(y, y, fmaf(z, z, w * w))));
  return make_float4(x * s, y * s, z * s, w * s);
}
__device__ __forceinline__ Xf quat_rows(const float4 q, const float4 t) {
  Xf o;
  const float x2 = q.x + q.x, y2 = q.y + q.y, z2 = q.z + q.z;
  const float xx = q.x * x2, yy = q.y * y2, zz = q.z * z2, xy = q.x * y2, xz = q.x * z2, yz = q.y * z2, wx = q.w * x2, wy = q.w * y2, wz = q.w * z2;
  o.r[0] = 1.f - (yy + zz); o.r[1] = xy - wz;         o.r[2] = xz + wy;
  o.r[3] = xy + wz;         o.r[4] = 1.f - (xx + zz); o.r[5] = yz - wx;
  o.r[6] = xz - wy;         o.r[7] = yz + wx;         o.r[8] = 1.f - (xx + yy);
  o.p[0] = t.x; o.p[1] = t.y; o.p[2] = t.z;
  return o;
}

// The joint and moving-object records of the kernel travel as a kernel parameter (constant bank, uniform loads):
// fetched from global memory they cost ~220 per-lane LDG per warp and made the L1/LSU data path the busiest unit of
// the kernel (75 % of peak in the full ncu capture of round 1) although every lane reads the same address.
struct FkObjC {
  float4 rec[3];  // OBB frame in the owner joint frame (cull record words 0..2)
  float bcj[3];   // bounding-sphere centre in the owner joint frame
  int dyn;        // moving-object slot
};
struct FkTab {
  DevJoint j[MPB_MAX_JOINTS];  // dyn_first indexes o[] (objects grouped by owner joint)
  FkObjC o[MPB_MAX_DYN];
};
// (no minimum-blocks hint: with one, ptxas spends 72 registers instead of 56 and the kernel loses a fifth of its occupancy)
__global__ void __launch_bounds__(128) fk_table_kernel(const __grid_constant__ SceneImg img, const __grid_constant__ FkTab tab,
                                                       const float *__restrict__ qconst, const float *__restrict__ q, long long n, float *__restrict__ pose, size_t ld,
                                                       float4 *__restrict__ obb) {
  // per warp: one object's 32 OBB rows, transposed for full-sector stores.  Quaternions at float4 [0, 32), translations
  // at [36, 68): the 4-word shift puts the two halves of a quarter-warp's interleaved read on disjoint banks
  __shared__ __align__(16) float s_tr[4][68 * 4];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long tile0 = i - lane;
  if (tile0 >= n) return;                   // whole warps stay: the tables are padded to a multiple of 32 configurations
  const long long ic = i < n ? i : n - 1;   // lanes past the end recompute the last configuration into the padding
  const int cbase = img.w_sphere, jbase = img.w_joint;
  float *tr = s_tr[warp];
  Xf cur = xf_identity();
  const float *qrow = q + ic * img.state_dim;
  for (int j = 0; j < img.n_joints; j++) {
    const DevJoint &J = tab.j[j];
    const float qj = J.slot >= 0 ? qrow[J.slot] : __ldg(qconst + j);  // q_base values live in a device table: changing them
                                                                      // leaves the by-value records (and captured graphs) alone
    if (J.parent == 0) cur = joint_step(J, nullptr, qj);
    else if (J.pjslot >= 0) { const Xf par = pose_joint(pose, ld, jbase + J.pjslot * 12, i); cur = joint_step(J, &par, qj); }
    else cur = joint_step(J, &cur, qj);
    if (J.jslot >= 0) {
#pragma unroll
      for (int k = 0; k < 9; k++) POSE(jbase + J.jslot * 12 + k, i) = cur.r[k];
#pragma unroll
      for (int k = 0; k < 3; k++) POSE(jbase + J.jslot * 12 + 9 + k, i) = cur.p[k];
    }
    for (int e = 0; e < J.dyn_count; e++) {
      const FkObjC &O = tab.o[J.dyn_first + e];
      float c[3];
      xf_point(cur, O.bcj, c);  // bounding-sphere centre in the joint frame -> world
#pragma unroll
      for (int k = 0; k < 3; k++) POSE(cbase + O.dyn * 3 + k, i) = c[k];
      const Xf E = xf_mul(cur, xf_rows(O.rec[0], O.rec[1], O.rec[2]));  // OBB frame in the joint frame -> world
      // OBB table [object][configuration][quaternion, translation]: the warp's 32 frames are 1024 contiguous bytes
      float4 *tr4 = reinterpret_cast<float4 *>(tr);
      tr4[lane] = rot_to_quat(E.r);
      tr4[36 + lane] = make_float4(E.p[0], E.p[1], E.p[2], 0.f);
      __syncwarp();
      float4 *dst = obb + ((size_t)O.dyn * ld + tile0) * 2;
#pragma unroll
      for (int r = 0; r < 2; r++) {
        const unsigned o = lane + 32 * r;  // output word: configuration o / 2, half o % 2
        dst[o] = tr4[(o & 1u) * 36 + (o >> 1)];
      }
      __syncwarp();
    }
  }
}

// Cull table in parameter space (constant bank): the entries are the same for every lane, and fetched with per-lane
// LDG.128 each one writes 512 B back through the LSU.  NQ entries / NR runs; larger scenes use the global tables.
template <int NQ, int NR>
struct CullTabC {
  float4 q[NQ];   // fixed partner: (-2c, |c|^2 - rr2 - slack); moving partner: (slot as int bits, 0, 0, rr2)
  int4 runs[NR];
};
template <>
struct CullTabC<0, 0> {};
template <int MODE, int NQ, int NR>
__global__ void __launch_bounds__(128) sphere_kernel(const __grid_constant__ SceneImg img, const __grid_constant__ CullTabC<NQ, NR> ct,
                                                     const float *__restrict__ pose, size_t ld, long long n, long long cfg_base,
                                                     uint2 *__restrict__ sq, unsigned sq_cap, unsigned *__restrict__ totals) {
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31;
  const int cbase = img.w_sphere;
  extern __shared__ float s_ctr[];  // NQ > 0: [warp][n_dyn * 3][32] sphere centres of the warp's tile
  float *ctr = s_ctr + (size_t)(threadIdx.x >> 5) * (img.n_dyn * 3 * 32) + lane;
  unsigned wbase = 0, wused = 0, wcap = 0;  // warp-uniform survivor-queue chunk, kept across the warp's tiles
  const long long n_tiles = (n + 31) / 32;
  const long long warp_stride = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long tile = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); tile < n_tiles; tile += warp_stride) {
  const long long i = tile * 32 + lane;  // whole warps run; lanes past n push nothing
  const bool valid = i < n;
  const long long ic = valid ? i : n - 1;
  if constexpr (NQ > 0) {
    // parameter-space tier (few pairs, mostly moving x moving): the tile's sphere centres are staged once in the
    // warp's [word][lane] block of shared memory; a partner's centre is then three LDS at a 32-bit address instead of
    // three global loads with 64-bit index arithmetic (41 -> 16 instructions per moving x moving pair)
    const float *pcol = pose + (size_t)cbase * ld + ic;
#pragma unroll 4
    for (int w = 0; w < img.n_dyn * 3; w++) ctr[w * 32] = pcol[(size_t)w * ld];
  }
  float ub_min = FLT_MAX;
  if (MODE == 1) {
    for (int k = 0; k < img.n_pairs; k++) {
      const float4 c = __ldg(img.cullc + k);
      const int mm = __ldg(&img.cullm[k].y);
      const int m1 = mm & 0xffff, m2 = mm >> 16;  // biased by one; 0 = none
      float ax = 0.f, ay = 0.f, az = 0.f, bx = c.x, by = c.y, bz = c.z;
      if (m1) { ax = POSE(cbase + (m1 - 1) * 3, ic); ay = POSE(cbase + (m1 - 1) * 3 + 1, ic); az = POSE(cbase + (m1 - 1) * 3 + 2, ic); }
      if (m2) { bx = POSE(cbase + (m2 - 1) * 3, ic); by = POSE(cbase + (m2 - 1) * 3 + 1, ic); bz = POSE(cbase + (m2 - 1) * 3 + 2, ic); }
      const float dx = ax - bx, dy = ay - by, dz = az - bz;
      ub_min = fminf(ub_min, sqrtf(fmaf(dx, dx, fmaf(dy, dy, dz * dz))) + sqrtf(c.w));
    }
    ub_min *= 1.000001f;
  }
  int r = 0;  // run cursor
  for (int gbase = 0; gbase < img.n_pairs; gbase += 32) {
    const int gend = min(gbase + 32, img.n_pairs);
    unsigned mask = 0;
    int k = gbase;
    while (k < gend) {
      int4 run;  // {moving slot or -1, first, count, partner kind}
      if constexpr (NQ > 0) run = ct.runs[r];
      else run = __ldg(img.runs + r);
      const int rlast = run.y + run.z;
      const int rend = min(rlast, gend);
      float ax = 0.f, ay = 0.f, az = 0.f;
      if constexpr (NQ > 0) {
        if (run.x >= 0) { ax = ctr[run.x * 96]; ay = ctr[run.x * 96 + 32]; az = ctr[run.x * 96 + 64]; }
      } else {
        if (run.x >= 0) { ax = POSE(cbase + run.x * 3, ic); ay = POSE(cbase + run.x * 3 + 1, ic); az = POSE(cbase + run.x * 3 + 2, ic); }
      }
      if constexpr (NQ > 0) {  // MODE 0 only
        const float na2 = -fmaf(ax, ax, fmaf(ay, ay, az * az));
        if (run.w == 0) {
#pragma unroll 8
          for (; k < rend; k++) {
            const float4 c = ct.q[k];
            mask |= (unsigned)(fmaf(ax, c.x, fmaf(ay, c.y, fmaf(az, c.z, c.w))) <= na2) << (k - gbase);
          }
        } else {
#pragma unroll 4
          for (; k < rend; k++) {
            const float4 c = ct.q[k];
            const int m2 = __float_as_int(c.x);
            const float dx = ax - ctr[m2 * 96], dy = ay - ctr[m2 * 96 + 32], dz = az - ctr[m2 * 96 + 64];
            mask |= (unsigned)(fmaf(dx, dx, fmaf(dy, dy, dz * dz)) <= c.w) << (k - gbase);
          }
        }
      } else if (run.w == 0 && MODE == 0) {
        // world-fixed partners, one uniform 16-byte load each.  |a - c|^2 <= rr2 in expanded form,
        // |a|^2 + (-2c).a + (|c|^2 - rr2) <= 0: three FFMA and a compare per pair; the host subtracts a slack that
        // covers the cancellation error, so the test stays a conservative cull
        const float na2 = -fmaf(ax, ax, fmaf(ay, ay, az * az));
#pragma unroll 8
        for (; k < rend; k++) {
          const float4 c = __ldg(img.cullq + k);
          mask |= (unsigned)(fmaf(ax, c.x, fmaf(ay, c.y, fmaf(az, c.z, c.w))) <= na2) << (k - gbase);
        }
      } else if (run.w == 0) {  // distance mode: centre + (R1 + R2)^2
#pragma unroll 8
        for (; k < rend; k++) {
          const float4 c = __ldg(img.cullc + k);
          const float dx = ax - c.x, dy = ay - c.y, dz = az - c.z;
          const float d2c = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
          mask |= (unsigned)(sqrtf(d2c) - sqrtf(c.w) <= ub_min) << (k - gbase);
        }
      } else {  // moving partners
        for (; k < rend; k++) {
          const float rr2 = __ldg(&img.cullc[k].w);
          const int m2 = (__ldg(&img.cullm[k].y) >> 16) - 1;
          const float dx = ax - POSE(cbase + m2 * 3, ic), dy = ay - POSE(cbase + m2 * 3 + 1, ic), dz = az - POSE(cbase + m2 * 3 + 2, ic);
          const float d2c = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
          const bool s = MODE == 0 ? (d2c <= rr2) : (sqrtf(d2c) - sqrtf(rr2) <= ub_min);
          mask |= (unsigned)s << (k - gbase);
        }
      }
      if (rend == rlast) r++;
    }
    if (!valid) mask = 0;
    if (!__any_sync(FULL, mask != 0)) continue;
    // push (configuration, entry) survivors: warp prefix sum into the warp's chunk of the survivor queue
    const unsigned c = __popc(mask);
    unsigned incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned t = __shfl_up_sync(FULL, incl, o);
      if ((int)lane >= o) incl += t;
    }
    const unsigned wtot = __shfl_sync(FULL, incl, 31);
    if (wused + wtot > wcap) {
      for (unsigned k2 = wused + lane; k2 < wcap; k2 += 32)
        if (wbase + k2 < sq_cap) sq[wbase + k2] = make_uint2(0xffffffffu, 0u);  // unused tail of the old chunk
      const unsigned want = max((unsigned)MPB_SQ_CHUNK, wtot);
      unsigned nb = 0;
      if (lane == 0) nb = atomicAdd(&totals[6], want);
      wbase = __shfl_sync(FULL, nb, 0);
      wused = 0;
      wcap = want;
    }
    unsigned pos = wbase + wused + incl - c;
    const unsigned cfg = (unsigned)(cfg_base + i);
    while (mask) {
      const int kk = __ffs(mask) - 1;
      mask &= mask - 1;
      if (pos < sq_cap) sq[pos] = make_uint2(cfg, (unsigned)(gbase + kk));
      pos++;
    }
    wused += wtot;
  }
  }  // tiles
  for (unsigned k2 = wused + lane; k2 < wcap; k2 += 32)
    if (wbase + k2 < sq_cap) sq[wbase + k2] = make_uint2(0xffffffffu, 0u);
}

// FK + sphere phase in one kernel, for scenes of the parameter-space tier (a robot against itself: C2's 19 pairs, C5's
// 150): thread per configuration, the joint chain in registers, the sphere centres of the thread's configuration in the
// warp's [word][lane] block of shared memory -- and NO sphere table, and OBB rows ON DEMAND: fk_table_kernel writes 352 B
// of OBB rows per Panda configuration of which the ~1.6 survivors per configuration read < 10 % (80 % of the DRAM traffic of
// a C2 pass).  Here a thread collects, while it tests the pairs, the set of moving objects its survivors name, then walks
// the joint chain a second time (arithmetic is cheap next to the HBM writes it replaces) and writes the rows of those
// objects only; the dense cull reads them where it always did.  A first version re-derived the frames per survivor inside
// the dense cull (two divergent chains per survivor, 110 registers): C2 0.375 -> 0.400 ms, C5 3.9 -> 5.2 ms, dropped.
// MEASURED (profiles/r02_ncu_c2_fk_sphere.txt): DRAM traffic of a C2 pass 790 -> 290 MB, but 186 us against 109 + 65 us for
// fk_table_kernel + sphere_kernel -- FK is bound by issue slots (fk_table_kernel: 74 % issue-active), not by its writes, and
// the second walk costs more instructions than the stores it replaces.  Kept behind MPB_LAZY_FK=1, off by default.
// Branch-point joint frames (pjslot) that a later child reads stay in shared memory (`jneed`: bit per joint-table slot
// that some joint names as its parent).  Same arithmetic as fk_table_kernel + sphere_kernel<0, 256, 64>: identical
// survivors and rows.
__global__ void __launch_bounds__(128) fk_sphere_kernel(const __grid_constant__ SceneImg img, const __grid_constant__ FkTab tab,
                                                        const __grid_constant__ CullTabC<256, 64> ct, const float *__restrict__ qconst,
                                                        const float *__restrict__ q, long long n, long long cfg_base, unsigned jneed,
                                                        float4 *__restrict__ obb, size_t ld, uint2 *__restrict__ sq, unsigned sq_cap,
                                                        unsigned *__restrict__ totals) {
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31;
  extern __shared__ float s_ctr[];  // [warp][n_dyn * 3 + 12 * popc(jneed)][32]
  const int n_need = __popc(jneed);
  float *ctr = s_ctr + (size_t)(threadIdx.x >> 5) * ((img.n_dyn * 3 + n_need * 12) * 32) + lane;
  float *jt = ctr + img.n_dyn * 3 * 32;
  unsigned wbase = 0, wused = 0, wcap = 0;  // warp-uniform survivor-queue chunk, kept across the warp's tiles
  const long long n_tiles = (n + 31) / 32;
  const long long warp_stride = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long tile = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); tile < n_tiles; tile += warp_stride) {
    const long long i = tile * 32 + lane;  // whole warps run; lanes past n push nothing
    const bool valid = i < n;
    const long long ic = valid ? i : n - 1;
    {  // forward kinematics down the forest (fk_table_kernel's loop; the frames stay on chip)
      Xf cur = xf_identity();
      const float *qrow = q + ic * img.state_dim;
      for (int j = 0; j < img.n_joints; j++) {
        const DevJoint &J = tab.j[j];
        const float qj = J.slot >= 0 ? qrow[J.slot] : __ldg(qconst + j);
        if (J.parent == 0) cur = joint_step(J, nullptr, qj);
        else if (J.pjslot >= 0) {
          const float *src = jt + __popc(jneed & ((1u << J.pjslot) - 1u)) * (12 * 32);
          Xf par;
#pragma unroll
          for (int k = 0; k < 9; k++) par.r[k] = src[k * 32];
#pragma unroll
          for (int k = 0; k < 3; k++) par.p[k] = src[(9 + k) * 32];
          cur = joint_step(J, &par, qj);
        } else cur = joint_step(J, &cur, qj);
        if (J.jslot >= 0 && ((jneed >> J.jslot) & 1u)) {
          float *dst = jt + __popc(jneed & ((1u << J.jslot) - 1u)) * (12 * 32);
#pragma unroll
          for (int k = 0; k < 9; k++) dst[k * 32] = cur.r[k];
#pragma unroll
          for (int k = 0; k < 3; k++) dst[(9 + k) * 32] = cur.p[k];
        }
        for (int e = 0; e < J.dyn_count; e++) {
          const FkObjC &O = tab.o[J.dyn_first + e];
          float c[3];
          xf_point(cur, O.bcj, c);  // bounding-sphere centre in the joint frame -> world
#pragma unroll
          for (int k = 0; k < 3; k++) ctr[(O.dyn * 3 + k) * 32] = c[k];
        }
      }
    }
    int r = 0;          // run cursor
    unsigned need = 0;  // moving objects (slots) named by this configuration's survivors
    for (int gbase = 0; gbase < img.n_pairs; gbase += 32) {
      const int gend = min(gbase + 32, img.n_pairs);
      unsigned mask = 0;
      int k = gbase;
      while (k < gend) {
        const int4 run = ct.runs[r];  // {moving slot or -1, first, count, partner kind}
        const int rlast = run.y + run.z;
        const int rend = min(rlast, gend);
        float ax = 0.f, ay = 0.f, az = 0.f;
        const unsigned before = mask;
        if (run.x >= 0) { ax = ctr[run.x * 96]; ay = ctr[run.x * 96 + 32]; az = ctr[run.x * 96 + 64]; }
        if (run.w == 0) {
          const float na2 = -fmaf(ax, ax, fmaf(ay, ay, az * az));
#pragma unroll 8
          for (; k < rend; k++) {
            const float4 c = ct.q[k];
            mask |= (unsigned)(fmaf(ax, c.x, fmaf(ay, c.y, fmaf(az, c.z, c.w))) <= na2) << (k - gbase);
          }
        } else {
#pragma unroll 4
          for (; k < rend; k++) {
            const float4 c = ct.q[k];
            const int m2 = __float_as_int(c.x);
            const float dx = ax - ctr[m2 * 96], dy = ay - ctr[m2 * 96 + 32], dz = az - ctr[m2 * 96 + 64];
            const unsigned hit = (unsigned)(fmaf(dx, dx, fmaf(dy, dy, dz * dz)) <= c.w);
            mask |= hit << (k - gbase);
            need |= hit << m2;
          }
        }
        if (mask != before && run.x >= 0) need |= 1u << run.x;  // the segment's bits are new ones: any hit changes the mask
        if (rend == rlast) r++;
      }
      if (!valid) mask = 0;
      if (!__any_sync(FULL, mask != 0)) continue;
      // push (configuration, entry) survivors: warp prefix sum into the warp's chunk of the survivor queue
      const unsigned c = __popc(mask);
      unsigned incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned t = __shfl_up_sync(FULL, incl, o);
        if ((int)lane >= o) incl += t;
      }
      const unsigned wtot = __shfl_sync(FULL, incl, 31);
      if (wused + wtot > wcap) {
        for (unsigned k2 = wused + lane; k2 < wcap; k2 += 32)
          if (wbase + k2 < sq_cap) sq[wbase + k2] = make_uint2(0xffffffffu, 0u);  // unused tail of the old chunk
        const unsigned want = max((unsigned)MPB_SQ_CHUNK, wtot);
        unsigned nb = 0;
        if (lane == 0) nb = atomicAdd(&totals[6], want);
        wbase = __shfl_sync(FULL, nb, 0);
        wused = 0;
        wcap = want;
      }
      unsigned pos = wbase + wused + incl - c;
      const unsigned cfg = (unsigned)(cfg_base + i);
      while (mask) {
        const int kk = __ffs(mask) - 1;
        mask &= mask - 1;
        if (pos < sq_cap) sq[pos] = make_uint2(cfg, (unsigned)(gbase + kk));
        pos++;
      }
      wused += wtot;
    }
    // ---- OBB rows of the objects the survivors name: the chain once more, rows written where dense2_kernel reads them
    if (!valid) need = 0;
    unsigned wneed = need;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wneed |= __shfl_xor_sync(FULL, wneed, o);
    if (wneed) {
      Xf cur = xf_identity();
      const float *qrow = q + ic * img.state_dim;
      for (int j = 0; j < img.n_joints; j++) {
        const DevJoint &J = tab.j[j];
        const float qj = J.slot >= 0 ? qrow[J.slot] : __ldg(qconst + j);
        if (J.parent == 0) cur = joint_step(J, nullptr, qj);
        else if (J.pjslot >= 0) {  // stored by the first walk
          const float *src = jt + __popc(jneed & ((1u << J.pjslot) - 1u)) * (12 * 32);
          Xf par;
#pragma unroll
          for (int k = 0; k < 9; k++) par.r[k] = src[k * 32];
#pragma unroll
          for (int k = 0; k < 3; k++) par.p[k] = src[(9 + k) * 32];
          cur = joint_step(J, &par, qj);
        } else cur = joint_step(J, &cur, qj);
        for (int e = 0; e < J.dyn_count; e++) {
          const FkObjC &O = tab.o[J.dyn_first + e];
          if (!((wneed >> O.dyn) & 1u)) continue;  // warp-uniform
          const Xf E = xf_mul(cur, xf_rows(O.rec[0], O.rec[1], O.rec[2]));  // OBB frame in the joint frame -> world
          if ((need >> O.dyn) & 1u) {
            float4 *dst = obb + ((size_t)O.dyn * ld + (size_t)(cfg_base + i)) * 2;
            dst[0] = rot_to_quat(E.r);
            dst[1] = make_float4(E.p[0], E.p[1], E.p[2], 0.f);
          }
        }
      }
    }
  }  // tiles
  for (unsigned k2 = wused + lane; k2 < wcap; k2 += 32)
    if (wbase + k2 < sq_cap) sq[wbase + k2] = make_uint2(0xffffffffu, 0u);
}

// sphere phase through a broadphase GRID for large cull tables whose partners are world-fixed (C3: 12 moving objects x 130
// obstacles).  sphere_blk_kernel tests every admitted pair of every configuration, 1587 tests of which 2.6 % pass; here
// the host rasterises, per run (= moving object), every partner's sphere inflated by the object's radius into a uniform
// grid, and a configuration tests only the partners listed in the cell that holds the moving object's centre (mean 7,
// warp maximum ~11 per object instead of 130).  The survivors are exactly those of the full loop (the cell lists are
// conservative), in a different order.  Runs of moving partners (self pairs) keep the direct loop.
#define POSE(w, i) pose[(size_t)(w) * ld + (i)]
#ifndef MPB_SQ_CHUNK_GRID
#define MPB_SQ_CHUNK_GRID 1024  // survivor-queue slots a warp of sphere_grid_kernel reserves per atomic: a trip pushes up to 128 entries,
                                // and what is left of a chunk when the next push does not fit is padding dense_kernel has to skip
#endif
#ifndef MPB_GRID_MINB
#define MPB_GRID_MINB 10  // 48 registers (16 bytes of spills): ten blocks per SM; cull leg of a C3 step 7.58 -> 7.33 ms against 56 registers
#endif
__global__ void __launch_bounds__(128, MPB_GRID_MINB) sphere_grid_kernel(const __grid_constant__ SceneImg img, const float *__restrict__ pose, size_t ld, long long n,
                                                          long long cfg_base, uint2 *__restrict__ sq, unsigned sq_cap, unsigned *__restrict__ totals) {
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31;
  const int cbase = img.w_sphere;
  unsigned wbase = 0, wused = 0, wcap = 0;  // warp-uniform survivor-queue chunk, kept across the warp's tiles
  const long long n_tiles = (n + 31) / 32;
  const long long warp_stride = (long long)gridDim.x * (blockDim.x >> 5);
  const int cells = img.grid_n[0] * img.grid_n[1] * img.grid_n[2];
  for (long long tile = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); tile < n_tiles; tile += warp_stride) {
    const long long i = tile * 32 + lane;
    const bool valid = i < n;
    const long long ic = valid ? i : n - 1;
    const unsigned cfg = (unsigned)(cfg_base + i);
    for (int r = 0;; r++) {
      const int4 run = __ldg(img.runs + r);  // {moving slot or -1, first, count, partner kind}
      if (run.z == 0) break;                  // sentinel
      float ax = 0.f, ay = 0.f, az = 0.f;
      if (run.x >= 0) { ax = POSE(cbase + run.x * 3, ic); ay = POSE(cbase + run.x * 3 + 1, ic); az = POSE(cbase + run.x * 3 + 2, ic); }
      const int2 gr = __ldg(img.grid_run + r);
      int j0 = 0, j1 = 0;  // this lane's slice of candidate slots
      const unsigned short *list = nullptr;
      if (gr.x >= 0) {
        const int cx = (int)floorf((ax - img.grid_o[0]) * img.grid_inv_h), cy = (int)floorf((ay - img.grid_o[1]) * img.grid_inv_h),
                  cz = (int)floorf((az - img.grid_o[2]) * img.grid_inv_h);
        if (valid && cx >= 0 && cy >= 0 && cz >= 0 && cx < img.grid_n[0] && cy < img.grid_n[1] && cz < img.grid_n[2]) {
          const int cell = (cz * img.grid_n[1] + cy) * img.grid_n[0] + cx;
          j0 = __ldg(img.grid_start + gr.x + cell);
          j1 = __ldg(img.grid_start + gr.x + cell + 1);
        }
        list = img.grid_list + gr.y;
      } else if (valid) j1 = run.z;  // no grid for this run (moving partners): every entry is a candidate
      (void)cells;
      // four candidates per trip: the cell lists are padded to multiples of four entries (0xffff = none), so one aligned
      // 8-byte load brings four entry numbers and the four partner records are fetched independently -- the kernel is
      // bound by its chain of dependent loads, not by instructions
      const int len = __reduce_max_sync(FULL, j1 - j0);
      for (int t = 0; t < len; t += 4) {
        unsigned pmask = 0;
        int kk[4] = {0, 0, 0, 0};
        if (j0 + t < j1) {
          if (list) {
            const uint2 pk = __ldg(reinterpret_cast<const uint2 *>(list + j0 + t));
            kk[0] = (int)(pk.x & 0xffffu); kk[1] = (int)(pk.x >> 16); kk[2] = (int)(pk.y & 0xffffu); kk[3] = (int)(pk.y >> 16);
          } else {
#pragma unroll
            for (int u = 0; u < 4; u++) kk[u] = t + u < j1 ? t + u : 0xffff;
          }
          float4 c[4];
#pragma unroll
          for (int u = 0; u < 4; u++) c[u] = __ldg(img.cullc + run.y + (kk[u] == 0xffff ? 0 : kk[u]));
#pragma unroll
          for (int u = 0; u < 4; u++) {
            if (kk[u] == 0xffff) continue;
            float bx = c[u].x, by = c[u].y, bz = c[u].z;
            if (run.w == 1) {  // moving partner
              const int m2 = (__ldg(&img.cullm[run.y + kk[u]].y) >> 16) - 1;
              bx = POSE(cbase + m2 * 3, ic); by = POSE(cbase + m2 * 3 + 1, ic); bz = POSE(cbase + m2 * 3 + 2, ic);
            }
            const float dx = ax - bx, dy = ay - by, dz = az - bz;
            if (fmaf(dx, dx, fmaf(dy, dy, dz * dz)) <= c[u].w) pmask |= 1u << u;
          }
        }
        if (!__any_sync(FULL, pmask != 0)) continue;
        const unsigned cntl = __popc(pmask);
        unsigned incl = cntl;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const unsigned tt = __shfl_up_sync(FULL, incl, o);
          if ((int)lane >= o) incl += tt;
        }
        const unsigned cnt = __shfl_sync(FULL, incl, 31);
        if (wused + cnt > wcap) {
          for (unsigned k2 = wused + lane; k2 < wcap; k2 += 32)
            if (wbase + k2 < sq_cap) sq[wbase + k2] = make_uint2(0xffffffffu, 0u);  // unused tail of the old chunk
          unsigned nb = 0;
          if (lane == 0) nb = atomicAdd(&totals[6], (unsigned)MPB_SQ_CHUNK_GRID);
          wbase = __shfl_sync(FULL, nb, 0);
          wused = 0;
          wcap = MPB_SQ_CHUNK_GRID;
        }
        unsigned pos = wbase + wused + incl - cntl;
#pragma unroll
        for (int u = 0; u < 4; u++)
          if (pmask & (1u << u)) {
            if (pos < sq_cap) sq[pos] = make_uint2(cfg, (unsigned)(run.y + kk[u]));
            pos++;
          }
        wused += cnt;
      }
    }
  }
  for (unsigned k2 = wused + lane; k2 < wcap; k2 += 32)
    if (wbase + k2 < sq_cap) sq[wbase + k2] = make_uint2(0xffffffffu, 0u);
}
#undef POSE

// sphere_kernel<0> for cull tables too large for the parameter space (C3: 1587 entries), register-blocked over CB
// tiles of 32 configurations: an entry fetched with a per-lane LDG.128 writes 512 B back through the LSU (4 cycles per
// SM) although every lane reads the same address, which is more than the ~8 instructions of its test cost to issue;
// applying each fetched entry to CB configurations per lane divides that cost by CB.
__device__ __forceinline__ void push_survivors(unsigned mask, unsigned cfg, int gbase, unsigned lane, uint2 *__restrict__ sq, unsigned sq_cap,
                                               unsigned *__restrict__ totals, unsigned &wbase, unsigned &wused, unsigned &wcap) {
  const unsigned FULL = 0xffffffffu;
  if (!__any_sync(FULL, mask != 0)) return;
  const unsigned c = __popc(mask);
  unsigned incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned t = __shfl_up_sync(FULL, incl, o);
    if ((int)lane >= o) incl += t;
  }
  const unsigned wtot = __shfl_sync(FULL, incl, 31);
  if (wused + wtot > wcap) {
    for (unsigned k2 = wused + lane; k2 < wcap; k2 += 32)
      if (wbase + k2 < sq_cap) sq[wbase + k2] = make_uint2(0xffffffffu, 0u);  // unused tail of the old chunk
    const unsigned want = max((unsigned)MPB_SQ_CHUNK, wtot);
    unsigned nb = 0;
    if (lane == 0) nb = atomicAdd(&totals[6], want);
    wbase = __shfl_sync(FULL, nb, 0);
    wused = 0;
    wcap = want;
  }
  unsigned pos = wbase + wused + incl - c;
  while (mask) {
    const int kk = __ffs(mask) - 1;
    mask &= mask - 1;
    if (pos < sq_cap) sq[pos] = make_uint2(cfg, (unsigned)(gbase + kk));
    pos++;
  }
  wused += wtot;
}

#define POSE(w, i) pose[(size_t)(w) * ld + (i)]
template <int CB>
__global__ void __launch_bounds__(128) sphere_blk_kernel(const __grid_constant__ SceneImg img, const float *__restrict__ pose, size_t ld, long long n,
                                                         long long cfg_base, uint2 *__restrict__ sq, unsigned sq_cap, unsigned *__restrict__ totals) {
  const unsigned lane = threadIdx.x & 31;
  const int cbase = img.w_sphere;
  unsigned wbase = 0, wused = 0, wcap = 0;  // warp-uniform survivor-queue chunk, kept across the warp's tiles
  const long long n_tiles = (n + 31) / 32, n_groups = (n_tiles + CB - 1) / CB;
  const long long warp_stride = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long grp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); grp < n_groups; grp += warp_stride) {
    long long i[CB], ic[CB];
    bool valid[CB];
#pragma unroll
    for (int c = 0; c < CB; c++) {
      i[c] = (grp * CB + c) * 32 + lane;
      valid[c] = i[c] < n;
      ic[c] = valid[c] ? i[c] : n - 1;
    }
    int r = 0;  // run cursor
    for (int gbase = 0; gbase < img.n_pairs; gbase += 32) {
      const int gend = min(gbase + 32, img.n_pairs);
      unsigned mask[CB];
#pragma unroll
      for (int c = 0; c < CB; c++) mask[c] = 0;
      int k = gbase;
      while (k < gend) {
        const int4 run = __ldg(img.runs + r);  // {moving slot or -1, first, count, partner kind}
        const int rlast = run.y + run.z;
        const int rend = min(rlast, gend);
        float ax[CB], ay[CB], az[CB], na2[CB];
#pragma unroll
        for (int c = 0; c < CB; c++) {
          ax[c] = ay[c] = az[c] = 0.f;
          if (run.x >= 0) { ax[c] = POSE(cbase + run.x * 3, ic[c]); ay[c] = POSE(cbase + run.x * 3 + 1, ic[c]); az[c] = POSE(cbase + run.x * 3 + 2, ic[c]); }
          na2[c] = -fmaf(ax[c], ax[c], fmaf(ay[c], ay[c], az[c] * az[c]));
        }
        if (run.w == 0) {  // world-fixed partners, expanded form (see sphere_kernel)
#pragma unroll 4
          for (; k < rend; k++) {
            const float4 q = __ldg(img.cullq + k);
#pragma unroll
            for (int c = 0; c < CB; c++) mask[c] |= (unsigned)(fmaf(ax[c], q.x, fmaf(ay[c], q.y, fmaf(az[c], q.z, q.w))) <= na2[c]) << (k - gbase);
          }
        } else {  // moving partners
          for (; k < rend; k++) {
            const float rr2 = __ldg(&img.cullc[k].w);
            const int m2 = (__ldg(&img.cullm[k].y) >> 16) - 1;
#pragma unroll
            for (int c = 0; c < CB; c++) {
              const float dx = ax[c] - POSE(cbase + m2 * 3, ic[c]), dy = ay[c] - POSE(cbase + m2 * 3 + 1, ic[c]), dz = az[c] - POSE(cbase + m2 * 3 + 2, ic[c]);
              mask[c] |= (unsigned)(fmaf(dx, dx, fmaf(dy, dy, dz * dz)) <= rr2) << (k - gbase);
            }
          }
        }
        if (rend == rlast) r++;
      }
#pragma unroll
      for (int c = 0; c < CB; c++)
        push_survivors(valid[c] ? mask[c] : 0u, (unsigned)(cfg_base + i[c]), gbase, lane, sq, sq_cap, totals, wbase, wused, wcap);
    }
  }
  for (unsigned k2 = wused + lane; k2 < wcap; k2 += 32)
    if (wbase + k2 < sq_cap) sq[wbase + k2] = make_uint2(0xffffffffu, 0u);
}
#undef POSE

// ---- sufficient-hit test in front of the narrowphase.  Every object carries up to MPB_INNER_BALLS balls that lie inside
// it (host: primitive_balls / hull_balls).  If a ball of one shape intersects the other shape -- exactly, by the closed
// forms below, when that one is a primitive; a ball of it otherwise -- the shapes intersect, which is what MPR would
// report after three or more support evaluations.  In the C3 scene most narrowphase items are such collisions.
__device__ __forceinline__ bool ball_hits_primitive(int type, const float *prm, V3 c, float r) {  // c: ball centre in the primitive's frame
  // all four closed forms, selected at the end: the lanes of a warp meet different partner types, and a switch would run
  // each case for a handful of lanes
  const float rp = r + prm[0];
  const float d_sph = dot(c, c) - rp * rp;                                         // sphere: |c| < r + R
  const float bx = c.x - clampf(c.x, -prm[0], prm[0]), by = c.y - clampf(c.y, -prm[1], prm[1]), bz = c.z - clampf(c.z, -prm[2], prm[2]);
  const float d_box = fmaf(bx, bx, fmaf(by, by, bz * bz)) - r * r;                 // box: point-box distance < r
  const float cz = c.z - clampf(c.z, -prm[1], prm[1]);
  const float rad2 = fmaf(c.x, c.x, c.y * c.y);
  const float d_cap = rad2 + cz * cz - rp * rp;                                    // capsule: point-segment distance < r + R
  const float dr = fmaxf(sqrtf(rad2) - prm[0], 0.f), dz = fmaxf(fabsf(c.z) - prm[1], 0.f);
  const float d_cyl = fmaf(dr, dr, dz * dz) - r * r;                               // cylinder: radial / axial excess < r
  const float d = type == MPB_GEOM_SPHERE ? d_sph : (type == MPB_GEOM_BOX ? d_box : (type == MPB_GEOM_CAPSULE ? d_cap : d_cyl));
  return type >= MPB_GEOM_SPHERE && type <= MPB_GEOM_CYLINDER && d < 0.f;
}
// A, B: the two objects of the item; (R, t): B in A.  leaf >= 0: B is a BVH object and the item is its leaf primitive `leaf`.
__device__ __forceinline__ bool sufficient_hit(const SceneImg &img, const DevObject &A, const DevObject &B, const float *R, V3 t, int leaf) {
  if (leaf >= 0) {
    if (B.vcnt != MPB_GEOM_VBOX) return false;  // triangles: left to the narrowphase
    const float4 pc = __ldg(img.prims + 3 * leaf), ph = __ldg(img.prims + 3 * leaf + 1);
#pragma unroll
    for (int k = 0; k < MPB_INNER_BALLS; k++) {
      const float4 b = A.ball[k];
      if (b.w <= 0.f) break;
      const V3 c = rot_t(R, v3(b.x, b.y, b.z) - t);  // ball centre in the octree's frame
      const float dx = c.x - clampf(c.x, pc.x - ph.x, pc.x + ph.x), dy = c.y - clampf(c.y, pc.y - ph.y, pc.y + ph.y),
                  dz = c.z - clampf(c.z, pc.z - ph.z, pc.z + ph.z);
      if (fmaf(dx, dx, fmaf(dy, dy, dz * dz)) < b.w * b.w) return true;
    }
    return false;
  }
  const bool pa = A.type >= MPB_GEOM_SPHERE && A.type <= MPB_GEOM_CYLINDER, pb = B.type >= MPB_GEOM_SPHERE && B.type <= MPB_GEOM_CYLINDER;
  if (pb) {  // balls of A against the primitive B, exactly (an unused ball has radius 0 and a far centre: it cannot hit)
    bool hit = false;
#pragma unroll
    for (int k = 0; k < MPB_INNER_BALLS; k++) {
      const float4 b = A.ball[k];
      hit |= b.w > 0.f && ball_hits_primitive(B.type, B.prm, rot_t(R, v3(b.x, b.y, b.z) - t), b.w);
    }
    return hit;
  }
  if (pa) {  // balls of B against the primitive A
#pragma unroll
    for (int k = 0; k < MPB_INNER_BALLS; k++) {
      const float4 b = B.ball[k];
      if (b.w <= 0.f) break;
      if (ball_hits_primitive(A.type, A.prm, rot(R, v3(b.x, b.y, b.z)) + t, b.w)) return true;
    }
    return false;
  }
  if (A.type != MPB_GEOM_CONVEX || B.type != MPB_GEOM_CONVEX) return false;
#pragma unroll
  for (int j = 0; j < MPB_INNER_BALLS; j++) {
    const float4 bb = B.ball[j];
    if (bb.w <= 0.f) break;
    const V3 cb = rot(R, v3(bb.x, bb.y, bb.z)) + t;
#pragma unroll
    for (int k = 0; k < MPB_INNER_BALLS; k++) {
      const float4 ba = A.ball[k];
      if (ba.w <= 0.f) break;
      const V3 d = v3(ba.x, ba.y, ba.z) - cb;
      const float rr = ba.w + bb.w;
      if (dot(d, d) < rr * rr) return true;
    }
  }
  return false;
}

// sphere phase of a SMALL pass (a legacy single-state query, a few thousand states): thread per (configuration, cull-table
// entry) instead of thread per configuration -- one warp walking the 1587 entries of the C3 table for a single state
// took 0.21 ms, all of it load latency.  Same tests as sphere_kernel<0> (expanded form for world-fixed partners).
#define POSE(w, i) pose[(size_t)(w) * ld + (i)]
__global__ void __launch_bounds__(256) sphere_pairs_kernel(const __grid_constant__ SceneImg img, const float *__restrict__ pose, size_t ld,
                                                           long long n, long long cfg_base, uint2 *__restrict__ sq, unsigned sq_cap,
                                                           unsigned *__restrict__ totals) {
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31;
  const int cbase = img.w_sphere, P = img.n_pairs;
  const long long total = n * (long long)P, total_r = (total + 31) / 32 * 32;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total_r; idx += (long long)gridDim.x * blockDim.x) {
    bool pass = false;
    long long i = 0;
    int k = 0;
    if (idx < total) {
      i = idx / P;
      k = (int)(idx - i * P);
      const int mm = __ldg(&img.cullm[k].y);
      const int m1 = mm & 0xffff, m2 = mm >> 16;  // biased by one; 0 = none
      float ax = 0.f, ay = 0.f, az = 0.f;
      if (m1) { ax = POSE(cbase + (m1 - 1) * 3, i); ay = POSE(cbase + (m1 - 1) * 3 + 1, i); az = POSE(cbase + (m1 - 1) * 3 + 2, i); }
      if (m2) {
        const float rr2 = __ldg(&img.cullc[k].w);
        const float dx = ax - POSE(cbase + (m2 - 1) * 3, i), dy = ay - POSE(cbase + (m2 - 1) * 3 + 1, i), dz = az - POSE(cbase + (m2 - 1) * 3 + 2, i);
        pass = fmaf(dx, dx, fmaf(dy, dy, dz * dz)) <= rr2;
      } else {
        const float4 c = __ldg(img.cullq + k);
        pass = fmaf(ax, c.x, fmaf(ay, c.y, fmaf(az, c.z, c.w))) <= -fmaf(ax, ax, fmaf(ay, ay, az * az));
      }
    }
    const unsigned bm = __ballot_sync(FULL, pass);
    if (!bm) continue;
    unsigned base = 0;
    if (lane == 0) base = atomicAdd(&totals[6], (unsigned)__popc(bm));
    base = __shfl_sync(FULL, base, 0);
    if (pass) {
      const unsigned pos = base + __popc(bm & ((1u << lane) - 1u));
      if (pos < sq_cap) sq[pos] = make_uint2((unsigned)(cfg_base + i), (unsigned)k);
    }
  }
}
#undef POSE

#ifndef MPB_DENSE_MINB
#define MPB_DENSE_MINB 6  // 80 registers: six blocks of 128 threads per SM (measured best of 1 / 6: cull leg 13.6 -> 12.6 ms on C3)
#endif
template <int MODE>
__global__ void __launch_bounds__(128, MPB_DENSE_MINB) dense_kernel(SceneImg img, const float4 *__restrict__ obb, size_t ld, const uint2 *__restrict__ sq,
                                                    unsigned sq_cap, WorkQ wq, HitOut out, int presolve) {
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31;
  __shared__ unsigned s_bins[MPB_SMEM_BINS];  // per-block bin counts, flushed once at the end
  const bool smem_bins = img.n_bins <= MPB_SMEM_BINS;
  for (int b = threadIdx.x; b < MPB_SMEM_BINS; b += blockDim.x) s_bins[b] = 0;
  __syncthreads();
  unsigned wbase = 0xffffffffu, wused = MPB_STAGE_CHUNK;  // warp-uniform staging chunk state
  const unsigned total = min(wq.totals[6], sq_cap);
  const unsigned total_r = (total + 31u) & ~31u;
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&wq.counters[3], (unsigned long long)total);  // incl. chunk padding
  const unsigned e_first = blockIdx.x * blockDim.x + threadIdx.x, e_stride = gridDim.x * blockDim.x;
  uint2 ent_next = make_uint2(0xffffffffu, 0u);  // the survivor record of the next trip is fetched one trip ahead
  if (e_first < total) ent_next = sq[e_first];
  for (unsigned e = e_first; e < total_r; e += e_stride) {
    bool keep = false;
    Xf rel;
    int p = 0;
    unsigned cfg = 0;
    const uint2 ent = ent_next;
    ent_next = make_uint2(0xffffffffu, 0u);
    if (e + e_stride < total) ent_next = sq[e + e_stride];
    if (ent.x != 0xffffffffu) {
      cfg = ent.x;
      p = __ldg(&img.cullm[ent.y].x);
      const DevPair pr = img.pairs[p];
      // cull records (MPB_COBB_WORDS float4 per object): [0..2] OBB frame (world frame if the object is fixed,
      // owner-joint frame if it moves), [3] half extents + inscribed radius, [4] inscribed-ball centre in the OBB frame
      // + moving slot, [5..7] OBB frame in the object frame, [8..10] its inverse
      const float4 *ca = img.cobb + MPB_COBB_WORDS * pr.a, *cb = img.cobb + MPB_COBB_WORDS * pr.b;
      const float4 ha = __ldg(ca + 3), hb = __ldg(cb + 3), ia = __ldg(ca + 4), ib = __ldg(cb + 4);
      const int da = __float_as_int(ia.w), db = __float_as_int(ib.w);
      // moving: this configuration's row of the OBB table (quaternion, translation); fixed: the record's world frame
      Xf Ea, Eb;
      if (da >= 0) { const float4 *pa = obb + ((size_t)da * ld + cfg) * 2; Ea = quat_rows(pa[0], pa[1]); }
      else Ea = xf_rows(__ldg(ca), __ldg(ca + 1), __ldg(ca + 2));
      if (db >= 0) { const float4 *pb = obb + ((size_t)db * ld + cfg) * 2; Eb = quat_rows(pb[0], pb[1]); }
      else Eb = xf_rows(__ldg(cb), __ldg(cb + 1), __ldg(cb + 2));
      // box B in box A's frame:  Ra^T Rb,  Ra^T (cb - ca)  ==  (OBB frame A)^-1 (OBB frame B)
      Xf Rt;
      const float d0 = Eb.p[0] - Ea.p[0], d1 = Eb.p[1] - Ea.p[1], d2 = Eb.p[2] - Ea.p[2];
#pragma unroll
      for (int rr = 0; rr < 3; rr++) {
#pragma unroll
        for (int c = 0; c < 3; c++) Rt.r[3 * rr + c] = fmaf(Ea.r[rr], Eb.r[c], fmaf(Ea.r[3 + rr], Eb.r[3 + c], Ea.r[6 + rr] * Eb.r[6 + c]));
        Rt.p[rr] = fmaf(Ea.r[rr], d0, fmaf(Ea.r[3 + rr], d1, Ea.r[6 + rr] * d2));
      }
      keep = true;
      if (MODE == 0) {
        const float ah[3] = {ha.x, ha.y, ha.z}, bh[3] = {hb.x, hb.y, hb.z};
        keep = !obb_disjoint(Rt.r, Rt.p, ah, bh);
        if (keep && ha.w > 0.f && hb.w > 0.f) {
          // sufficient test: balls inscribed in the two shapes overlap -> the pair collides, no narrowphase needed
          const float ib3[3] = {ib.x, ib.y, ib.z};
          float icb[3];
          xf_point(Rt, ib3, icb);  // B's ball centre in A's OBB frame
          const float e0 = ia.x - icb[0], e1 = ia.y - icb[1], e2 = ia.z - icb[2];
          const float rs = 0.999f * (ha.w + hb.w);
          if (fmaf(e0, e0, fmaf(e1, e1, e2 * e2)) < rs * rs) {
            keep = false;
            report_hit(out, cfg, p);
          }
        }
      }
      if (keep) {  // relative pose of the two OBJECT frames for the narrowphase: Xa (A_obb^-1 B_obb) Xb^-1
        const Xf Xa = xf_rows(__ldg(ca + 5), __ldg(ca + 6), __ldg(ca + 7));
        const Xf Xbi = xf_rows(__ldg(cb + 8), __ldg(cb + 9), __ldg(cb + 10));
        rel = xf_mul(xf_mul(Xa, Rt), Xbi);
        // sufficient test with the inner balls (exact closed forms against a primitive partner): an item it settles is
        // never staged, binned or sent to the narrowphase (what matters: the narrowphase's prologue costs per ITEM)
        if (MODE == 0 && presolve && (pr.bin_dyn & 0xffff) != img.expand_bin && (pr.bin_dyn & 0xffff) != img.expand2_bin &&
            sufficient_hit(img, img.objs[pr.a], img.objs[pr.b], rel.r, v3(rel.p[0], rel.p[1], rel.p[2]), -1)) {
          keep = false;
          report_hit(out, cfg, p);
        }
      }
    }
    const unsigned km = __ballot_sync(FULL, keep);
    if (km == 0) continue;
    // warp-private chunk of staging slots: one same-address atomic per MPB_STAGE_CHUNK items
    const unsigned cntk = (unsigned)__popc(km);
    if (wused + cntk > MPB_STAGE_CHUNK) {
      if (wbase != 0xffffffffu)
        for (unsigned k = wused + lane; k < MPB_STAGE_CHUNK; k += 32)
          if (wbase + k < wq.cap) wq.skey[wbase + k] = MPB_KEY_HOLE;  // unused tail of the old chunk
      unsigned nb = 0;
      if (lane == 0) nb = atomicAdd(&wq.totals[0], (unsigned)MPB_STAGE_CHUNK);
      wbase = __shfl_sync(FULL, nb, 0);
      wused = 0;
    }
    if (keep) {
      const unsigned slot = wbase + wused + __popc(km & ((1u << lane) - 1u));
      const int bin = img.pairs[p].bin_dyn & 0xffff;
      if (slot < wq.cap) {
        item_store(wq, 0, slot, rel.r, rel.p, (int)cfg, p, 0, bin);
      }
      if (slot < wq.cap) {
        if (smem_bins) atomicAdd(&s_bins[bin], 1u);
        else atomicAdd(&wq.bin_count[bin], 1u);
      }
    }
    wused += cntk;
  }
  // invalidate the unused tail of the last chunk
  if (wbase != 0xffffffffu) {
    for (unsigned k = wused + lane; k < MPB_STAGE_CHUNK; k += 32)
      if (wbase + k < wq.cap) wq.skey[wbase + k] = MPB_KEY_HOLE;
  }
  __syncthreads();
  if (smem_bins)
    for (int b = threadIdx.x; b < img.n_bins; b += blockDim.x)
      if (s_bins[b]) atomicAdd(&wq.bin_count[b], s_bins[b]);
}
#undef POSE

// ---------------------------------------------------------------------------------------------- dense cull, two phases
// dense_kernel<0> does everything for a survivor in one go: SAT, relative pose, sufficient-hit test, staging.  ncu showed
// 11-14 of 32 lanes per instruction: half of a warp's survivors fail the SAT and idle through the longer second half
// (profiles/r02_ncu_dense_presolve_source.csv).  Here a warp runs PHASE 1 (frames, SAT, inscribed-ball test) on 32
// survivors and appends what is kept -- the relative OBB frame Rt, configuration and pair, 14 words -- to a per-warp
// buffer in shared memory; whenever the buffer holds 32 entries, PHASE 2 runs on them with all lanes busy: relative
// pose of the object frames, the (branch-free) sufficient-hit test, staging of what is left.  Nothing is loaded twice.
// (A first version buffered only (configuration, pair) and re-derived the frames in phase 2: the OBB rows are random
// 32-byte reads, and fetching them twice cost more than the idle lanes had.)
#define MPB_D2_CAP 64    // buffer entries per warp: phase 2 runs at 32, a phase-1 trip adds at most 32
#define MPB_D2_F 14      // Rt (12 floats), configuration, pair

#ifndef MPB_DENSE2_MINB
#define MPB_DENSE2_MINB 8  // 64 registers, no spills: eight blocks per SM (cull leg of a C3 step 7.40 -> 7.22 ms against six blocks at 72 registers)
#endif
__global__ void __launch_bounds__(128, MPB_DENSE2_MINB) dense2_kernel(SceneImg img, const float4 *__restrict__ obb, size_t ld, const uint2 *__restrict__ sq,
                                                                      unsigned sq_cap, WorkQ wq, HitOut out, int presolve) {
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ unsigned s_bins[MPB_SMEM_BINS];  // per-block bin counts, flushed once at the end
  __shared__ float s_buf[4][MPB_D2_F][MPB_D2_CAP];
  const bool smem_bins = img.n_bins <= MPB_SMEM_BINS;
  for (int b = threadIdx.x; b < MPB_SMEM_BINS; b += blockDim.x) s_bins[b] = 0;
  __syncthreads();
  float (*buf)[MPB_D2_CAP] = s_buf[warp];
  unsigned wbase = 0xffffffffu, wused = MPB_STAGE_CHUNK;  // warp-uniform staging chunk state
  int cnt = 0;                                             // entries in the warp's buffer (warp-uniform)
  const unsigned total = min(wq.totals[6], sq_cap);
  const unsigned total_r = (total + 31u) & ~31u;
  if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&wq.counters[3], (unsigned long long)total);  // incl. chunk padding

  // phase 2 on the top `nb` entries of the buffer
  auto phase2 = [&](int nb) {
    bool keep = false;
    unsigned cfg = 0;
    int p = 0;
    Xf rel;
    if ((int)lane < nb) {
      const int e = cnt - nb + (int)lane;
      Xf Rt;
#pragma unroll
      for (int k = 0; k < 9; k++) Rt.r[k] = buf[k][e];
#pragma unroll
      for (int k = 0; k < 3; k++) Rt.p[k] = buf[9 + k][e];
      cfg = __float_as_uint(buf[12][e]);
      p = __float_as_int(buf[13][e]);
      const DevPair pr = img.pairs[p];
      const float4 *ca = img.cobb + MPB_COBB_WORDS * pr.a, *cb = img.cobb + MPB_COBB_WORDS * pr.b;
      // relative pose of the two OBJECT frames for the narrowphase: Xa (A_obb^-1 B_obb) Xb^-1
      const Xf Xa = xf_rows(__ldg(ca + 5), __ldg(ca + 6), __ldg(ca + 7));
      const Xf Xbi = xf_rows(__ldg(cb + 8), __ldg(cb + 9), __ldg(cb + 10));
      rel = xf_mul(xf_mul(Xa, Rt), Xbi);
      keep = true;
      const int bin = pr.bin_dyn & 0xffff;
      if (presolve && bin != img.expand_bin && bin != img.expand2_bin &&
          sufficient_hit(img, img.objs[pr.a], img.objs[pr.b], rel.r, v3(rel.p[0], rel.p[1], rel.p[2]), -1)) {
        keep = false;
        report_hit(out, cfg, p);
      }
    }
    const unsigned km = __ballot_sync(FULL, keep);
    if (km) {
      // warp-private chunk of staging slots: one same-address atomic per MPB_STAGE_CHUNK items
      const unsigned cntk = (unsigned)__popc(km);
      if (wused + cntk > MPB_STAGE_CHUNK) {
        if (wbase != 0xffffffffu)
          for (unsigned k = wused + lane; k < MPB_STAGE_CHUNK; k += 32)
            if (wbase + k < wq.cap) wq.skey[wbase + k] = MPB_KEY_HOLE;  // unused tail of the old chunk
        unsigned nbs = 0;
        if (lane == 0) nbs = atomicAdd(&wq.totals[0], (unsigned)MPB_STAGE_CHUNK);
        wbase = __shfl_sync(FULL, nbs, 0);
        wused = 0;
      }
      if (keep) {
        const unsigned slot = wbase + wused + __popc(km & ((1u << lane) - 1u));
        if (slot < wq.cap) {
          const int bin = img.pairs[p].bin_dyn & 0xffff;
          item_store(wq, 0, slot, rel.r, rel.p, (int)cfg, p, 0, bin);
          if (smem_bins) atomicAdd(&s_bins[bin], 1u);
          else atomicAdd(&wq.bin_count[bin], 1u);
        }
      }
      wused += cntk;
    }
    cnt -= nb;
    __syncwarp();
  };

  const unsigned e_first = blockIdx.x * blockDim.x + threadIdx.x, e_stride = gridDim.x * blockDim.x;
  uint2 ent_next = make_uint2(0xffffffffu, 0u);  // the survivor record of the next trip is fetched one trip ahead
  if (e_first < total) ent_next = sq[e_first];
  for (unsigned e = e_first; e < total_r; e += e_stride) {
    // ---- phase 1: frames, SAT, inscribed balls
    bool keep = false;
    int p = 0;
    unsigned cfg = 0;
    Xf Rt;
    const uint2 ent = ent_next;
    ent_next = make_uint2(0xffffffffu, 0u);
    if (e + e_stride < total) ent_next = sq[e + e_stride];
    if (ent.x != 0xffffffffu) {
      cfg = ent.x;
      // one 16-byte load names the pair, its objects and their moving slots: the OBB rows (random 32-byte reads, the long
      // latency of this kernel) are requested together with the cull records instead of two dependent loads later
      const int4 cp = __ldg(img.cullp + ent.y);
      p = cp.x;
      const int da = (cp.w & 0xffff) - 1, db = (cp.w >> 16) - 1;
      // cull records (MPB_COBB_WORDS float4 per object): [0..2] OBB frame (world frame if the object is fixed,
      // owner-joint frame if it moves), [3] half extents + inscribed radius, [4] inscribed-ball centre in the OBB frame
      // + moving slot, [5..7] OBB frame in the object frame, [8..10] its inverse
      const float4 *ca = img.cobb + MPB_COBB_WORDS * cp.y, *cb = img.cobb + MPB_COBB_WORDS * cp.z;
      // moving: this configuration's row of the OBB table (quaternion, translation); fixed: the record's world frame
      const float4 *pa = da >= 0 ? obb + ((size_t)da * ld + cfg) * 2 : ca, *pb = db >= 0 ? obb + ((size_t)db * ld + cfg) * 2 : cb;
      const float4 a0 = __ldg(pa), a1 = __ldg(pa + 1), b0 = __ldg(pb), b1 = __ldg(pb + 1);
      const float4 ha = __ldg(ca + 3), hb = __ldg(cb + 3), ia = __ldg(ca + 4), ib = __ldg(cb + 4);
      Xf Ea, Eb;
      if (da >= 0) Ea = quat_rows(a0, a1);
      else Ea = xf_rows(a0, a1, __ldg(ca + 2));
      if (db >= 0) Eb = quat_rows(b0, b1);
      else Eb = xf_rows(b0, b1, __ldg(cb + 2));
      // box B in box A's frame:  Ra^T Rb,  Ra^T (cb - ca)  ==  (OBB frame A)^-1 (OBB frame B)
      const float d0 = Eb.p[0] - Ea.p[0], d1 = Eb.p[1] - Ea.p[1], d2 = Eb.p[2] - Ea.p[2];
#pragma unroll
      for (int rr = 0; rr < 3; rr++) {
#pragma unroll
        for (int c = 0; c < 3; c++) Rt.r[3 * rr + c] = fmaf(Ea.r[rr], Eb.r[c], fmaf(Ea.r[3 + rr], Eb.r[3 + c], Ea.r[6 + rr] * Eb.r[6 + c]));
        Rt.p[rr] = fmaf(Ea.r[rr], d0, fmaf(Ea.r[3 + rr], d1, Ea.r[6 + rr] * d2));
      }
      const float ah[3] = {ha.x, ha.y, ha.z}, bh[3] = {hb.x, hb.y, hb.z};
      keep = !obb_disjoint(Rt.r, Rt.p, ah, bh);
      if (keep && ha.w > 0.f && hb.w > 0.f) {
        // sufficient test: balls inscribed in the two shapes overlap -> the pair collides, no narrowphase needed
        const float ib3[3] = {ib.x, ib.y, ib.z};
        float icb[3];
        xf_point(Rt, ib3, icb);  // B's ball centre in A's OBB frame
        const float e0 = ia.x - icb[0], e1 = ia.y - icb[1], e2 = ia.z - icb[2];
        const float rs = 0.999f * (ha.w + hb.w);
        if (fmaf(e0, e0, fmaf(e1, e1, e2 * e2)) < rs * rs) {
          keep = false;
          report_hit(out, cfg, p);
        }
      }
    }
    // ---- kept survivors -> the warp's buffer; 32 entries run phase 2
    const unsigned m = __ballot_sync(FULL, keep);
    if (!m) continue;
    if (keep) {
      const int e2 = cnt + __popc(m & ((1u << lane) - 1u));
#pragma unroll
      for (int k = 0; k < 9; k++) buf[k][e2] = Rt.r[k];
#pragma unroll
      for (int k = 0; k < 3; k++) buf[9 + k][e2] = Rt.p[k];
      buf[12][e2] = __uint_as_float(cfg);
      buf[13][e2] = __int_as_float(p);
    }
    cnt += __popc(m);
    __syncwarp();
    if (cnt >= 32) phase2(32);
  }
  while (cnt > 0) phase2(min(cnt, 32));
  // invalidate the unused tail of the last chunk
  if (wbase != 0xffffffffu) {
    for (unsigned k = wused + lane; k < MPB_STAGE_CHUNK; k += 32)
      if (wbase + k < wq.cap) wq.skey[wbase + k] = MPB_KEY_HOLE;
  }
  __syncthreads();
  if (smem_bins)
    for (int b = threadIdx.x; b < img.n_bins; b += blockDim.x)
      if (s_bins[b]) atomicAdd(&wq.bin_count[b], s_bins[b]);
}

// start of a pass (or of its second narrowphase round): bin counts, staged-item and survivor counters back to zero
__global__ void pass_reset_kernel(WorkQ wq, int n_bins) {
  for (int b = threadIdx.x; b < n_bins; b += blockDim.x) wq.bin_count[b] = 0;
  if (threadIdx.x == 0) { wq.totals[0] = 0; wq.totals[6] = 0; }
}

// the same for a small pass, plus the zeroing of its outputs (and, at the start of a call, of the call's counters): one
// launch instead of a reset kernel and up to five memsets -- a legacy single-state query is bounded by its node count
__global__ void __launch_bounds__(256) small_reset_kernel(WorkQ wq, int n_bins, unsigned *__restrict__ verdict, unsigned v_words,
                                                          unsigned *__restrict__ pair_bits, unsigned p_words, unsigned *__restrict__ n_hits,
                                                          int reset_call) {
  if (blockIdx.x == 0) {
    for (int b = threadIdx.x; b < n_bins; b += blockDim.x) wq.bin_count[b] = 0;
    if (threadIdx.x < 8 && (reset_call || threadIdx.x == 0 || threadIdx.x == 6)) wq.totals[threadIdx.x] = 0;
    if (reset_call && threadIdx.x < 4) wq.counters[threadIdx.x] = 0;
    if (reset_call && n_hits && threadIdx.x == 0) *n_hits = 0;
  }
  const unsigned stride = gridDim.x * blockDim.x, i0 = blockIdx.x * blockDim.x + threadIdx.x;
  for (unsigned i = i0; i < v_words; i += stride) verdict[i] = 0;
  if (pair_bits)
    for (unsigned i = i0; i < p_words; i += stride) pair_bits[i] = 0;
}

__global__ void scan_bins_kernel(WorkQ wq, int n_bins, int skip_bin, int skip_bin2, unsigned n_warps, unsigned sq_cap) {
  // one warp: bins in groups of 32 (count -> warp prefix sums), the running totals carried in registers
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  if (blockIdx.x != 0 || threadIdx.x >= 32) return;
  // unit = slice of one bin a warp owns: as large as possible (persistent lanes keep all 32 lanes busy across the
  // spread of MPR run lengths; the tail of a slice runs in the cooperative mode) while still feeding n_warps
  unsigned all = 0;
  for (int b = lane; b < n_bins; b += 32)
    if (b != skip_bin && b != skip_bin2) all += wq.bin_count[b];
  all = __reduce_add_sync(FULL, all);
  unsigned unit = MPB_UNIT_ITEMS;
  while (unit > 2u && all / unit < n_warps - n_warps / 4) unit >>= 1;  // few items (a planning round): short slices -> cooperative supports
  unsigned items = 0, units = 0;
  for (int b0 = 0; b0 < n_bins; b0 += 32) {
    const int b = b0 + lane;
    const unsigned c = b < n_bins ? wq.bin_count[b] : 0u;
    const unsigned u = (b < n_bins && b != skip_bin && b != skip_bin2) ? (c + unit - 1u) / unit : 0u;  // BVH pairs go to expand_kernel / expand2_kernel instead
    unsigned ic = c, iu = u;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned t1 = __shfl_up_sync(FULL, ic, o), t2 = __shfl_up_sync(FULL, iu, o);
      if (lane >= o) { ic += t1; iu += t2; }
    }
    if (b < n_bins) {
      wq.bin_start[b] = items + ic - c;
      wq.unit_start[b] = units + iu - u;
      wq.bin_cursor[b] = 0;
    }
    items += __shfl_sync(FULL, ic, 31);
    units += __shfl_sync(FULL, iu, 31);
  }
  if (lane == 0) {
    wq.totals[4] = unit;
    wq.bin_start[n_bins] = items;
    wq.unit_start[n_bins] = units;
    wq.totals[1] = units;
    wq.totals[3] = 0;
    wq.totals[5] = 0;  // expand_kernel's item ticket
    wq.totals[7] = 0;  // expand2_kernel's item ticket
    if (wq.totals[0] > wq.cap || wq.totals[6] > sq_cap) wq.totals[2] = 1u;  // queue overflow: the host redoes the pass in smaller chunks
    atomicAdd(wq.counters, (unsigned long long)items);
  }
}

// Produced order -> binned order.  A block ranks the bin keys of its slots in shared memory (the shared atomic returns the
// rank), reserves one range per bin per trip in the global cursors, then writes order[position] = slot: six bytes per
// item.  The sufficient-hit test of the SECOND round (leaf items: inner balls of the shape against a voxel box) runs here
// (PRESOLVE), on the record read in produced order: an item it settles sets its result bits and gets no position;
// bin_cursor[b] ends as the number of items bin b really holds.  (First-round items are tested in the dense cull, before
// they are recorded at all.)
template <bool PRESOLVE>
__global__ void __launch_bounds__(PRESOLVE ? 512 : 1024, 2) scatter_kernel(SceneImg img, WorkQ wq, int round, HitOut out) {
  __shared__ unsigned s_cnt[MPB_SMEM_BINS], s_base[MPB_SMEM_BINS];
  const unsigned n = min(wq.totals[0], wq.cap);
  const int n_bins = round == 0 ? img.n_bins : img.n_bins2;
  const bool smem_bins = n_bins <= MPB_SMEM_BINS;
  const unsigned n_round = (n + blockDim.x - 1) / blockDim.x * blockDim.x;
  const unsigned lane = threadIdx.x & 31;
  for (unsigned s = blockIdx.x * blockDim.x + threadIdx.x; s < n_round; s += gridDim.x * blockDim.x) {
    int bin = s < n ? (int)wq.skey[s] : (int)MPB_KEY_HOLE;  // holes are the unused tails of warp chunks
    bool valid = bin != (int)MPB_KEY_HOLE;
    if (!valid) bin = -1;
    if constexpr (PRESOLVE) {
      if (valid && !(round == 0 && (bin == img.expand_bin || bin == img.expand2_bin))) {
        const float4 *r = wq.rec[round] + 4 * (size_t)s;
        float R[9], t3[3];
        int cfg, p, aux;
        item_pose(r, R, t3);
        item_ids(r, cfg, p, aux);
        const DevPair pr = img.pairs[p];
        if (sufficient_hit(img, img.objs[pr.a], img.objs[pr.b], R, v3(t3[0], t3[1], t3[2]), round == 0 ? -1 : aux)) {
          report_hit(out, (unsigned)cfg, p);
          valid = false;
          bin = -1;
        }
      }
    }
    unsigned dst = 0;
    if (smem_bins) {
      for (int b = threadIdx.x; b < n_bins; b += blockDim.x) s_cnt[b] = 0;
      __syncthreads();
      unsigned rank = 0;
      if (valid) rank = atomicAdd(&s_cnt[bin], 1u);
      __syncthreads();
      for (int b = threadIdx.x; b < n_bins; b += blockDim.x)
        if (s_cnt[b]) s_base[b] = atomicAdd(&wq.bin_cursor[b], s_cnt[b]);
      __syncthreads();
      if (valid) dst = wq.bin_start[bin] + s_base[bin] + rank;
      __syncthreads();  // s_cnt / s_base are rewritten by the next trip
    } else {
      if (!valid) bin = -1 - (int)lane;
      const unsigned grp = __match_any_sync(0xffffffffu, bin);
      const int leader = __ffs(grp) - 1;
      unsigned base = 0;
      if (valid && (int)lane == leader) base = atomicAdd(&wq.bin_cursor[bin], (unsigned)__popc(grp));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (valid) dst = wq.bin_start[bin] + base + __popc(grp & ((1u << lane) - 1u));
    }
    if (valid) wq.order[dst] = s;
  }
}

// ---------------------------------------------------------------------------------------------- supports
struct DevShape {
  int type, voff, vcnt, cell;
  float p0, p1, p2;
};
__device__ __forceinline__ DevShape shape_of(const DevObject &O) {
  DevShape s;
  s.type = O.type; s.voff = O.voff; s.vcnt = O.vcnt; s.cell = O.cell; s.p0 = O.prm[0]; s.p1 = O.prm[1]; s.p2 = O.prm[2];
  return s;
}
__device__ __forceinline__ float sgn_eps(float v) { return fabsf(v) < MPB_EPS ? 0.f : (v < 0.f ? -1.f : 1.f); }

// extreme point of the shape along `d` in the shape's own frame ([upstream FCL support* functions])
template <bool SMEM>
__device__ __forceinline__ V3 support_local(const DevShape &s, V3 d, const float4 *__restrict__ verts,
                                         const float4 *__restrict__ prims, const unsigned char *__restrict__ cells) {
  if (s.type == MPB_GEOM_CONVEX) {
    const float4 *v = verts + s.voff;
    float best = -FLT_MAX;
    int bi = 0;
    if (s.cell >= 0) {  // candidates of the direction's cube-map cell; same arg-max (and tie order) as the full scan
      const unsigned char *tb = cells + s.cell;
      const unsigned short *first = reinterpret_cast<const unsigned short *>(tb);
      const unsigned char *cand = tb + 2 * (MPB_CELL_COUNT + 1);
      const int cell = mpb_cube_cell(d.x, d.y, d.z);
      const int c0 = first[cell], c1 = first[cell + 1];
      bi = cand[c0];
      for (int c = c0; c < c1; c++) {
        const int i = cand[c];
        const float4 p = SMEM ? v[i] : __ldg(v + i);
        const float t = fmaf(p.x, d.x, fmaf(p.y, d.y, p.z * d.z));
        if (t > best) { best = t; bi = i; }
      }
      const float4 p = SMEM ? v[bi] : __ldg(v + bi);
      return v3(p.x, p.y, p.z);
    }
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
    case MPB_GEOM_TRI: {  // [upstream] supportTriangle: the vertex with the largest projection
      const float4 a = __ldg(prims + 3 * s.voff), b = __ldg(prims + 3 * s.voff + 1), c = __ldg(prims + 3 * s.voff + 2);
      const float da = fmaf(a.x, d.x, fmaf(a.y, d.y, a.z * d.z)), db = fmaf(b.x, d.x, fmaf(b.y, d.y, b.z * d.z)),
                  dc = fmaf(c.x, d.x, fmaf(c.y, d.y, c.z * d.z));
      float4 m = a;
      float dm = da;
      if (db > dm) { m = b; dm = db; }
      if (dc > dm) m = c;
      return v3(m.x, m.y, m.z);
    }
    case MPB_GEOM_VBOX: {  // voxel box: supportBox shifted to the cell centre
      const float4 c = __ldg(prims + 3 * s.voff), h = __ldg(prims + 3 * s.voff + 1);
      return v3(fmaf(sgn_eps(d.x), h.x, c.x), fmaf(sgn_eps(d.y), h.y, c.y), fmaf(sgn_eps(d.z), h.z, c.z));
    }
    default:
      return v3(0.f, 0.f, 0.f);
  }
}

__device__ __forceinline__ bool is_zero(float v) { return fabsf(v) < MPB_EPS; }

// Hull support evaluated by the whole warp for ONE query direction (same on every lane): lane-strided scan and a
// shuffle arg-max.  Ties resolve to the lowest vertex index, exactly like the serial scan of support_local.
template <bool SMEM>
__device__ __forceinline__ V3 support_hull_coop(const float4 *__restrict__ v, int vcnt, V3 d, unsigned lane) {
  float best = -FLT_MAX;
  int bi = 0x7fffffff;
  for (int i = (int)lane; i < vcnt; i += 32) {
    const float4 p = SMEM ? v[i] : __ldg(v + i);
    const float t = fmaf(p.x, d.x, fmaf(p.y, d.y, p.z * d.z));
    if (t > best) { best = t; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  const float4 p = SMEM ? v[bi] : __ldg(v + bi);
  return v3(p.x, p.y, p.z);
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
// ccdMPRIntersect [upstream libccd v2.1 mpr.c] in the frame of shape A; B sits at (R, t).
// Phases: 1 = v1, 2 = v2, 3 = v3 search (discoverPortal), 4 = refinePortal.  One support per trip.
#ifndef MPB_NARROW_MINB
#define MPB_NARROW_MINB 3
#endif
template <bool SMEM>
__global__ void __launch_bounds__(256, MPB_NARROW_MINB) narrow_kernel(SceneImg img, WorkQ wq, HitOut out, int n_bins, int round, int refill_min) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) unsigned long long bar;
  const float4 *verts = img.verts;
  if (SMEM) {
    const unsigned total = (unsigned)img.pool_bytes;
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
    verts = reinterpret_cast<const float4 *>(smem_raw);
  }
  const unsigned char *cells = reinterpret_cast<const unsigned char *>(verts + img.n_verts);
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31;
  const unsigned n_units = wq.totals[1], unit_items = wq.totals[4];
  unsigned long long it_sum = 0, cap_sum = 0;
  for (;;) {
    // dynamic unit ticket (one atomic per warp per MPB_UNIT_ITEMS items)
    unsigned u = 0;
    if (lane == 0) u = atomicAdd(&wq.totals[3], 1u);
    u = __shfl_sync(FULL, u, 0);
    if (u >= n_units) break;
    int lo = 0, hi = n_bins;  // bin of this unit: largest b with unit_start[b] <= u
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (wq.unit_start[mid] <= u) lo = mid; else hi = mid;
    }
    unsigned next = wq.bin_start[lo] + (u - wq.unit_start[lo]) * unit_items;
    // the bin holds bin_cursor items (the scatter's sufficient-hit test may have settled some of the counted ones)
    const unsigned end = min(min(wq.bin_start[lo] + wq.bin_cursor[lo], next + unit_items), wq.cap);
    if (next >= end) continue;
    const unsigned first = next;

    // per-lane item state
    bool busy = false;
    float R[9];
    V3 t = v3(0.f, 0.f, 0.f), v0 = t, v1 = t, v2 = t, v3p = t, dir = v3(1.f, 0.f, 0.f), cA = t, cBl = t;
    DevShape A, B;
    int cfg = 0, p = 0, phase = 1, iters = 0;
    {
      // shapes are uniform over the bin: read them once from the first item of the slice
      int c0_, p0_, a0_;
      item_ids(item_rec(wq, round, first), c0_, p0_, a0_);
      const DevPair pr0 = img.pairs[p0_];
      const DevObject &OA = img.objs[pr0.a];
      const DevObject &OB = img.objs[pr0.b];
      A = shape_of(OA); B = shape_of(OB);
      if (round == 1) B.type = OB.vcnt;  // leaf primitives of a BVH object: MPB_GEOM_TRI / MPB_GEOM_VBOX
#pragma unroll
      for (int k = 0; k < 9; k++) R[k] = (k % 4 == 0) ? 1.f : 0.f;
    }
    for (;;) {
      // ---- refill idle lanes from the slice
      // the refill block is as long as a support evaluation: it runs when at least MPB_REFILL_MIN lanes are idle (or none has
      // work), not every time one lane finishes
      const unsigned need = __ballot_sync(FULL, !busy);
      if (need == FULL || (__popc(need) >= refill_min && next < end)) {
        const unsigned my = next + __popc(need & ((1u << lane) - 1u));
        if (!busy && my < end) {
          bool busy_ok = true;
          const float4 *rc = item_rec(wq, round, my);
          float t3[3];
          int aux;
          item_pose(rc, R, t3);
          item_ids(rc, cfg, p, aux);
          t = v3(t3[0], t3[1], t3[2]);
          const DevPair pr = img.pairs[p];
          const DevObject &OA = img.objs[pr.a];
          const DevObject &OB = img.objs[pr.b];
          A.p0 = OA.prm[0]; A.p1 = OA.prm[1]; A.p2 = OA.prm[2];  // primitive parameters differ per item
          B.p0 = OB.prm[0]; B.p1 = OB.prm[1]; B.p2 = OB.prm[2];
          cA = v3(OA.ctr[0], OA.ctr[1], OA.ctr[2]);
          cBl = v3(OB.ctr[0], OB.ctr[1], OB.ctr[2]);
          if (round == 1) {  // [upstream] centerTriangle = centroid; voxel Box centre = its pose
            B.voff = aux;
            const float4 a = __ldg(img.prims + 3 * B.voff);
            if (B.type == MPB_GEOM_TRI) {
              const float4 b = __ldg(img.prims + 3 * B.voff + 1), c = __ldg(img.prims + 3 * B.voff + 2);
              cBl = v3((a.x + b.x + c.x) * (1.f / 3.f), (a.y + b.y + c.y) * (1.f / 3.f), (a.z + b.z + c.z) * (1.f / 3.f));
            } else cBl = v3(a.x, a.y, a.z);
            // early exit: this (configuration, pair) is already known to collide
            busy_ok = !hit_known(out, (unsigned)cfg, p);
          }
          v0 = cA - (rot(R, cBl) + t);
          if (is_zero(v0.x) && is_zero(v0.y) && is_zero(v0.z)) v0.x += 10.f * MPB_EPS;
          dir = normalized(v0 * -1.f);
          phase = 1;
          iters = 0;
          busy = !(round == 1 && !busy_ok);
        }
        next += __popc(need);
        if (__ballot_sync(FULL, busy) == 0) {
          // no lane has work: either the slice is exhausted, or every item just fetched was skipped as already known
          // (second round: the leaf items of one (configuration, pair) sit next to each other) -- then fetch again
          if (next >= end) break;
          continue;
        }
      }
      // ---- one Minkowski support for every lane: sA(dir) - (R sB(R^T(-dir)) + t)
      const unsigned bm = __ballot_sync(FULL, busy);
      const V3 dB = rot_t(R, dir * -1.f);
      V3 sa, sb;
      if (__popc(bm) <= MPB_COOP_MAX) {
        // few lanes left (the tail of a slice): the whole warp scans the hulls for one busy lane at a time
        // (lane-strided vertices + shuffle arg-max) instead of every lane scanning every vertex for itself
        // warp-uniform (bin-uniform shapes); hulls with a support cell table scan a handful of candidates per lane anyway
        const bool ca = A.type == MPB_GEOM_CONVEX && A.cell < 0, cb = B.type == MPB_GEOM_CONVEX && B.cell < 0;
        if (!ca) sa = support_local<SMEM>(A, dir, verts, img.prims, cells);
        if (!cb) sb = support_local<SMEM>(B, dB, verts, img.prims, cells);
        for (unsigned m = bm; m; m &= m - 1) {
          const int src = __ffs(m) - 1;
          if (ca) {
            const V3 d = v3(__shfl_sync(FULL, dir.x, src), __shfl_sync(FULL, dir.y, src), __shfl_sync(FULL, dir.z, src));
            const V3 r = support_hull_coop<SMEM>(verts + A.voff, A.vcnt, d, lane);
            if ((int)lane == src) sa = r;
          }
          if (cb) {
            const V3 d = v3(__shfl_sync(FULL, dB.x, src), __shfl_sync(FULL, dB.y, src), __shfl_sync(FULL, dB.z, src));
            const V3 r = support_hull_coop<SMEM>(verts + B.voff, B.vcnt, d, lane);
            if ((int)lane == src) sb = r;
          }
        }
      } else {
        sa = support_local<SMEM>(A, dir, verts, img.prims, cells);
        sb = support_local<SMEM>(B, dB, verts, img.prims, cells);
      }
      const V3 v = sa - (rot(R, sb) + t);
      if (!busy) continue;
      iters++;
      bool done = false, hit = false;
      const float dv = dot(v, dir);
      if (phase == 1) {
        v1 = v;
        if (is_zero(dv) || dv < 0.f) done = true;
        else {
          dir = cross(v0, v1);
          // libccd tests |v0 x v1|^2 against DBL_EPSILON (effectively exact zero); in fp32 the test must be
          // relative to the operand lengths or centimetre-sized shapes alias to "origin on the segment"
          if (dot(dir, dir) <= 1e-12f * dot(v0, v0) * dot(v1, v1)) { done = true; hit = true; }
          else { dir = normalized(dir); phase = 2; }
        }
      } else if (phase == 2) {
        v2 = v;
        if (is_zero(dv) || dv < 0.f) done = true;
        else {
          dir = normalized(cross(v1 - v0, v2 - v0));
          if (dot(dir, v0) > 0.f) { V3 tmp = v1; v1 = v2; v2 = tmp; dir = dir * -1.f; }
          phase = 3;
        }
      } else if (phase == 3) {
        v3p = v;
        if (is_zero(dv) || dv < 0.f) done = true;
        else {
          bool cont = false;
          const float d1 = dot(cross(v1, v3p), v0);
          if (d1 < 0.f) { v2 = v3p; cont = true; }  // triple products: exact sign (DBL_EPSILON-scale zero band upstream)
          if (!cont) {
            const float d2 = dot(cross(v3p, v2), v0);
            if (d2 < 0.f) { v1 = v3p; cont = true; }
          }
          if (cont) dir = normalized(cross(v1 - v0, v2 - v0));
          else {
            phase = 4;
            dir = normalized(cross(v2 - v1, v3p - v1));
            const float de = dot(dir, v1);
            if (is_zero(de) || de > 0.f) { done = true; hit = true; }
          }
        }
      } else {
        // refinePortal: can v4 expand the portal past the origin, and is there still progress?
        if (!(is_zero(dv) || dv > 0.f)) done = true;
        else {
          const float m = fminf(dv - dot(v1, dir), fminf(dv - dot(v2, dir), dv - dot(v3p, dir)));
          if (m <= MPB_MPR_TOL) done = true;
          else {
            const V3 v4v0 = cross(v, v0);
            if (dot(v1, v4v0) > 0.f) {
              if (dot(v2, v4v0) > 0.f) v1 = v; else v3p = v;
            } else {
              if (dot(v3p, v4v0) > 0.f) v2 = v; else v1 = v;
            }
            dir = normalized(cross(v2 - v1, v3p - v1));
            const float de = dot(dir, v1);
            if (is_zero(de) || de > 0.f) { done = true; hit = true; }
          }
        }
      }
      if (!done && iters >= MPB_MPR_MAX_ITERS) { done = true; cap_sum++; }
      if (done) {
        busy = false;
        it_sum += (unsigned)iters;
        if (hit) report_hit(out, (unsigned)cfg, p);
      }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    it_sum += __shfl_xor_sync(FULL, it_sum, o);
    cap_sum += __shfl_xor_sync(FULL, cap_sum, o);
  }
  if (lane == 0 && it_sum) atomicAdd(&wq.counters[1], it_sum);
  if (lane == 0 && cap_sum) atomicAdd(&wq.counters[2], cap_sum);
}

// ---------------------------------------------------------------------------------------------- narrowphase, phase-split
// narrow_kernel runs the four MPR phases as one per-lane state machine; ncu showed 15 of 32 lanes active per instruction
// (profiles/r02_ncu_full_c3_pre_expand_lane_raw.csv): the phase bodies serialise, and the long item-fetch block runs for
// the few lanes that just finished.  Here a warp alternates between two CONVERGED regions over its slice of a bin:
//   A  prologue: 32 fresh items, one per lane -- fetch, v0, the first two supports (v1, v2) and their exits.  More
//      than half of the items that do not collide end here; what is left has a triangle (v0, v1, v2) to search from.
//   B  portal loop over the survivors of A (kept in a per-warp buffer in shared memory: 13 floats + item index):
//      discoverPortal's v3 search and refinePortal, one support per trip, lanes refilled from the buffer.
// The arithmetic is that of narrow_kernel (libccd's mpr.c order); only the schedule differs.
#define MPB_SURV_CAP 64
#define MPB_SURV_F 13  // v0 v1 v2 dir (12 floats) + iteration count

struct MprItem {  // what a lane needs of its item for a support evaluation
  float R[9];
  V3 t;
  int cfg, p;
};

template <bool SMEM>
__device__ __forceinline__ V3 mink_support(const DevShape &A, const DevShape &B, const MprItem &I, V3 dir, const float4 *verts,
                                           const float4 *prims, const unsigned char *cells) {
  const V3 sa = support_local<SMEM>(A, dir, verts, prims, cells);
  const V3 sb = support_local<SMEM>(B, rot_t(I.R, dir * -1.f), verts, prims, cells);
  return sa - (rot(I.R, sb) + I.t);
}

// item `my` of the binned queue -> lane state; false: nothing to do (already known to collide)
__device__ __forceinline__ bool mpr_fetch(const SceneImg &img, const WorkQ &wq, const HitOut &out, unsigned my, int round, DevShape &A, DevShape &B,
                                          MprItem &I, V3 &cA, V3 &cBl, int2 *objs = nullptr) {
  const float4 *rc = item_rec(wq, round, my);
  float t3[3];
  int aux;
  item_pose(rc, I.R, t3);
  item_ids(rc, I.cfg, I.p, aux);
  I.t = v3(t3[0], t3[1], t3[2]);
  const DevPair pr = img.pairs[I.p];
  if (objs) *objs = make_int2(pr.a, pr.b);
  const DevObject &OA = img.objs[pr.a];
  const DevObject &OB = img.objs[pr.b];
  A.p0 = OA.prm[0]; A.p1 = OA.prm[1]; A.p2 = OA.prm[2];  // primitive parameters differ per item
  B.p0 = OB.prm[0]; B.p1 = OB.prm[1]; B.p2 = OB.prm[2];
  cA = v3(OA.ctr[0], OA.ctr[1], OA.ctr[2]);
  cBl = v3(OB.ctr[0], OB.ctr[1], OB.ctr[2]);
  if (round == 1) {  // [upstream] centerTriangle = centroid; voxel Box centre = its pose
    B.voff = aux;
    const float4 a = __ldg(img.prims + 3 * B.voff);
    if (B.type == MPB_GEOM_TRI) {
      const float4 b = __ldg(img.prims + 3 * B.voff + 1), c = __ldg(img.prims + 3 * B.voff + 2);
      cBl = v3((a.x + b.x + c.x) * (1.f / 3.f), (a.y + b.y + c.y) * (1.f / 3.f), (a.z + b.z + c.z) * (1.f / 3.f));
    } else cBl = v3(a.x, a.y, a.z);
    return !hit_known(out, (unsigned)I.cfg, I.p);  // early exit: this (configuration, pair) is already known to collide
  }
  return true;
}

template <bool SMEM>
__global__ void __launch_bounds__(256, MPB_NARROW_MINB) narrow2_kernel(SceneImg img, WorkQ wq, HitOut out, int n_bins, int round, int refill_min, int presolve) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ __align__(8) unsigned long long bar;
  const float4 *verts = img.verts;
  const unsigned pool = ((unsigned)img.pool_bytes + 127u) & ~127u;
  if (SMEM) {
    const unsigned total = (unsigned)img.pool_bytes;
    if (threadIdx.x == 0) {
      mbar_init(&bar, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0 && total > 0) {
      mbar_expect_tx(&bar, total);
      for (unsigned off = 0; off < total; off += 32768u) tma_bulk_g2s(smem_raw + off, reinterpret_cast<const unsigned char *>(img.verts) + off, min(32768u, total - off), &bar);
    }
    if (total > 0) mbar_wait(&bar, 0);
    verts = reinterpret_cast<const float4 *>(smem_raw);
  }
  const unsigned char *cells = reinterpret_cast<const unsigned char *>(verts + img.n_verts);
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // survivor buffer of this warp: [field][slot] floats, then [slot] item indices
  float *sv = reinterpret_cast<float *>(smem_raw + (SMEM ? pool : 0u)) + (size_t)warp * (MPB_SURV_F + 1) * MPB_SURV_CAP;
  unsigned *sv_item = reinterpret_cast<unsigned *>(sv + MPB_SURV_F * MPB_SURV_CAP);
  const unsigned n_units = wq.totals[1], unit_items = wq.totals[4];
  unsigned long long it_sum = 0, cap_sum = 0;
  for (;;) {
    unsigned u = 0;
    if (lane == 0) u = atomicAdd(&wq.totals[3], 1u);
    u = __shfl_sync(FULL, u, 0);
    if (u >= n_units) break;
    int lo = 0, hi = n_bins;  // bin of this unit: largest b with unit_start[b] <= u
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (wq.unit_start[mid] <= u) lo = mid; else hi = mid;
    }
    unsigned next = wq.bin_start[lo] + (u - wq.unit_start[lo]) * unit_items;
    const unsigned end = min(min(wq.bin_start[lo] + wq.bin_cursor[lo], next + unit_items), wq.cap);
    if (next >= end) continue;
    DevShape A, B;
    {  // shapes are uniform over the bin: read them once from the first item of the slice
      int c0_, p0_, a0_;
      item_ids(item_rec(wq, round, next), c0_, p0_, a0_);
      const DevPair pr0 = img.pairs[p0_];
      A = shape_of(img.objs[pr0.a]);
      B = shape_of(img.objs[pr0.b]);
      if (round == 1) B.type = img.objs[pr0.b].vcnt;  // leaf primitives of a BVH object: MPB_GEOM_TRI / MPB_GEOM_VBOX
    }
    int cnt = 0;  // survivors in the buffer (warp-uniform)
    for (;;) {
      // ---------------- region A: prologue batches until the buffer holds a warp's worth of survivors
      while (cnt <= MPB_SURV_CAP - 32 && next < end) {
        const unsigned my = next + lane;
        next += 32;
        MprItem I;
        V3 cA = v3(0.f, 0.f, 0.f), cBl = cA;
        bool act = false, hit = false;
        if (my < end) {
          int2 ob;
          act = mpr_fetch(img, wq, out, my, round, A, B, I, cA, cBl, &ob);
          // sufficient test with the inner balls (exact closed forms against a primitive partner; leaf items: against the voxel
          // box): here the lanes are converged and the geometry types are those of the bin, so the type switch is uniform.
          // (Inside dense_kernel, on unsorted survivors, the same code ran on 2-4 lanes of 32: profiles/r02_ncu_dense_presolve_source.csv.)
          if (act && presolve && sufficient_hit(img, img.objs[ob.x], img.objs[ob.y], I.R, I.t, round == 1 ? B.voff : -1)) { act = false; hit = true; }
        } else {
#pragma unroll
          for (int k = 0; k < 9; k++) I.R[k] = (k % 4 == 0) ? 1.f : 0.f;
          I.t = cA; I.cfg = 0; I.p = 0;
          if (round == 1) B.voff = 0;
        }
        V3 v0 = cA - (rot(I.R, cBl) + I.t);
        if (is_zero(v0.x) && is_zero(v0.y) && is_zero(v0.z)) v0.x += 10.f * MPB_EPS;
        V3 dir = normalized(v0 * -1.f);
        int iters = 0;
        const V3 v1 = mink_support<SMEM>(A, B, I, dir, verts, img.prims, cells);
        V3 v2 = v1;
        if (act) {
          iters = 1;
          const float dv = dot(v1, dir);
          if (is_zero(dv) || dv < 0.f) act = false;
          else {
            dir = cross(v0, v1);
            // libccd tests |v0 x v1|^2 against DBL_EPSILON (effectively exact zero); in fp32 the test must be
            // relative to the operand lengths or centimetre-sized shapes alias to "origin on the segment"
            if (dot(dir, dir) <= 1e-12f * dot(v0, v0) * dot(v1, v1)) { act = false; hit = true; }
            else dir = normalized(dir);
          }
        }
        V3 w1 = v1;
        if (__any_sync(FULL, act)) {
          v2 = mink_support<SMEM>(A, B, I, dir, verts, img.prims, cells);
          if (act) {
            iters = 2;
            const float dv = dot(v2, dir);
            if (is_zero(dv) || dv < 0.f) act = false;
            else {
              dir = normalized(cross(v1 - v0, v2 - v0));
              if (dot(dir, v0) > 0.f) { w1 = v2; v2 = v1; dir = dir * -1.f; }
            }
          }
        }
        if (hit) report_hit(out, (unsigned)I.cfg, I.p);
        if (!act) it_sum += (unsigned)iters;
        const unsigned sm = __ballot_sync(FULL, act);
        if (act) {
          const int slot = cnt + __popc(sm & ((1u << lane) - 1u));
          sv[0 * MPB_SURV_CAP + slot] = v0.x; sv[1 * MPB_SURV_CAP + slot] = v0.y; sv[2 * MPB_SURV_CAP + slot] = v0.z;
          sv[3 * MPB_SURV_CAP + slot] = w1.x; sv[4 * MPB_SURV_CAP + slot] = w1.y; sv[5 * MPB_SURV_CAP + slot] = w1.z;
          sv[6 * MPB_SURV_CAP + slot] = v2.x; sv[7 * MPB_SURV_CAP + slot] = v2.y; sv[8 * MPB_SURV_CAP + slot] = v2.z;
          sv[9 * MPB_SURV_CAP + slot] = dir.x; sv[10 * MPB_SURV_CAP + slot] = dir.y; sv[11 * MPB_SURV_CAP + slot] = dir.z;
          sv[12 * MPB_SURV_CAP + slot] = __int_as_float(iters);
          sv_item[slot] = my;
        }
        cnt += __popc(sm);
        __syncwarp();
      }
      if (cnt == 0) break;  // slice exhausted (the loop above ran out of items) and nothing to refine
      // ---------------- region B: portal search / refinement over the buffered survivors
      bool busy = false;
      MprItem I;
      V3 v0 = v3(0.f, 0.f, 0.f), v1 = v0, v2 = v0, v3p = v0, dir = v3(1.f, 0.f, 0.f);
      int phase = 3, iters = 0;
#pragma unroll
      for (int k = 0; k < 9; k++) I.R[k] = (k % 4 == 0) ? 1.f : 0.f;
      I.t = v0; I.cfg = 0; I.p = 0;
      for (;;) {
        const unsigned idle = __ballot_sync(FULL, !busy);
        if (cnt > 0 && (idle == FULL || __popc(idle) >= refill_min)) {
          const int take = min(cnt, __popc(idle));
          const int rank = __popc(idle & ((1u << lane) - 1u));
          if (!busy && rank < take) {
            const int slot = cnt - 1 - rank;
            v0 = v3(sv[0 * MPB_SURV_CAP + slot], sv[1 * MPB_SURV_CAP + slot], sv[2 * MPB_SURV_CAP + slot]);
            v1 = v3(sv[3 * MPB_SURV_CAP + slot], sv[4 * MPB_SURV_CAP + slot], sv[5 * MPB_SURV_CAP + slot]);
            v2 = v3(sv[6 * MPB_SURV_CAP + slot], sv[7 * MPB_SURV_CAP + slot], sv[8 * MPB_SURV_CAP + slot]);
            dir = v3(sv[9 * MPB_SURV_CAP + slot], sv[10 * MPB_SURV_CAP + slot], sv[11 * MPB_SURV_CAP + slot]);
            iters = __float_as_int(sv[12 * MPB_SURV_CAP + slot]);
            V3 cA, cBl;
            mpr_fetch(img, wq, out, sv_item[slot], round, A, B, I, cA, cBl);
            phase = 3;
            busy = true;
          }
          cnt -= take;
          __syncwarp();
        }
        if (__ballot_sync(FULL, busy) == 0) break;  // buffer empty (or it would have refilled) and every lane finished
        const V3 v = mink_support<SMEM>(A, B, I, dir, verts, img.prims, cells);
        if (!busy) continue;
        iters++;
        bool done = false, hit = false;
        const float dv = dot(v, dir);
        if (phase == 3) {
          v3p = v;
          if (is_zero(dv) || dv < 0.f) done = true;
          else {
            bool cont = false;
            const float d1 = dot(cross(v1, v3p), v0);
            if (d1 < 0.f) { v2 = v3p; cont = true; }  // triple products: exact sign (DBL_EPSILON-scale zero band upstream)
            if (!cont) {
              const float d2 = dot(cross(v3p, v2), v0);
              if (d2 < 0.f) { v1 = v3p; cont = true; }
            }
            if (cont) dir = normalized(cross(v1 - v0, v2 - v0));
            else {
              phase = 4;
              dir = normalized(cross(v2 - v1, v3p - v1));
              const float de = dot(dir, v1);
              if (is_zero(de) || de > 0.f) { done = true; hit = true; }
            }
          }
        } else {
          // refinePortal: can v4 expand the portal past the origin, and is there still progress?
          if (!(is_zero(dv) || dv > 0.f)) done = true;
          else {
            const float m = fminf(dv - dot(v1, dir), fminf(dv - dot(v2, dir), dv - dot(v3p, dir)));
            if (m <= MPB_MPR_TOL) done = true;
            else {
              const V3 v4v0 = cross(v, v0);
              if (dot(v1, v4v0) > 0.f) {
                if (dot(v2, v4v0) > 0.f) v1 = v; else v3p = v;
              } else {
                if (dot(v3p, v4v0) > 0.f) v2 = v; else v1 = v;
              }
              dir = normalized(cross(v2 - v1, v3p - v1));
              const float de = dot(dir, v1);
              if (is_zero(de) || de > 0.f) { done = true; hit = true; }
            }
          }
        }
        if (!done && iters >= MPB_MPR_MAX_ITERS) { done = true; cap_sum++; }
        if (done) {
          busy = false;
          it_sum += (unsigned)iters;
          if (hit) report_hit(out, (unsigned)I.cfg, I.p);
        }
      }
      if (next >= end && cnt == 0) break;
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    it_sum += __shfl_xor_sync(FULL, it_sum, o);
    cap_sum += __shfl_xor_sync(FULL, cap_sum, o);
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

// ---------------------------------------------------------------------------------------------- BVH expansion
// Warp per (configuration, shape x BVH-object) item of the first round.  Walks the object's AABB tree
// (mesh triangles, or octree cells merged into boxes) against the shape's OBB, both in the object's frame
// (replaces the OBBRSS descent of MeshShapeCollisionTraversalNodeOBBRSS and OcTreeSolver::
// OcTreeShapeIntersectRecurse that ::fcl::collide runs for src/planning_world.cpp:419-434, 464-478), and
// emits one leaf-level item per surviving primitive for the second narrowphase round.  A leaf whose closest
// point lies inside the shape's inscribed ball is an immediate hit (no item needed).
#include <cooperative_groups.h>
namespace cg = cooperative_groups;


#define MPB_EXP_STACK 1024  // node frontier per warp

// Warp per item: the frontier of BVH nodes lives in a shared-memory stack; every trip pops up to 32 nodes (one per
// lane), tests them against the shape's OBB, pushes the children of the surviving inner nodes (ballot compaction) and
// lets the lanes that hold surviving leaves test / emit their primitives.  No divergence inside the 15-axis test and
// a short critical path even for small batches (a planning round); when the stack runs nearly full the warp pops
// one node per trip (depth-first), which bounds the growth by the tree depth.
#define MPB_EXPL_MIN_ITEMS 32768u  // items in the bin from which expand_lane_kernel runs instead of expand_kernel
__global__ void __launch_bounds__(128, 8) expand_kernel(SceneImg img, WorkQ wq, HitOut out, int lane_variant_on) {
  __shared__ int s_stack[4][MPB_EXP_STACK];
  const int eb = img.expand_bin;
  if (eb < 0) return;
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int *stack = s_stack[warp];
  const unsigned first = wq.bin_start[eb], end = min(wq.bin_start[eb + 1], wq.cap);
  if (lane_variant_on && end - first >= MPB_EXPL_MIN_ITEMS) return;  // expand_lane_kernel takes large bins
  for (;;) {
    // four items per ticket; lanes 0..3 fetch their (configuration, pair, already-known) in parallel
    unsigned it0 = 0;
    if (lane == 0) it0 = atomicAdd(&wq.totals[5], 4u);
    it0 = first + __shfl_sync(FULL, it0, 0);
    if (it0 >= end) break;
    int my_cfg = 0, my_p = 0;
    bool my_known = true;
    if (lane < 4 && it0 + lane < end) {
      int aux_;
      item_ids(item_rec(wq, 0, it0 + lane), my_cfg, my_p, aux_);
      my_known = hit_known(out, (unsigned)my_cfg, my_p);
    }
    for (int sub = 0; sub < 4; sub++) {
    const unsigned it = it0 + sub;
    const int cfg = __shfl_sync(FULL, my_cfg, sub), p = __shfl_sync(FULL, my_p, sub);
    if (__shfl_sync(FULL, (int)my_known, sub)) continue;  // also covers slots past the end of the bin
    const DevPair pr = img.pairs[p];
    const DevObject &A = img.objs[pr.a];  // the shape
    const DevObject &B = img.objs[pr.b];  // the BVH object
    if (B.voff < 0) continue;             // empty octree
    float R[9], t[3];
    item_pose(item_rec(wq, 0, it), R, t);
    // shape OBB expressed in the object frame B:  x_B = R^T (x_A - t)
    float Ra[9], ca[3];  // OBB axes (columns) and centre in B
    {
      const float d0 = A.oc[0] - t[0], d1 = A.oc[1] - t[1], d2 = A.oc[2] - t[2];
#pragma unroll
      for (int r = 0; r < 3; r++) {
        ca[r] = fmaf(R[r], d0, fmaf(R[3 + r], d1, R[6 + r] * d2));
#pragma unroll
        for (int c = 0; c < 3; c++) Ra[3 * r + c] = fmaf(R[r], A.oR[c], fmaf(R[3 + r], A.oR[3 + c], R[6 + r] * A.oR[6 + c]));
      }
    }
    // node AABB (axes = identity) in the OBB's frame: Rrel = Ra^T
    float Rrel[9];
#pragma unroll
    for (int r = 0; r < 3; r++)
#pragma unroll
      for (int c = 0; c < 3; c++) Rrel[3 * r + c] = Ra[3 * c + r];
    // inscribed-ball centre of the shape in B
    float cin[3];
    {
      const float d0 = A.ctr[0] - t[0], d1 = A.ctr[1] - t[1], d2 = A.ctr[2] - t[2];
#pragma unroll
      for (int r = 0; r < 3; r++) cin[r] = fmaf(R[r], d0, fmaf(R[3 + r], d1, R[6 + r] * d2));
    }
    const float rin2 = A.rin * A.rin;
    const float ah[3] = {A.oh[0], A.oh[1], A.oh[2]};
    // axis-aligned bounds of the shape's OBB in the object frame: the cheap test of the descent (six compares per node);
    // the exact 15-axis test runs per leaf primitive only
    float bh[3];
#pragma unroll
    for (int r = 0; r < 3; r++) bh[r] = fmaf(fabsf(Ra[3 * r]), ah[0], fmaf(fabsf(Ra[3 * r + 1]), ah[1], fabsf(Ra[3 * r + 2]) * ah[2])) + 1e-5f;
    const int leaf_type = B.vcnt;
    const int bin2 = img.pair_bin2[p];
    int sp = 1;
    if (lane == 0) stack[0] = B.voff;
    __syncwarp();
    bool hit = false, dropped = false;
    while (sp > 0) {
      const int take = sp > MPB_EXP_STACK - 64 ? 1 : min(sp, 32);
      int node = -1;
      if ((int)lane < take) node = stack[sp - 1 - (int)lane];
      sp -= take;
      __syncwarp();
      bool pass = false;
      DevBvhNode nd;
      nd.left = 0; nd.right = 0;
      if (node >= 0) {
        nd = img.nodes[node];
        pass = fabsf(nd.c[0] - ca[0]) <= nd.h[0] + bh[0] && fabsf(nd.c[1] - ca[1]) <= nd.h[1] + bh[1] &&
               fabsf(nd.c[2] - ca[2]) <= nd.h[2] + bh[2];
      }
      const bool inner = pass && nd.left >= 0;
      const unsigned im = __ballot_sync(FULL, inner);
      if (inner) {
        const int pos = sp + 2 * __popc(im & ((1u << lane) - 1u));
        if (pos + 1 < MPB_EXP_STACK) { stack[pos] = nd.left; stack[pos + 1] = nd.right; }
        else { atomicAdd(&wq.counters[2], 1ull); dropped = true; }  // dropped subtree: counted, and the item reads as colliding
      }
      sp = min(sp + 2 * __popc(im), MPB_EXP_STACK);  // (a frontier beyond the stack would need a tree deeper than ~60)
      __syncwarp();
      bool lhit = false;
      if (pass && nd.left < 0) {
        const int pfirst = ~nd.left;
        for (int k = 0; k < nd.right && !lhit; k++) {
          const int prim = pfirst + k;
          const float4 a = __ldg(img.prims + 3 * prim), b = __ldg(img.prims + 3 * prim + 1);
          float pc[3], ph[3];
          if (leaf_type == MPB_GEOM_VBOX) {
            pc[0] = a.x; pc[1] = a.y; pc[2] = a.z; ph[0] = b.x; ph[1] = b.y; ph[2] = b.z;
          } else {
            const float4 c = __ldg(img.prims + 3 * prim + 2);
            const float lo0 = fminf(a.x, fminf(b.x, c.x)), hi0 = fmaxf(a.x, fmaxf(b.x, c.x));
            const float lo1 = fminf(a.y, fminf(b.y, c.y)), hi1 = fmaxf(a.y, fmaxf(b.y, c.y));
            const float lo2 = fminf(a.z, fminf(b.z, c.z)), hi2 = fmaxf(a.z, fmaxf(b.z, c.z));
            pc[0] = 0.5f * (lo0 + hi0); pc[1] = 0.5f * (lo1 + hi1); pc[2] = 0.5f * (lo2 + hi2);
            ph[0] = 0.5f * (hi0 - lo0) + 1e-6f; ph[1] = 0.5f * (hi1 - lo1) + 1e-6f; ph[2] = 0.5f * (hi2 - lo2) + 1e-6f;
          }
          {  // primitive-level reject: the exact OBB test
            const float g0 = pc[0] - ca[0], g1 = pc[1] - ca[1], g2 = pc[2] - ca[2];
            float tr[3];
#pragma unroll
            for (int r = 0; r < 3; r++) tr[r] = fmaf(Ra[r], g0, fmaf(Ra[3 + r], g1, Ra[6 + r] * g2));
            if (obb_disjoint(Rrel, tr, ah, ph)) continue;
          }
          if (leaf_type == MPB_GEOM_VBOX && rin2 > 0.f) {  // sufficient test: cell touches the inscribed ball
            const float q0 = clampf(cin[0], pc[0] - ph[0], pc[0] + ph[0]) - cin[0];
            const float q1 = clampf(cin[1], pc[1] - ph[1], pc[1] + ph[1]) - cin[1];
            const float q2 = clampf(cin[2], pc[2] - ph[2], pc[2] + ph[2]) - cin[2];
            if (fmaf(q0, q0, fmaf(q1, q1, q2 * q2)) < rin2) { lhit = true; break; }
          }
          // emit a leaf-level item (warp-aggregated slot allocation)
          cg::coalesced_group g = cg::coalesced_threads();
          unsigned base = 0;
          if (g.thread_rank() == 0) base = atomicAdd(&wq.totals[0], (unsigned)g.size());
          base = g.shfl(base, 0);
          const unsigned slot = base + g.thread_rank();
          if (slot < wq.cap) {
            item_store(wq, 1, slot, R, t, cfg, p, prim, bin2);
            atomicAdd(&wq.bin_count[bin2], 1u);
          }
        }
      }
      if (__any_sync(FULL, lhit || dropped)) { hit = true; break; }  // a subtree that could not be held: conservative verdict
    }
    if (hit && lane == 0) report_hit(out, (unsigned)cfg, p);
    __syncwarp();
    }  // sub-items of the ticket
  }
}

// Lane per item: the throughput variant of expand_kernel for passes with many items (a collide call over 10^5+
// configurations).  A lane walks its own item's tree depth-first with a private stack in shared memory ([depth][thread],
// conflict-free), one sibling pair per trip: the stack holds the first-child index of inner nodes whose bounds already
// passed; siblings are adjacent in the pool (host re-layout in mpb_scene_create), so one dependent 64-byte access
// brings both children's boxes (the first version popped a node, read it, then read its children: two dependent accesses
// per trip, and leaves cost a trip of their own).  Passing inner children are pushed, the one nearer to the shape's
// inscribed ball on top, so an item that collides meets its decisive leaves first; passing leaf children hand their
// primitives over in the same trip.  LEAVES are not tested by the lane that finds them: their
// primitives go to a per-warp queue and are tested 32 at a time by the whole warp (exact OBB test, inscribed-ball
// shortcut, emission of the leaf item), each lane reading the owner's item record from shared memory.  The first version
// tested leaves in place and ran its leaf code -- 60 % of its instructions -- on two lanes of 32
// (profiles/r02_ncu_expand_lane_v1_source.csv).  Lanes are persistent: idle lanes take the next items of the bin.
// The warp-per-item kernel stays for small passes (planning rounds), where the critical path of a single traversal
// matters more than throughput; both are launched, the item count of the bin decides which one runs.
#define MPB_EXPL_DEPTH 28     // per-lane stack entries: the walk holds at most one sibling pair per level (tree depth ~17 for 10^5 leaves; 28 covers 2^28)
#ifndef MPB_EXPL_REFILL
#define MPB_EXPL_REFILL 16    // idle lanes that trigger a refill (the refill block is long): 4: +0.65 ms per C3 step, 8: 0, 16: -0.18, 24: +0.05, 32: +4
#endif
#define MPB_EXPL_QCAP 128     // leaf-primitive queue entries per warp
#define MPB_EXPL_BINS 64      // second-round bins counted in shared memory
#define MPB_EXPL_F 31         // floats of an item record: R 9, t 3, Ra 9, ca 3, cin 3, ah 3, rin2

struct ExpShared {
  int stack[MPB_EXPL_DEPTH][128];
  float item[MPB_EXPL_F][128];  // [field][thread]: lanes of the warp read each other's record in the leaf phase
  int meta[4][128];             // cfg, pair, leaf type, second-round bin
  int hitflag[128];             // set by the leaf phase: this lane's item collides, stop its traversal
  uint2 queue[4][MPB_EXPL_QCAP];  // per warp: (owner thread, primitive)
  unsigned bins[MPB_EXPL_BINS];
};

// squared distance from point c to the box (centre, half extents) of a node
__device__ __forceinline__ float exp_node_dist2(const DevBvhNode &nd, const float *c) {
  const float d0 = fmaxf(fabsf(nd.c[0] - c[0]) - nd.h[0], 0.f), d1 = fmaxf(fabsf(nd.c[1] - c[1]) - nd.h[1], 0.f), d2 = fmaxf(fabsf(nd.c[2] - c[2]) - nd.h[2], 0.f);
  return fmaf(d0, d0, fmaf(d1, d1, d2 * d2));
}
__device__ __forceinline__ bool exp_box_pass(const DevBvhNode &nd, const float *ca, const float *bh) {
  return fabsf(nd.c[0] - ca[0]) <= nd.h[0] + bh[0] && fabsf(nd.c[1] - ca[1]) <= nd.h[1] + bh[1] && fabsf(nd.c[2] - ca[2]) <= nd.h[2] + bh[2];
}

// Warp-cooperative leaf phase: entries [q0, q0 + nb) of the warp's queue, one per lane.
__device__ __forceinline__ void exp_leaf_batch(const SceneImg &img, const WorkQ &wq, const HitOut &out, ExpShared &S, const uint2 *q, int nb,
                                               unsigned lane, bool smem_bins) {
  const unsigned FULL = 0xffffffffu;
  bool emit = false, hit = false;
  int o = 0, prim = 0;
  if ((int)lane < nb) {
    const uint2 e = q[lane];
    o = (int)e.x;
    prim = (int)e.y;
    if (!S.hitflag[o]) {
      const int leaf_type = S.meta[2][o];
      const float4 a = __ldg(img.prims + 3 * prim), b = __ldg(img.prims + 3 * prim + 1);
      float pc[3], ph[3];
      if (leaf_type == MPB_GEOM_VBOX) {
        pc[0] = a.x; pc[1] = a.y; pc[2] = a.z; ph[0] = b.x; ph[1] = b.y; ph[2] = b.z;
      } else {
        const float4 c = __ldg(img.prims + 3 * prim + 2);
        const float lo0 = fminf(a.x, fminf(b.x, c.x)), hi0 = fmaxf(a.x, fmaxf(b.x, c.x));
        const float lo1 = fminf(a.y, fminf(b.y, c.y)), hi1 = fmaxf(a.y, fmaxf(b.y, c.y));
        const float lo2 = fminf(a.z, fminf(b.z, c.z)), hi2 = fmaxf(a.z, fmaxf(b.z, c.z));
        pc[0] = 0.5f * (lo0 + hi0); pc[1] = 0.5f * (lo1 + hi1); pc[2] = 0.5f * (lo2 + hi2);
        ph[0] = 0.5f * (hi0 - lo0) + 1e-6f; ph[1] = 0.5f * (hi1 - lo1) + 1e-6f; ph[2] = 0.5f * (hi2 - lo2) + 1e-6f;
      }
      float Ra[9], ah[3], tr[3], Rrel[9];
#pragma unroll
      for (int k = 0; k < 9; k++) Ra[k] = S.item[12 + k][o];
#pragma unroll
      for (int k = 0; k < 3; k++) ah[k] = S.item[27 + k][o];
      const float g0 = pc[0] - S.item[21][o], g1 = pc[1] - S.item[22][o], g2 = pc[2] - S.item[23][o];
#pragma unroll
      for (int r = 0; r < 3; r++) {
        tr[r] = fmaf(Ra[r], g0, fmaf(Ra[3 + r], g1, Ra[6 + r] * g2));
#pragma unroll
        for (int c = 0; c < 3; c++) Rrel[3 * r + c] = Ra[3 * c + r];
      }
      if (!obb_disjoint(Rrel, tr, ah, ph)) {  // primitive-level reject: the exact OBB test
        emit = true;
        const float rin2 = S.item[30][o];
        if (leaf_type == MPB_GEOM_VBOX && rin2 > 0.f) {  // sufficient test: cell touches the inscribed ball
          const float c0 = S.item[24][o], c1 = S.item[25][o], c2 = S.item[26][o];
          const float q0 = clampf(c0, pc[0] - ph[0], pc[0] + ph[0]) - c0;
          const float q1 = clampf(c1, pc[1] - ph[1], pc[1] + ph[1]) - c1;
          const float q2 = clampf(c2, pc[2] - ph[2], pc[2] + ph[2]) - c2;
          if (fmaf(q0, q0, fmaf(q1, q1, q2 * q2)) < rin2) { hit = true; emit = false; }
        }
      }
    }
  }
  if (hit) {
    S.hitflag[o] = 1;
    report_hit(out, (unsigned)S.meta[0][o], S.meta[1][o]);
  }
  const unsigned em = __ballot_sync(FULL, emit);
  if (em) {
    unsigned base = 0;
    if (lane == 0) base = atomicAdd(&wq.totals[0], (unsigned)__popc(em));
    base = __shfl_sync(FULL, base, 0);
    if (emit) {
      const unsigned slot = base + __popc(em & ((1u << lane) - 1u));
      if (slot < wq.cap) {
        float R9[9], t3[3];
#pragma unroll
        for (int k2 = 0; k2 < 9; k2++) R9[k2] = S.item[k2][o];
#pragma unroll
        for (int k2 = 0; k2 < 3; k2++) t3[k2] = S.item[9 + k2][o];
        const int bin2 = S.meta[3][o];
        item_store(wq, 1, slot, R9, t3, S.meta[0][o], S.meta[1][o], prim, bin2);
        if (smem_bins) atomicAdd(&S.bins[bin2], 1u);
        else atomicAdd(&wq.bin_count[bin2], 1u);
      }
    }
  }
  __syncwarp();
}

#ifndef MPB_EXPL_MINB
#define MPB_EXPL_MINB 1
#endif
__global__ void __launch_bounds__(128, MPB_EXPL_MINB) expand_lane_kernel(SceneImg img, WorkQ wq, HitOut out) {
  __shared__ ExpShared S;
  const int eb = img.expand_bin;
  if (eb < 0) return;
  const unsigned first = wq.bin_start[eb], end = min(wq.bin_start[eb + 1], wq.cap);
  if (end - first < MPB_EXPL_MIN_ITEMS) return;  // expand_kernel (warp per item) handles small bins
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tid = threadIdx.x;
  const bool smem_bins = img.n_bins2 <= MPB_EXPL_BINS;
  for (int b = tid; b < MPB_EXPL_BINS; b += blockDim.x) S.bins[b] = 0;
  S.hitflag[tid] = 0;
  __syncthreads();
  int *stack = &S.stack[0][tid];  // entry d at stack[d * 128]
  uint2 *queue = S.queue[warp];
  unsigned long long drop_sum = 0;
  bool busy = false, drained = false;  // drained: the bin has no more items (warp-uniform)
  int sp = 0, qn = 0;                  // qn: entries in the warp's leaf queue (warp-uniform)
  float ca[3] = {0.f, 0.f, 0.f}, bh[3] = {0.f, 0.f, 0.f}, cin[3] = {0.f, 0.f, 0.f};
  for (;;) {
    if (busy && S.hitflag[tid]) { busy = false; sp = 0; }  // a leaf of this lane's item was an immediate hit
    int n_leaf = 0, leaf_first = 0;
    // ---- refill: idle lanes take the next items of the bin (one ticket per warp and refill; the refill block is long, so
    // it runs only when a quarter of the warp is idle, or nobody has work)
    const unsigned need = __ballot_sync(FULL, !busy);
    if (!drained && (__popc(need) >= MPB_EXPL_REFILL || need == FULL)) {
      // queued leaves refer to the records of their owner lanes: test them before any record is replaced
      while (qn > 0) { const int nb = min(qn, 32); exp_leaf_batch(img, wq, out, S, queue + qn - nb, nb, lane, smem_bins); qn -= nb; }
      unsigned base = 0;
      if (lane == 0) base = atomicAdd(&wq.totals[5], (unsigned)__popc(need));
      base = first + __shfl_sync(FULL, base, 0);
      if (base + __popc(need) >= end) drained = true;
      const unsigned my = base + __popc(need & ((1u << lane) - 1u));
      if (!busy && my < end) {
        const float4 *rc = item_rec(wq, 0, my);
        int cfg, p, aux_;
        item_ids(rc, cfg, p, aux_);
        const DevPair pr = img.pairs[p];
        const DevObject &A = img.objs[pr.a];  // the shape
        const DevObject &B = img.objs[pr.b];  // the BVH object
        if (B.voff >= 0 && !hit_known(out, (unsigned)cfg, p)) {
          float R[9], t[3], Ra[9];
          item_pose(rc, R, t);
          {  // shape OBB in the object frame B:  x_B = R^T (x_A - t)
            const float d0 = A.oc[0] - t[0], d1 = A.oc[1] - t[1], d2 = A.oc[2] - t[2];
#pragma unroll
            for (int r = 0; r < 3; r++) {
              ca[r] = fmaf(R[r], d0, fmaf(R[3 + r], d1, R[6 + r] * d2));
#pragma unroll
              for (int c = 0; c < 3; c++) Ra[3 * r + c] = fmaf(R[r], A.oR[c], fmaf(R[3 + r], A.oR[3 + c], R[6 + r] * A.oR[6 + c]));
            }
            const float e0 = A.ctr[0] - t[0], e1 = A.ctr[1] - t[1], e2 = A.ctr[2] - t[2];
#pragma unroll
            for (int r = 0; r < 3; r++) cin[r] = fmaf(R[r], e0, fmaf(R[3 + r], e1, R[6 + r] * e2));
          }
#pragma unroll
          for (int r = 0; r < 3; r++)
            bh[r] = fmaf(fabsf(Ra[3 * r]), A.oh[0], fmaf(fabsf(Ra[3 * r + 1]), A.oh[1], fabsf(Ra[3 * r + 2]) * A.oh[2])) + 1e-5f;
          const DevBvhNode root = img.nodes[B.voff];
          if (exp_box_pass(root, ca, bh)) {
#pragma unroll
            for (int k = 0; k < 9; k++) { S.item[k][tid] = R[k]; S.item[12 + k][tid] = Ra[k]; }
#pragma unroll
            for (int k = 0; k < 3; k++) { S.item[9 + k][tid] = t[k]; S.item[21 + k][tid] = ca[k]; S.item[24 + k][tid] = cin[k]; S.item[27 + k][tid] = A.oh[k]; }
            S.item[30][tid] = A.rin * A.rin;
            S.meta[0][tid] = cfg; S.meta[1][tid] = p; S.meta[2][tid] = B.vcnt; S.meta[3][tid] = img.pair_bin2[p];
            S.hitflag[tid] = 0;
            if (root.left >= 0) {
              stack[0] = root.left;  // the stack holds sibling pairs: index of the first child of a passing inner node
              sp = 1;
              busy = true;
            } else {  // one-leaf tree: its primitives go to the queue in this trip; the record stays until the next refill
              leaf_first = ~root.left;
              n_leaf = root.right;
            }
          }
        }
      }
      __syncwarp();
    }
    if (__ballot_sync(FULL, busy || n_leaf > 0) == 0) {
      while (qn > 0) { const int nb = min(qn, 32); exp_leaf_batch(img, wq, out, S, queue + qn - nb, nb, lane, smem_bins); qn -= nb; }
      if (drained) break;
      continue;
    }
    // ---- one sibling pair of every busy lane's traversal (siblings are adjacent: one 64-byte access reads both boxes)
    if (busy) {
      const int pair = stack[(sp - 1) * 128];
      sp--;
      const DevBvhNode nl = img.nodes[pair], nr = img.nodes[pair + 1];
      const bool pl = exp_box_pass(nl, ca, bh), prt = exp_box_pass(nr, ca, bh);
      const bool il = pl && nl.left >= 0, ir = prt && nr.left >= 0;   // inner children to descend into
      const bool ll = pl && nl.left < 0, lr = prt && nr.left < 0;     // leaf children: primitives to the queue
      if (sp + 2 > MPB_EXPL_DEPTH) {  // cannot happen for a tree of depth < MPB_EXPL_DEPTH; conservative if it does
        drop_sum++;
        S.hitflag[tid] = 1;
        report_hit(out, (unsigned)S.meta[0][tid], S.meta[1][tid]);
      } else {
        const bool left_first = exp_node_dist2(nl, cin) <= exp_node_dist2(nr, cin);
        const int c_far = left_first ? nr.left : nl.left, c_near = left_first ? nl.left : nr.left;
        const bool p_far = left_first ? ir : il, p_near = left_first ? il : ir;
        if (p_far) { stack[sp * 128] = c_far; sp++; }
        if (p_near) { stack[sp * 128] = c_near; sp++; }
        // sibling leaves cover adjacent primitive ranges (build_bvh splits [first, mid) | [mid, end)): one range when both pass
        if (ll) { leaf_first = ~nl.left; n_leaf = nl.right; }
        if (lr) {
          if (!ll) leaf_first = ~nr.left;
          n_leaf += nr.right;
        }
      }
      if (sp == 0) busy = false;  // its queued leaves are tested before the lane's record is replaced
    }
    // ---- leaf primitives -> the warp's queue (prefix sum of the per-lane counts)
    unsigned incl = (unsigned)n_leaf;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned tt = __shfl_up_sync(FULL, incl, o);
      if ((int)lane >= o) incl += tt;
    }
    const int tot = (int)__shfl_sync(FULL, incl, 31);
    if (tot) {
      if (qn + tot > MPB_EXPL_QCAP) {  // make room first (a trip adds at most 32 lanes x 2 leaves x 2 primitives = MPB_EXPL_QCAP entries)
        while (qn > 0) { const int nb = min(qn, 32); exp_leaf_batch(img, wq, out, S, queue + qn - nb, nb, lane, smem_bins); qn -= nb; }
      }
      int pos = qn + (int)incl - n_leaf;
      for (int k = 0; k < n_leaf; k++)
        if (pos + k < MPB_EXPL_QCAP) queue[pos + k] = make_uint2((unsigned)tid, (unsigned)(leaf_first + k));
      qn = min(qn + tot, MPB_EXPL_QCAP);
      __syncwarp();
      while (qn >= 32) { exp_leaf_batch(img, wq, out, S, queue + qn - 32, 32, lane, smem_bins); qn -= 32; }
    }
  }
  if (drop_sum) atomicAdd(&wq.counters[2], drop_sum);
  __syncthreads();
  if (smem_bins)
    for (int b = tid; b < img.n_bins2; b += blockDim.x)
      if (S.bins[b]) atomicAdd(&wq.bin_count[b], S.bins[b]);
}

// ---------------------------------------------------------------------------------------------- BVH x BVH
// Mesh x mesh (BVHModel links of a `convex=False` robot, mesh obstacles) and mesh x octree pairs: replaces
// MeshCollisionTraversalNodeOBBRSS (recursive OBB-OBB descent, triangle-triangle leaves) and
// OcTreeSolver::OcTreeMeshIntersectRecurse (box-triangle leaves) behind ::fcl::collide.  Warp per item; the
// frontier of (node of A, node of B) pairs lives in a shared-memory stack, 32 pairs per trip.  The leaf tests are
// exact closed forms evaluated in place (no second narrowphase round): the verdict is "some primitive pair
// intersects", which is what FCL returns with num_max_contacts = 1.
#define MPB_EXP2_STACK 768

// triangle x triangle [upstream FCL Intersect::intersect_Triangle = PQP TriContact]: 17-axis separating-axis test (two
// normals, nine edge x edge products, six in-plane edge normals, which decide coplanar triangles); touching = hit
__device__ __forceinline__ bool tri_axis_overlap(V3 ax, V3 p2, V3 p3, V3 q1, V3 q2, V3 q3) {  // p1 is the origin
  const float P2 = dot(ax, p2), P3 = dot(ax, p3), Q1 = dot(ax, q1), Q2 = dot(ax, q2), Q3 = dot(ax, q3);
  const float mx1 = fmaxf(0.f, fmaxf(P2, P3)), mn1 = fminf(0.f, fminf(P2, P3)), mx2 = fmaxf(Q1, fmaxf(Q2, Q3)), mn2 = fminf(Q1, fminf(Q2, Q3));
  return !(mn1 > mx2 || mn2 > mx1);
}
__device__ __forceinline__ bool tri_tri_hit(const V3 *A, const V3 *B) {
  const V3 p2 = A[1] - A[0], p3 = A[2] - A[0], q1 = B[0] - A[0], q2 = B[1] - A[0], q3 = B[2] - A[0];
  const V3 e[3] = {p2, p3 - p2, p3 * -1.f}, f[3] = {q2 - q1, q3 - q2, q1 - q3};
  const V3 n1 = cross(e[0], e[1]), m1 = cross(f[0], f[1]);
  if (!tri_axis_overlap(n1, p2, p3, q1, q2, q3)) return false;
  if (!tri_axis_overlap(m1, p2, p3, q1, q2, q3)) return false;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++)
      if (!tri_axis_overlap(cross(e[i], f[j]), p2, p3, q1, q2, q3)) return false;
#pragma unroll
  for (int i = 0; i < 3; i++)
    if (!tri_axis_overlap(cross(e[i], n1), p2, p3, q1, q2, q3)) return false;
#pragma unroll
  for (int j = 0; j < 3; j++)
    if (!tri_axis_overlap(cross(f[j], m1), p2, p3, q1, q2, q3)) return false;
  return true;
}
// axis-aligned box (centre c, half h) vs triangle, 13-axis SAT (closed sets touch = hit)
__device__ __forceinline__ bool box_tri_hit(V3 c, V3 h, V3 p0, V3 p1, V3 p2) {
  const V3 v0 = p0 - c, v1 = p1 - c, v2 = p2 - c;
  const V3 e0 = v1 - v0, e1 = v2 - v1, e2 = v0 - v2;
  if (fminf(v0.x, fminf(v1.x, v2.x)) > h.x || fmaxf(v0.x, fmaxf(v1.x, v2.x)) < -h.x) return false;
  if (fminf(v0.y, fminf(v1.y, v2.y)) > h.y || fmaxf(v0.y, fmaxf(v1.y, v2.y)) < -h.y) return false;
  if (fminf(v0.z, fminf(v1.z, v2.z)) > h.z || fmaxf(v0.z, fmaxf(v1.z, v2.z)) < -h.z) return false;
  const V3 n = cross(e0, e1);
  if (fabsf(dot(n, v0)) > h.x * fabsf(n.x) + h.y * fabsf(n.y) + h.z * fabsf(n.z)) return false;
  const V3 es[3] = {e0, e1, e2};
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const V3 e = es[i];
    // axes (1,0,0) x e, (0,1,0) x e, (0,0,1) x e
    const V3 ax[3] = {v3(0.f, -e.z, e.y), v3(e.z, 0.f, -e.x), v3(-e.y, e.x, 0.f)};
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const float q0 = dot(ax[k], v0), q1 = dot(ax[k], v1), q2 = dot(ax[k], v2);
      const float r = h.x * fabsf(ax[k].x) + h.y * fabsf(ax[k].y) + h.z * fabsf(ax[k].z);
      if (fminf(q0, fminf(q1, q2)) > r || fmaxf(q0, fmaxf(q1, q2)) < -r) return false;
    }
  }
  return true;
}

__global__ void __launch_bounds__(128) expand2_kernel(SceneImg img, WorkQ wq, HitOut out) {
  __shared__ int2 s_stack[4][MPB_EXP2_STACK];
  const int eb = img.expand2_bin;
  if (eb < 0) return;
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int2 *stack = s_stack[warp];
  const unsigned first = wq.bin_start[eb], end = min(wq.bin_start[eb + 1], wq.cap);
  for (;;) {
    unsigned it = 0;
    if (lane == 0) it = atomicAdd(&wq.totals[7], 1u);
    it = first + __shfl_sync(FULL, it, 0);
    if (it >= end) break;
    const float4 *rc = item_rec(wq, 0, it);
    int cfg, p, aux_;
    item_ids(rc, cfg, p, aux_);
    if (hit_known(out, (unsigned)cfg, p)) continue;
    const DevPair pr = img.pairs[p];
    const DevObject &A = img.objs[pr.a];  // mesh
    const DevObject &B = img.objs[pr.b];  // mesh or octree
    if (B.voff < 0) continue;             // empty octree
    float R[9], t[3];  // B in A
    item_pose(rc, R, t);
    const int leaf_b = B.vcnt;  // MPB_GEOM_TRI or MPB_GEOM_VBOX (A's leaves are triangles)
    int sp = 1;
    if (lane == 0) stack[0] = make_int2(A.voff, B.voff);
    __syncwarp();
    bool hit = false, dropped = false;
    while (sp > 0) {
      const int take = sp > MPB_EXP2_STACK - 64 ? 1 : min(sp, 32);
      int2 np = make_int2(-1, -1);
      if ((int)lane < take) np = stack[sp - 1 - (int)lane];
      sp -= take;
      __syncwarp();
      bool pass = false;
      DevBvhNode na, nb;
      na.left = nb.left = 0; na.right = nb.right = 0;
      if (np.x >= 0) {
        na = img.nodes[np.x];
        nb = img.nodes[np.y];
        float tt[3];
#pragma unroll
        for (int r = 0; r < 3; r++) tt[r] = fmaf(R[3 * r], nb.c[0], fmaf(R[3 * r + 1], nb.c[1], fmaf(R[3 * r + 2], nb.c[2], t[r]))) - na.c[r];
        pass = !obb_disjoint(R, tt, na.h, nb.h);
      }
      const bool both_leaf = na.left < 0 && nb.left < 0;
      const bool inner = pass && !both_leaf;
      const unsigned im = __ballot_sync(FULL, inner);
      if (inner) {
        // descend the inner node (the larger one when both are inner)
        const float sa = na.h[0] + na.h[1] + na.h[2], sb = nb.h[0] + nb.h[1] + nb.h[2];
        const bool split_a = nb.left < 0 || (na.left >= 0 && sa >= sb);
        const int pos = sp + 2 * __popc(im & ((1u << lane) - 1u));
        if (pos + 1 < MPB_EXP2_STACK) {
          stack[pos] = split_a ? make_int2(na.left, np.y) : make_int2(np.x, nb.left);
          stack[pos + 1] = split_a ? make_int2(na.right, np.y) : make_int2(np.x, nb.right);
        } else { atomicAdd(&wq.counters[2], 1ull); dropped = true; }  // dropped subtree pair: counted, and the item reads as colliding
      }
      sp = min(sp + 2 * __popc(im), MPB_EXP2_STACK);
      __syncwarp();
      bool lhit = false;
      if (pass && both_leaf) {
        const int fa = ~na.left, fb = ~nb.left;
        for (int i = 0; i < na.right && !lhit; i++) {
          const float4 a0 = __ldg(img.prims + 3 * (fa + i)), a1 = __ldg(img.prims + 3 * (fa + i) + 1), a2 = __ldg(img.prims + 3 * (fa + i) + 2);
          const V3 TA[3] = {v3(a0.x, a0.y, a0.z), v3(a1.x, a1.y, a1.z), v3(a2.x, a2.y, a2.z)};
          for (int j = 0; j < nb.right && !lhit; j++) {
            const float4 b0 = __ldg(img.prims + 3 * (fb + j)), b1 = __ldg(img.prims + 3 * (fb + j) + 1);
            if (leaf_b == MPB_GEOM_TRI) {
              const float4 b2 = __ldg(img.prims + 3 * (fb + j) + 2);
              const V3 TB[3] = {rot(R, v3(b0.x, b0.y, b0.z)) + v3(t[0], t[1], t[2]), rot(R, v3(b1.x, b1.y, b1.z)) + v3(t[0], t[1], t[2]),
                                rot(R, v3(b2.x, b2.y, b2.z)) + v3(t[0], t[1], t[2])};
              lhit = tri_tri_hit(TA, TB);
            } else {  // triangle of A into B's frame, box stays axis-aligned
              const V3 tv = v3(t[0], t[1], t[2]);
              lhit = box_tri_hit(v3(b0.x, b0.y, b0.z), v3(b1.x, b1.y, b1.z), rot_t(R, TA[0] - tv), rot_t(R, TA[1] - tv), rot_t(R, TA[2] - tv));
            }
          }
        }
      }
      if (__any_sync(FULL, lhit || dropped)) { hit = true; break; }
    }
    if (hit && lane == 0) report_hit(out, (unsigned)cfg, p);
    __syncwarp();
  }
}
