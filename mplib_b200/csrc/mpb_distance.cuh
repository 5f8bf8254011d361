// mpb_distance.cuh -- GJK separation distance for the candidate (configuration, pair) items of the distance path.
//
// Replaces ::fcl::distance with the default DistanceRequest (enable_signed_distance = false, tolerance 1e-6,
// python/pybind_fcl.hpp:311-315) as called by PlanningWorldTpl::distanceSelf / distanceOthers
// (src/planning_world.cpp:492-720): the separation distance of a pair, or -1 when the shapes touch or intersect.
// Per configuration the minimum over the ACM-admitted pairs and its arg-min (lowest pair index among ties, the
// reference keeps the first pair that is strictly smaller) are folded with one 64-bit atomicMin.
#pragma once
#include "mpb_kernels.cuh"

// closest point to the origin on segment ab; mask: bit 0 = a, bit 1 = b support the closest point
__device__ __forceinline__ V3 gjk_closest_seg(V3 a, V3 b, int &mask) {
  const V3 ab = b - a;
  const float t = -dot(a, ab), dd = dot(ab, ab);
  if (t <= 0.f || dd <= 0.f) { mask = 1; return a; }
  if (t >= dd) { mask = 2; return b; }
  mask = 3;
  return a + ab * (t / dd);
}

// closest point to the origin on triangle abc (Ericson, Real-Time Collision Detection 5.1.5); mask bits a, b, c
__device__ __forceinline__ V3 gjk_closest_tri(V3 a, V3 b, V3 c, int &mask) {
  const V3 ab = b - a, ac = c - a;
  const float d1 = -dot(ab, a), d2 = -dot(ac, a);
  if (d1 <= 0.f && d2 <= 0.f) { mask = 1; return a; }
  const float d3 = -dot(ab, b), d4 = -dot(ac, b);
  if (d3 >= 0.f && d4 <= d3) { mask = 2; return b; }
  const float vc = d1 * d4 - d3 * d2;
  if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) { mask = 3; return a + ab * (d1 / (d1 - d3)); }
  const float d5 = -dot(ab, c), d6 = -dot(ac, c);
  if (d6 >= 0.f && d5 <= d6) { mask = 4; return c; }
  const float vb = d5 * d2 - d1 * d6;
  if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) { mask = 5; return a + ac * (d2 / (d2 - d6)); }
  const float va = d3 * d6 - d5 * d4;
  if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) { mask = 6; return b + (c - b) * ((d4 - d3) / ((d4 - d3) + (d5 - d6))); }
  const float den = 1.f / (va + vb + vc);
  mask = 7;
  return a + ab * (vb * den) + ac * (vc * den);
}

struct GjkSimplex {
  V3 p[4];
  int n;
};

// Closest point of the simplex to the origin; the simplex is reduced to the sub-simplex that supports it.
// Returns false when the origin is inside the tetrahedron (the shapes intersect).
__device__ __forceinline__ bool gjk_closest(GjkSimplex &S, V3 &v) {
  int mask = 1;
  if (S.n == 1) { v = S.p[0]; return true; }
  if (S.n == 2) v = gjk_closest_seg(S.p[0], S.p[1], mask);
  else if (S.n == 3) v = gjk_closest_tri(S.p[0], S.p[1], S.p[2], mask);
  else {
    float best = FLT_MAX;
    bool outside = false;
    mask = 0;
#pragma unroll
    for (int f = 0; f < 4; f++) {
      // face f leaves out vertex 3 - f
      const int i0 = f == 3 ? 1 : 0, i1 = f >= 2 ? 2 : 1, i2 = f == 0 ? 2 : 3, io = 3 - f;
      const V3 a = S.p[i0], b = S.p[i1], c = S.p[i2];
      const V3 nrm = cross(b - a, c - a);
      const V3 ad = S.p[io] - a;
      const float so = -dot(nrm, a), sd = dot(nrm, ad);
      // origin beyond this face, or a (nearly) flat tetrahedron whose side tests mean nothing in fp32
      if (so * sd < 0.f || sd * sd <= 1e-6f * dot(nrm, nrm) * dot(ad, ad)) {
        outside = true;
        int m;
        const V3 w = gjk_closest_tri(a, b, c, m);
        const float d2 = dot(w, w);
        if (d2 < best) {
          best = d2;
          v = w;
          mask = ((m & 1) ? (1 << i0) : 0) | ((m & 2) ? (1 << i1) : 0) | ((m & 4) ? (1 << i2) : 0);
        }
      }
    }
    if (!outside) return false;
  }
  int k = 0;
#pragma unroll
  for (int i = 0; i < 4; i++)
    if (i < S.n && ((mask >> i) & 1)) S.p[k++] = S.p[i];
  S.n = k;
  return true;
}

#define MPB_GJK_MAX_ITERS 48
#define MPB_GJK_TOL 1e-6f

// Lane per item; a warp owns 32 consecutive items of one bin (same two hull classes), so the hull scans of
// support_local are broadcast shared-memory loads.  best[cfg] = min over items of (key << 32 | pair), where
// key = 0 for touching / intersecting pairs (reported as -1) and the bits of the non-negative distance + 2^31 else.
template <bool SMEM>
__global__ void __launch_bounds__(256) distance_kernel(SceneImg img, WorkQ wq, unsigned long long *__restrict__ best, int n_bins) {
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
      for (unsigned off = 0; off < total; off += 32768u)
        tma_bulk_g2s(smem_raw + off, reinterpret_cast<const unsigned char *>(img.verts) + off, min(32768u, total - off), &bar);
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
    unsigned u = 0;
    if (lane == 0) u = atomicAdd(&wq.totals[3], 1u);
    u = __shfl_sync(FULL, u, 0);
    if (u >= n_units) break;
    int lo = 0, hi = n_bins;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (wq.unit_start[mid] <= u) lo = mid; else hi = mid;
    }
    const unsigned first = wq.bin_start[lo] + (u - wq.unit_start[lo]) * unit_items;
    const unsigned end = min(min(wq.bin_start[lo + 1], first + unit_items), wq.cap);
    for (unsigned i0 = first; i0 < end; i0 += 32) {
      const unsigned my = i0 + lane;
      bool busy = my < end;
      const unsigned src = busy ? my : first;  // idle lanes shadow a real item so the hull scans stay uniform
      float R[9];
      const float4 *rc = item_rec(wq, 0, src);
      float t3[3];
      int cfg, p, aux_;
      item_pose(rc, R, t3);
      item_ids(rc, cfg, p, aux_);
      const V3 t = v3(t3[0], t3[1], t3[2]);
      const DevPair pr = img.pairs[p];
      const DevObject &OA = img.objs[pr.a];
      const DevObject &OB = img.objs[pr.b];
      DevShape A = shape_of(OA), B = shape_of(OB);
      // core + margin: a sphere is a point and a capsule a segment inflated by r, so GJK runs on polytopes
      // (finite convergence, no curved-surface stalling) and the radius is subtracted at the end
      float margin = 0.f;
      if (A.type == MPB_GEOM_SPHERE) { margin += A.p0; A.p0 = 0.f; }
      else if (A.type == MPB_GEOM_CAPSULE) { margin += A.p0; A.p0 = 0.f; }
      if (B.type == MPB_GEOM_SPHERE) { margin += B.p0; B.p0 = 0.f; }
      else if (B.type == MPB_GEOM_CAPSULE) { margin += B.p0; B.p0 = 0.f; }
      // Minkowski difference A - B in the frame of A; first direction: from B's centre towards A's
      const V3 cB = rot(R, v3(OB.ctr[0], OB.ctr[1], OB.ctr[2])) + t;
      V3 v = v3(OA.ctr[0], OA.ctr[1], OA.ctr[2]) - cB;
      if (dot(v, v) < 1e-20f) v = v3(1.f, 0.f, 0.f);
      GjkSimplex S;
      S.n = 0;
      float dist = -1.f;
      int iters = 0;
      bool have = false;
      while (__any_sync(FULL, busy)) {
        // support of A - B along -v
        const V3 dir = normalized(v) * -1.f;
        const V3 sa = support_local<SMEM>(A, dir, verts, img.prims, cells);
        const V3 sb = support_local<SMEM>(B, rot_t(R, dir * -1.f), verts, img.prims, cells);
        const V3 w = sa - (rot(R, sb) + t);
        if (!busy) continue;
        iters++;
#ifdef MPB_DEBUG_GJK
        if (cfg == MPB_DEBUG_GJK) printf("it %d n %d v %g %g %g w %g %g %g sa %g %g %g sb %g %g %g types %d %d\n", iters, S.n, v.x, v.y, v.z, w.x, w.y, w.z, sa.x, sa.y, sa.z, sb.x, sb.y, sb.z, A.type, B.type);
#endif
        if (have) {
          const float vl = sqrtf(dot(v, v));
          const float lb = -dot(w, dir);  // supporting-plane lower bound of the distance
          bool dup = false;
          for (int i = 0; i < S.n; i++) {
            const V3 e = S.p[i] - w;
            dup |= dot(e, e) <= 1e-14f;
          }
          if (vl - lb <= MPB_GJK_TOL || dup) { dist = vl; busy = false; continue; }
        }
        S.p[S.n++] = w;
        if (!gjk_closest(S, v)) { dist = -1.f; busy = false; continue; }
        have = true;
        if (dot(v, v) < 1e-12f) { dist = -1.f; busy = false; continue; }  // touching
        if (iters >= MPB_GJK_MAX_ITERS) { dist = sqrtf(dot(v, v)); cap_sum++; busy = false; }
      }
      if (my < end) {
        it_sum += (unsigned)iters;
        if (dist >= 0.f) { dist -= margin; if (dist <= 0.f) dist = -1.f; }
        const unsigned key = dist < 0.f ? 0u : (__float_as_uint(dist) | 0x80000000u);
        atomicMin(&best[cfg], ((unsigned long long)key << 32) | (unsigned)p);
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

// ---------------------------------------------------------------------------------------------- BVH distance
// ::fcl::distance for pairs that involve a BVHModel or an OcTree (src/planning_world.cpp:492-712 reach them through
// distanceSelf for `convex=False` links and distanceOthers for mesh obstacles / the point-cloud octree): the minimum over
// the leaf primitives of the leaf distance, found by branch and bound over the object's AABB tree -- the role of
// MeshShapeDistanceTraversalNodeOBBRSS / OcTreeSolver::OcTreeShapeDistanceRecurse / MeshDistanceTraversalNodeOBBRSS /
// OcTreeMeshDistanceRecurse [upstream FCL 0.7.0].  A node is pruned when a lower bound of its distance exceeds the
// smallest distance known for the configuration (from ANY admitted pair: best[cfg] is re-read every trip), so the
// convex / primitive pairs, which run first, already cut most of the tree.

// One GJK run per LANE (leaf primitives differ per lane; the loop is divergent by nature).  Same iteration as
// distance_kernel.  Returns the separation distance, or -1 when the two shapes touch or intersect.
__device__ __forceinline__ float gjk_lane(DevShape A, DevShape B, const float *R, V3 t, V3 cA, V3 cBl, const float4 *__restrict__ verts,
                                          const float4 *__restrict__ prims, const unsigned char *__restrict__ cells, int &iters, bool &capped) {
  float margin = 0.f;
  if (A.type == MPB_GEOM_SPHERE || A.type == MPB_GEOM_CAPSULE) { margin += A.p0; A.p0 = 0.f; }
  if (B.type == MPB_GEOM_SPHERE || B.type == MPB_GEOM_CAPSULE) { margin += B.p0; B.p0 = 0.f; }
  V3 v = cA - (rot(R, cBl) + t);
  if (dot(v, v) < 1e-20f) v = v3(1.f, 0.f, 0.f);
  GjkSimplex S;
  S.n = 0;
  bool have = false;
  float dist = -1.f;
  for (;;) {
    const V3 dir = normalized(v) * -1.f;
    const V3 sa = support_local<false>(A, dir, verts, prims, cells);
    const V3 sb = support_local<false>(B, rot_t(R, dir * -1.f), verts, prims, cells);
    const V3 w = sa - (rot(R, sb) + t);
    iters++;
    if (have) {
      const float vl = sqrtf(dot(v, v));
      const float lb = -dot(w, dir);
      bool dup = false;
      for (int i = 0; i < S.n; i++) {
        const V3 e = S.p[i] - w;
        dup |= dot(e, e) <= 1e-14f;
      }
      if (vl - lb <= MPB_GJK_TOL || dup) { dist = vl; break; }
    }
    S.p[S.n++] = w;
    if (!gjk_closest(S, v)) { dist = -1.f; break; }
    have = true;
    if (dot(v, v) < 1e-12f) { dist = -1.f; break; }
    if (iters >= MPB_GJK_MAX_ITERS) { dist = sqrtf(dot(v, v)); capped = true; break; }
  }
  if (dist >= 0.f) { dist -= margin; if (dist <= 0.f) dist = -1.f; }
  return dist;
}

// Smallest distance known so far for the configuration, as a pruning bound for pair p.  An intersecting pair (-1) found
// by a pair of HIGHER index does not end the search: the arg-min is the first intersecting pair in list order
// (the reference keeps the first pair that attains the minimum), so p must still find out whether it intersects --
// the bound is then 0 (only overlapping nodes survive).  -> false: a pair of lower index already intersects, p cannot matter.
__device__ __forceinline__ bool best_bound(const unsigned long long *best, int cfg, int p, float &ub) {
  const unsigned long long b = *reinterpret_cast<const volatile unsigned long long *>(best + cfg);
  if (b == ~0ull) return true;
  const unsigned key = (unsigned)(b >> 32);
  if (key == 0u) {
    if ((unsigned)(b & 0xffffffffu) < (unsigned)p) return false;
    ub = fminf(ub, 0.f);
    return true;
  }
  ub = fminf(ub, __uint_as_float(key & 0x7fffffffu));
  return true;
}
__device__ __forceinline__ float gap3(float d0, float d1, float d2) {  // length of the positive parts
  d0 = fmaxf(d0, 0.f); d1 = fmaxf(d1, 0.f); d2 = fmaxf(d2, 0.f);
  return sqrtf(fmaf(d0, d0, fmaf(d1, d1, d2 * d2)));
}
// bounds of a leaf primitive in its object's frame: centre, half extents
__device__ __forceinline__ void prim_box(const float4 *__restrict__ prims, int prim, int leaf_type, float *pc, float *ph) {
  const float4 a = __ldg(prims + 3 * prim), b = __ldg(prims + 3 * prim + 1);
  if (leaf_type == MPB_GEOM_VBOX) { pc[0] = a.x; pc[1] = a.y; pc[2] = a.z; ph[0] = b.x; ph[1] = b.y; ph[2] = b.z; return; }
  const float4 c = __ldg(prims + 3 * prim + 2);
  const float lo0 = fminf(a.x, fminf(b.x, c.x)), hi0 = fmaxf(a.x, fmaxf(b.x, c.x));
  const float lo1 = fminf(a.y, fminf(b.y, c.y)), hi1 = fmaxf(a.y, fmaxf(b.y, c.y));
  const float lo2 = fminf(a.z, fminf(b.z, c.z)), hi2 = fmaxf(a.z, fmaxf(b.z, c.z));
  pc[0] = 0.5f * (lo0 + hi0); pc[1] = 0.5f * (lo1 + hi1); pc[2] = 0.5f * (lo2 + hi2);
  ph[0] = 0.5f * (hi0 - lo0); ph[1] = 0.5f * (hi1 - lo1); ph[2] = 0.5f * (hi2 - lo2);
}
__device__ __forceinline__ V3 prim_centre(const float4 *__restrict__ prims, int prim, int leaf_type) {  // centerTriangle / box pose
  const float4 a = __ldg(prims + 3 * prim);
  if (leaf_type != MPB_GEOM_TRI) return v3(a.x, a.y, a.z);
  const float4 b = __ldg(prims + 3 * prim + 1), c = __ldg(prims + 3 * prim + 2);
  return v3((a.x + b.x + c.x) * (1.f / 3.f), (a.y + b.y + c.y) * (1.f / 3.f), (a.z + b.z + c.z) * (1.f / 3.f));
}
#define MPB_LB_SLACK 1e-6f  // fp32 rounding of the bounds: a node is kept unless its bound clears the best distance by this

// warp per (configuration, shape x BVH object) item of the expand bin
__global__ void __launch_bounds__(128) distance_expand_kernel(SceneImg img, WorkQ wq, unsigned long long *__restrict__ best) {
  __shared__ int s_stack[4][MPB_EXP_STACK];
  const int eb = img.expand_bin;
  if (eb < 0) return;
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int *stack = s_stack[warp];
  const unsigned char *cells = reinterpret_cast<const unsigned char *>(img.verts + img.n_verts);
  const unsigned first = wq.bin_start[eb], end = min(wq.bin_start[eb + 1], wq.cap);
  unsigned long long it_sum = 0, cap_sum = 0;
  for (;;) {
    unsigned it = 0;
    if (lane == 0) it = atomicAdd(&wq.totals[5], 1u);
    it = first + __shfl_sync(FULL, it, 0);
    if (it >= end) break;
    const float4 *rc = item_rec(wq, 0, it);
    int cfg, p, aux_;
    item_ids(rc, cfg, p, aux_);
    float ub = FLT_MAX;
    if (!best_bound(best, cfg, p, ub)) continue;  // a pair of lower index already intersects: this one cannot change the result
    const DevPair pr = img.pairs[p];
    const DevObject &A = img.objs[pr.a];  // the shape
    const DevObject &B = img.objs[pr.b];  // the BVH object
    if (B.voff < 0) continue;             // empty octree: no leaf, no distance
    float R[9], t3[3];
    item_pose(rc, R, t3);
    const V3 t = v3(t3[0], t3[1], t3[2]);
    // the shape seen from the object's frame:  x_B = R^T (x_A - t)
    const V3 cs = rot_t(R, v3(A.bc[0], A.bc[1], A.bc[2]) - t);   // bounding-sphere centre
    const V3 co = rot_t(R, v3(A.oc[0], A.oc[1], A.oc[2]) - t);   // OBB centre
    float bh[3];  // axis-aligned bounds of the shape's OBB in the object's frame
#pragma unroll
    for (int r = 0; r < 3; r++) {
      float acc = 0.f;
#pragma unroll
      for (int c = 0; c < 3; c++) {
        const float m = fmaf(R[r], A.oR[c], fmaf(R[3 + r], A.oR[3 + c], R[6 + r] * A.oR[6 + c]));  // (R^T oR)[r][c]
        acc = fmaf(fabsf(m), A.oh[c], acc);
      }
      bh[r] = acc;
    }
    const int leaf_type = B.vcnt;
    DevShape SA = shape_of(A), SB;
    SB.type = leaf_type; SB.voff = 0; SB.vcnt = 0; SB.cell = -1; SB.p0 = SB.p1 = SB.p2 = 0.f;
    const V3 cA = v3(A.ctr[0], A.ctr[1], A.ctr[2]);
    float local = FLT_MAX;
    int sp = 1;
    if (lane == 0) stack[0] = B.voff;
    __syncwarp();
    while (sp > 0) {
      if (!best_bound(best, cfg, p, ub)) break;
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
        const float lb1 = gap3(fabsf(nd.c[0] - cs.x) - nd.h[0], fabsf(nd.c[1] - cs.y) - nd.h[1], fabsf(nd.c[2] - cs.z) - nd.h[2]) - A.br;
        const float lb2 = gap3(fabsf(nd.c[0] - co.x) - nd.h[0] - bh[0], fabsf(nd.c[1] - co.y) - nd.h[1] - bh[1], fabsf(nd.c[2] - co.z) - nd.h[2] - bh[2]);
        pass = fmaxf(lb1, lb2) <= ub + MPB_LB_SLACK;
      }
      const bool inner = pass && nd.left >= 0;
      const unsigned im = __ballot_sync(FULL, inner);
      bool dropped = false;
      if (inner) {
        const int pos = sp + 2 * __popc(im & ((1u << lane) - 1u));
        if (pos + 1 < MPB_EXP_STACK) { stack[pos] = nd.left; stack[pos + 1] = nd.right; }
        else dropped = true;
      }
      sp = min(sp + 2 * __popc(im), MPB_EXP_STACK);
      __syncwarp();
      if (__any_sync(FULL, dropped)) cap_sum++;  // a subtree that could not be held: counted (narrow_capped), the minimum may miss it
      float dmin = FLT_MAX;
      if (pass && nd.left < 0) {
        const int pfirst = ~nd.left;
        for (int k = 0; k < nd.right; k++) {
          const int prim = pfirst + k;
          float pc[3], ph[3];
          prim_box(img.prims, prim, leaf_type, pc, ph);
          const float lb1 = gap3(fabsf(pc[0] - cs.x) - ph[0], fabsf(pc[1] - cs.y) - ph[1], fabsf(pc[2] - cs.z) - ph[2]) - A.br;
          const float lb2 = gap3(fabsf(pc[0] - co.x) - ph[0] - bh[0], fabsf(pc[1] - co.y) - ph[1] - bh[1], fabsf(pc[2] - co.z) - ph[2] - bh[2]);
          if (fmaxf(lb1, lb2) > fminf(ub, dmin) + MPB_LB_SLACK) continue;
          SB.voff = prim;
          int iters = 0;
          bool capped = false;
          const float d = gjk_lane(SA, SB, R, t, cA, prim_centre(img.prims, prim, leaf_type), img.verts, img.prims, cells, iters, capped);
          it_sum += (unsigned)iters;
          cap_sum += capped ? 1u : 0u;
          dmin = fminf(dmin, d);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) dmin = fminf(dmin, __shfl_xor_sync(FULL, dmin, o));
      if (dmin < local) local = dmin;
      if (dmin < 0.f) break;  // this pair intersects: nothing smaller to find
      ub = fminf(ub, dmin);
    }
    if (lane == 0 && local < FLT_MAX) {
      const unsigned key = local < 0.f ? 0u : (__float_as_uint(local) | 0x80000000u);
      atomicMin(&best[cfg], ((unsigned long long)key << 32) | (unsigned)p);
    }
    __syncwarp();
  }
  for (int o = 16; o > 0; o >>= 1) {
    it_sum += __shfl_xor_sync(FULL, it_sum, o);
    cap_sum += __shfl_xor_sync(FULL, cap_sum, o);
  }
  if (lane == 0 && it_sum) atomicAdd(&wq.counters[1], it_sum);
  if (lane == 0 && cap_sum) atomicAdd(&wq.counters[2], cap_sum);
}

// warp per (configuration, mesh x mesh | mesh x octree) item of the expand2 bin; A = mesh, B = mesh or octree
__global__ void __launch_bounds__(128) distance_expand2_kernel(SceneImg img, WorkQ wq, unsigned long long *__restrict__ best) {
  __shared__ int2 s_stack[4][MPB_EXP2_STACK];
  const int eb = img.expand2_bin;
  if (eb < 0) return;
  const unsigned FULL = 0xffffffffu;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int2 *stack = s_stack[warp];
  const unsigned char *cells = reinterpret_cast<const unsigned char *>(img.verts + img.n_verts);
  const unsigned first = wq.bin_start[eb], end = min(wq.bin_start[eb + 1], wq.cap);
  unsigned long long it_sum = 0, cap_sum = 0;
  for (;;) {
    unsigned it = 0;
    if (lane == 0) it = atomicAdd(&wq.totals[7], 1u);
    it = first + __shfl_sync(FULL, it, 0);
    if (it >= end) break;
    const float4 *rc = item_rec(wq, 0, it);
    int cfg, p, aux_;
    item_ids(rc, cfg, p, aux_);
    float ub = FLT_MAX;
    if (!best_bound(best, cfg, p, ub)) continue;
    const DevPair pr = img.pairs[p];
    const DevObject &A = img.objs[pr.a];
    const DevObject &B = img.objs[pr.b];
    if (B.voff < 0) continue;
    float R[9], t3[3];  // B in A
    item_pose(rc, R, t3);
    const V3 t = v3(t3[0], t3[1], t3[2]);
    const int leaf_b = B.vcnt;
    DevShape SA, SB;
    SA.type = MPB_GEOM_TRI; SA.voff = 0; SA.vcnt = 0; SA.cell = -1; SA.p0 = SA.p1 = SA.p2 = 0.f;
    SB = SA;
    SB.type = leaf_b;
    float local = FLT_MAX;
    int sp = 1;
    if (lane == 0) stack[0] = make_int2(A.voff, B.voff);
    __syncwarp();
    while (sp > 0) {
      if (!best_bound(best, cfg, p, ub)) break;
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
        const V3 cb = rot(R, v3(nb.c[0], nb.c[1], nb.c[2])) + t;  // B node centre in A's frame; its box lies inside the sphere |h|
        const float rb = sqrtf(fmaf(nb.h[0], nb.h[0], fmaf(nb.h[1], nb.h[1], nb.h[2] * nb.h[2])));
        const float lb = gap3(fabsf(cb.x - na.c[0]) - na.h[0], fabsf(cb.y - na.c[1]) - na.h[1], fabsf(cb.z - na.c[2]) - na.h[2]) - rb;
        pass = lb <= ub + MPB_LB_SLACK;
      }
      const bool both_leaf = na.left < 0 && nb.left < 0;
      const bool inner = pass && !both_leaf;
      const unsigned im = __ballot_sync(FULL, inner);
      bool dropped = false;
      if (inner) {
        const float sa = na.h[0] + na.h[1] + na.h[2], sb = nb.h[0] + nb.h[1] + nb.h[2];
        const bool split_a = nb.left < 0 || (na.left >= 0 && sa >= sb);
        const int pos = sp + 2 * __popc(im & ((1u << lane) - 1u));
        if (pos + 1 < MPB_EXP2_STACK) {
          stack[pos] = split_a ? make_int2(na.left, np.y) : make_int2(np.x, nb.left);
          stack[pos + 1] = split_a ? make_int2(na.right, np.y) : make_int2(np.x, nb.right);
        } else dropped = true;
      }
      sp = min(sp + 2 * __popc(im), MPB_EXP2_STACK);
      __syncwarp();
      if (__any_sync(FULL, dropped)) cap_sum++;
      float dmin = FLT_MAX;
      if (pass && both_leaf) {
        const int fa = ~na.left, fb = ~nb.left;
        for (int i = 0; i < na.right; i++)
          for (int j = 0; j < nb.right; j++) {
            SA.voff = fa + i;
            SB.voff = fb + j;
            int iters = 0;
            bool capped = false;
            float d = gjk_lane(SA, SB, R, t, prim_centre(img.prims, fa + i, MPB_GEOM_TRI), prim_centre(img.prims, fb + j, leaf_b), img.verts, img.prims,
                               cells, iters, capped);
            if (d < 0.f && leaf_b == MPB_GEOM_TRI) d = 0.f;  // [upstream] TriangleDistance::triDistance returns 0 for overlapping triangles
            it_sum += (unsigned)iters;
            cap_sum += capped ? 1u : 0u;
            dmin = fminf(dmin, d);
          }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) dmin = fminf(dmin, __shfl_xor_sync(FULL, dmin, o));
      if (dmin < local) local = dmin;
      if (dmin <= 0.f) break;  // this pair intersects (0: overlapping triangles of two meshes): nothing smaller to find
      ub = fminf(ub, dmin);
    }
    if (lane == 0 && local < FLT_MAX) {
      const unsigned key = local < 0.f ? 0u : (__float_as_uint(local) | 0x80000000u);
      atomicMin(&best[cfg], ((unsigned long long)key << 32) | (unsigned)p);
    }
    __syncwarp();
  }
  for (int o = 16; o > 0; o >>= 1) {
    it_sum += __shfl_xor_sync(FULL, it_sum, o);
    cap_sum += __shfl_xor_sync(FULL, cap_sum, o);
  }
  if (lane == 0 && it_sum) atomicAdd(&wq.counters[1], it_sum);
  if (lane == 0 && cap_sum) atomicAdd(&wq.counters[2], cap_sum);
}

// best[] -> (min_distance, argmin pair); a configuration with no admitted pair keeps FLT_MAX / -1
__global__ void distance_unpack_kernel(const unsigned long long *__restrict__ best, long long n, float *__restrict__ min_dist,
                                       int *__restrict__ argmin) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long b = best[i];
  const unsigned key = (unsigned)(b >> 32);
  float d;
  int p = (int)(unsigned)(b & 0xffffffffu);
  if (b == ~0ull) { d = FLT_MAX; p = -1; }
  else if (key == 0u) d = -1.f;
  else d = __uint_as_float(key & 0x7fffffffu);
  min_dist[i] = d;
  if (argmin) argmin[i] = p;
}
