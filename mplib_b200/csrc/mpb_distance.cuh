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
#pragma unroll
      for (int k = 0; k < 9; k++) R[k] = wq.itf[(size_t)k * wq.cap + src];
      const V3 t = v3(wq.itf[(size_t)9 * wq.cap + src], wq.itf[(size_t)10 * wq.cap + src], wq.itf[(size_t)11 * wq.cap + src]);
      const int cfg = wq.icfg[src], p = wq.ipair[src];
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
