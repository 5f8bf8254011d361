// mpb_plan.cuh -- kernels of the planning layer on top of the batched validity check.
//
//   motion_count_kernel / motion_expand_kernel / motion_reduce_kernel
//       Batched ompl::base::DiscreteMotionValidator::checkMotion [upstream OMPL 1.6.0] as the reference sets it up
//       (src/ompl_planner.cpp:62-70, validity resolution 1 %): an edge (a, b) is valid iff b and the interior samples
//       a + (j/nd)(b - a), j = 1..nd-1, are valid, with nd = max over the 1-D subspaces of
//       ceil(|b_j - a_j| / (0.01 * extent_j))  (CompoundStateSpace::validSegmentCount).  All samples of all edges
//       become ONE batch for the validity kernels; a per-edge AND folds the verdicts.
//   rrt_advance_kernel / rrt_extract_kernel
//       RRTConnect [upstream OMPL 1.6.0 geometric/planners/rrt/src/RRTConnect.cpp] for Q independent queries in
//       lock-step (replaces the RRTConnect branch of OMPLPlannerTpl::plan, src/ompl_planner.cpp:97-243).  One
//       warp per query; one growTree() call per query per round.  A round is: advance (consume the previous
//       edge verdict, update the trees, emit the next edge) -> expand the edges into samples -> validity kernels.
//       The per-round kernel sequence has fixed shapes, so the host captures it once in a CUDA graph.
#pragma once
#include "mpb_kernels.cuh"

// distance / interpolation of the compound space: 1-D RealVector subspaces (L1 sum) and SO(2) subspaces
__device__ __forceinline__ float ss_delta(float a, float b, bool so2) {
  float d = b - a;
  if (so2) {
    const float PI_F = 3.14159265358979323846f;
    if (d > PI_F) d -= 2.f * PI_F;
    else if (d < -PI_F) d += 2.f * PI_F;
  }
  return d;
}
__device__ __forceinline__ float ss_interp(float a, float b, float t, bool so2) {
  float v = fmaf(ss_delta(a, b, so2), t, a);
  if (so2) {
    const float PI_F = 3.14159265358979323846f;
    if (v > PI_F) v -= 2.f * PI_F;
    else if (v < -PI_F) v += 2.f * PI_F;
  }
  return v;
}

struct SpaceDesc {
  int dim;
  const float *lo, *hi;      // [dim] sampling bounds
  const float *seg;          // [dim] longest valid segment per subspace (0.01 * extent)
  const unsigned char *so2;  // [dim] 1 = SO(2) subspace (continuous joint)
};

__device__ __forceinline__ int ss_segment_count(const SpaceDesc &sp, const float *a, const float *b) {
  int nd = 0;
  for (int j = 0; j < sp.dim; j++) {
    const float d = fabsf(ss_delta(a[j], b[j], sp.so2[j] != 0));
    nd = max(nd, (int)ceilf(d / sp.seg[j]));
  }
  return nd;
}

// ---------------------------------------------------------------------------------------------- motion validation
// pass 1: samples per edge (max(nd, 1): b itself plus nd-1 interior points) and unordered slot allocation
__global__ void __launch_bounds__(256) motion_count_kernel(SpaceDesc sp, const float *__restrict__ a, const float *__restrict__ b,
                                                           long long n_edges, int *__restrict__ nd_out,
                                                           unsigned *__restrict__ first_out, unsigned *__restrict__ total) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int nd = 0;
  if (e < n_edges) nd = ss_segment_count(sp, a + e * sp.dim, b + e * sp.dim);
  const unsigned cnt = e < n_edges ? (unsigned)max(nd, 1) : 0u;
  // warp-aggregated allocation
  unsigned incl = cnt;
  const unsigned lane = threadIdx.x & 31;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned t = __shfl_up_sync(0xffffffffu, incl, o);
    if ((int)lane >= o) incl += t;
  }
  const unsigned wtot = __shfl_sync(0xffffffffu, incl, 31);
  unsigned base = 0;
  if (lane == 31 && wtot) base = atomicAdd(total, wtot);
  base = __shfl_sync(0xffffffffu, base, 31);
  if (e < n_edges) {
    nd_out[e] = nd;
    first_out[e] = base + incl - cnt;
  }
}

// pass 2: write the samples.  Fixed-stride mode (first == nullptr): edge e owns slots [e*stride, (e+1)*stride) and
// unused slots repeat b, so the per-edge OR over all its slots is unchanged.
__global__ void __launch_bounds__(256) motion_expand_kernel(SpaceDesc sp, const float *__restrict__ a, const float *__restrict__ b,
                                                            long long n_edges, const int *__restrict__ nd_in,
                                                            const unsigned *__restrict__ first, int stride,
                                                            float *__restrict__ states, int *__restrict__ owner) {
  const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n_edges) return;
  const float *pa = a + e * sp.dim, *pb = b + e * sp.dim;
  const int nd = nd_in ? nd_in[e] : ss_segment_count(sp, pa, pb);
  const int cnt = first ? max(nd, 1) : stride;
  const size_t s0 = first ? (size_t)first[e] : (size_t)e * stride;
  for (int k = 0; k < cnt; k++) {
    const int j = k + 1;  // sample j/nd; j >= nd is b itself
    float *o = states + (s0 + k) * sp.dim;
    if (j >= nd) {
      for (int c = 0; c < sp.dim; c++) o[c] = pb[c];
    } else {
      const float t = (float)j / (float)nd;
      for (int c = 0; c < sp.dim; c++) o[c] = ss_interp(pa[c], pb[c], t, sp.so2[c] != 0);
    }
    if (owner) owner[s0 + k] = (int)e;
  }
}

// pass 3: valid_bits starts all-ones; a colliding sample clears the bit of its edge
__global__ void __launch_bounds__(256) motion_reduce_kernel(const unsigned *__restrict__ verdict_bits, const int *__restrict__ owner,
                                                            unsigned n_states, unsigned *__restrict__ valid_bits) {
  const unsigned s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_states) return;
  if ((verdict_bits[s >> 5] >> (s & 31)) & 1u) {
    const int e = owner[s];
    atomicAnd(&valid_bits[e >> 5], ~(1u << (e & 31)));
  }
}

__global__ void fill_valid_bits_kernel(unsigned *__restrict__ valid_bits, long long n_edges) {
  const long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long words = (n_edges + 31) / 32;
  if (w >= words) return;
  const long long rem = n_edges - w * 32;
  valid_bits[w] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u);
}

// ---------------------------------------------------------------------------------------------- RRTConnect, lock-step
// Queries run in SLOTS (one warp each).  A slot whose query finishes -- solved, out of its per-query round budget, tree
// full, no goal -- writes the query's result (status, path) at once and takes the next pending query in the following
// round, so a batch of Q queries flows through n_slots <= Q concurrent searches and no slot idles while work is left.
struct RrtState {
  int n_queries, n_slots, cap, dim, smax, max_goals, max_path;
  int S;               // sub-edges a slot may emit per round (connect steps validated speculatively)
  int max_rounds_q;    // rounds a query may occupy its slot (<= 0: unlimited); the per-query budget
  // ---- per slot
  float *nodes;        // [slots][2][cap][dim]   tree 0 = tStart_, tree 1 = tGoal_
  int *parent;         // [slots][2][cap]        -1 = root
  int *size;           // [slots][2]
  int *slot_query;     // [slots]  query in the slot, -1 = idle
  int *rounds_used;    // [slots]
  int *goals_used;     // [slots]  pis_.getSampledGoalsCount()
  int *phase;          // [slots]  0 = nothing pending, 3 = extend edge + first connect steps pending, 2 = further connect steps pending
  int *flag_start;     // [slots]  startTree_
  int *tsel;           // [slots]  tree the pending edges grow
  int *nearest;        // [slots]  nmotion of the first pending edge
  int *nearest2;       // [slots]  phase 3: nmotion of the first connect step (other tree)
  int *reach;          // [slots]  the last pending edge ends on the target
  int *n_steps;        // [slots]  pending sub-edges (1 .. S)
  int *added;          // [slots]  addedMotion (index in tree 1 - tsel while connecting)
  int *iters;          // [slots]  growTree calls consumed by the slot's query
  float *target;       // [slots][dim]  rstate
  float *edge_a;       // [slots][S][dim]  sub-edge k runs from edge_a[k] to edge_b[k]; edge_a[k] = edge_b[k-1]
  float *edge_b;       // [slots][S][dim]
  // ---- per query
  const float *start;  // [Q][dim]
  const float *goals;  // [Q][max_goals][dim]
  const int *n_goals;  // [Q]
  int *q_status;       // [Q]  MPB_RRT_*
  int *q_pathlen;      // [Q]
  int *q_iters;        // [Q]  growTree calls
  float *q_paths;      // [Q][max_path][dim]
  unsigned *pending;   // [0] next query to start   [1] queries finished
  unsigned long long seed;
};

__device__ __forceinline__ float rrt_uniform(unsigned long long seed, unsigned q, unsigned counter, unsigned j) {
  unsigned long long z = seed + 0x9E3779B97F4A7C15ull * ((((unsigned long long)q << 32) | counter) * 64ull + j + 1ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;  // splitmix64 finaliser
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (float)(z >> 40) * (1.0f / 16777216.0f);
}

// A slot that is idle (or a speculative sub-edge that is not used this round) must not cost narrowphase work:
// its edge slots are set to NaN, which every bounding-sphere comparison of the sphere kernels rejects.
__device__ __forceinline__ void rrt_blank_edges(float *ea, float *eb, int n, int lane) {
  const float nanv = __int_as_float(0x7fc00000);
  for (int i = lane; i < n; i += 32) { ea[i] = nanv; eb[i] = nanv; }
}

// nearest node of a tree to `tg` (L1 over the 1-D subspaces, SO(2) wrap); ties -> lowest index
__device__ __forceinline__ int rrt_nearest(const float *base, int n, int D, const float *tg, const SpaceDesc &sp, int lane, float *best_out) {
  float best = FLT_MAX;
  int bi = 0x7fffffff;
  for (int i = lane; i < n; i += 32) {
    float d = 0.f;
    for (int j = 0; j < D; j++) d += fabsf(ss_delta(base[(size_t)i * D + j], tg[j], sp.so2[j] != 0));
    if (d < best) { best = d; bi = i; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
  }
  *best_out = best;
  return bi;
}

// Sub-edges of the walk from node `from` towards `tg` in steps of `range` (the last one lands on `tg`): at most
// `max_steps` of them into slots slot0.. of (ea, eb); the other slots up to slot_end are blanked.  -> steps written
__device__ __forceinline__ int rrt_emit_walk(float *ea, float *eb, int slot0, int slot_end, int max_steps, const float *from, float best,
                                             const float *tg, float range, int D, const SpaceDesc &sp, int lane, bool *reach) {
  const int k_total = best > range ? (int)ceilf(best / range) : 1;
  const int ns = min(max_steps, k_total);
  *reach = ns == k_total;
  const float nanv = __int_as_float(0x7fc00000);
  for (int idx = lane; idx < (slot_end - slot0) * D; idx += 32) {
    const int k = idx / D, j = idx - k * D;
    const float a = from[j];
    const bool so2 = sp.so2[j] != 0;
    // point k of the walk (k = 0: the start node)
    const float pk = k == 0 ? a : ((k == k_total) ? tg[j] : ss_interp(a, tg[j], (float)k * range / best, so2));
    const float pk1 = (k + 1 >= k_total) ? tg[j] : ss_interp(a, tg[j], (float)(k + 1) * range / best, so2);
    ea[slot0 * D + idx] = k < ns ? pk : nanv;
    eb[slot0 * D + idx] = k < ns ? pk1 : nanv;
  }
  return ns;
}

// The query of slot w ends with `status`: its result is written now (the slot's trees are about to be reused).
// Solved: the path is the start chain (root first) followed by the goal chain (towards its root).
__device__ __forceinline__ void rrt_finish(const RrtState &r, int w, int q, int status, int sol_s, int sol_g, int lane) {
  const int D = r.dim;
  int plen = 0;
  if (status == MPB_RRT_SOLVED) {
    const int *p0 = r.parent + ((size_t)w * 2 + 0) * r.cap, *p1 = r.parent + ((size_t)w * 2 + 1) * r.cap;
    int ls = 0, lg = 0;
    for (int i = sol_s; i >= 0; i = p0[i]) ls++;
    for (int i = sol_g; i >= 0; i = p1[i]) lg++;
    if (ls + lg > r.max_path) status = MPB_RRT_PATH_TOO_LONG;
    else {
      float *out = r.q_paths + (size_t)q * r.max_path * D;
      const float *n0 = r.nodes + ((size_t)w * 2 + 0) * r.cap * D, *n1 = r.nodes + ((size_t)w * 2 + 1) * r.cap * D;
      int k = ls - 1;
      for (int i = sol_s; i >= 0; i = p0[i], k--)
        for (int j = lane; j < D; j += 32) out[(size_t)k * D + j] = n0[(size_t)i * D + j];
      k = ls;
      for (int i = sol_g; i >= 0; i = p1[i], k++)
        for (int j = lane; j < D; j += 32) out[(size_t)k * D + j] = n1[(size_t)i * D + j];
      plen = ls + lg;
    }
  }
  if (lane == 0) {
    r.q_status[q] = status;
    r.q_pathlen[q] = plen;
    r.q_iters[q] = r.iters[w];
    r.slot_query[w] = -1;
    atomicAdd(&r.pending[1], 1u);
  }
  rrt_blank_edges(r.edge_a + (size_t)w * r.S * D, r.edge_b + (size_t)w * r.S * D, r.S * D, lane);
}

// One warp per slot.  A round consumes the verdicts of the slot's pending sub-edges and emits the next ones.
// An iteration of RRTConnect::solve is  growTree(tree, rstate)  followed, unless TRAPPED, by the connect loop
// `while (gsc == ADVANCED) gsc = growTree(otherTree, tgi, rmotion)`.  Both are emitted in ONE round (phase 3): slot 0
// carries the extend edge; its end state q_new is known before it is validated, so slots 1.. carry the first steps
// of the connect walk from the other tree's nearest node towards q_new.  The walk is a straight line in steps of
// `range` -- after every ADVANCED step the node just added is the nearest one (its distance to the target dropped by
// exactly `range`, every other node is at least as far as the previous nearest) -- so its steps can be validated
// speculatively; the valid prefix is added node by node exactly as the sequential loop would, the first invalid step
// is the TRAPPED call, and a TRAPPED extend discards the speculative connect results.  A walk longer than the slots
// continues in the following rounds (phase 2) with all S slots.
__global__ void __launch_bounds__(256) rrt_advance_kernel(RrtState r, SpaceDesc sp, float range, const unsigned *__restrict__ verdict_bits) {
  const unsigned FULL = 0xffffffffu;
  const int w = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  const int lane = threadIdx.x & 31;
  if (w >= r.n_slots) return;
  const int D = r.dim, S = r.S;
  float *ea = r.edge_a + (size_t)w * S * D, *eb = r.edge_b + (size_t)w * S * D, *tg = r.target + (size_t)w * D;
  int q = r.slot_query[w];
  if (q < 0) {
    // ---- idle slot: take the next pending query (the trees restart from its start state)
    unsigned nxt = 0;
    if (lane == 0) nxt = r.pending[0] < (unsigned)r.n_queries ? atomicAdd(&r.pending[0], 1u) : 0xffffffffu;
    nxt = __shfl_sync(FULL, nxt, 0);
    if (nxt >= (unsigned)r.n_queries) return;  // nothing left: the slot stays idle (its edges are already blank)
    q = (int)nxt;
    float *root = r.nodes + ((size_t)w * 2 + 0) * r.cap * D;
    for (int j = lane; j < D; j += 32) root[j] = r.start[(size_t)q * D + j];
    if (lane == 0) {
      r.parent[((size_t)w * 2 + 0) * r.cap] = -1;
      r.size[w * 2] = 1; r.size[w * 2 + 1] = 0;   // tStart_ holds the start state
      r.goals_used[w] = 0; r.phase[w] = 0; r.flag_start[w] = 1; r.iters[w] = 0; r.rounds_used[w] = 0;
      r.slot_query[w] = q;
    }
    __syncwarp();
  }
  if (r.max_rounds_q > 0) {
    const int used = r.rounds_used[w];
    if (used >= r.max_rounds_q) { rrt_finish(r, w, q, MPB_RRT_ACTIVE, -1, -1, lane); return; }  // per-query budget spent: "Timeout"
    __syncwarp();
    if (lane == 0) r.rounds_used[w] = used + 1;
  }
  const int phase = r.phase[w];
  int T = r.tsel[w];  // phase 3: the tree of the extend edge; phase 2: the connecting tree
  bool new_iter = phase == 0;
  if (phase != 0) {
    // ---- verdicts of the pending sub-edges: sub-edge k is invalid if any of its samples collides
    const int ns_all = r.n_steps[w];
    unsigned bad = 0;
    for (int k = lane; k < ns_all * r.smax; k += 32) {
      const unsigned s = (unsigned)w * S * r.smax + k;
      if ((verdict_bits[s >> 5] >> (s & 31)) & 1u) bad |= 1u << (k / r.smax);
    }
    bad = __reduce_or_sync(FULL, bad);
    int first_slot = 0, calls = 0;
    bool go = true;
    if (phase == 3) {
      calls = 1;
      if (bad & 1u) { go = false; new_iter = true; }  // the extend is TRAPPED: the speculative connect steps are void
      else {
        const int n = r.size[w * 2 + T];
        if (n + 1 > r.cap) { rrt_finish(r, w, q, MPB_RRT_CAPACITY, -1, -1, lane); return; }
        float *node = r.nodes + (((size_t)w * 2 + T) * r.cap + n) * D;
        for (int j = lane; j < D; j += 32) node[j] = eb[j];
        if (lane == 0) { r.parent[((size_t)w * 2 + T) * r.cap + n] = r.nearest[w]; r.size[w * 2 + T] = n + 1; r.added[w] = n; }
        T = 1 - T;  // the connect loop grows the other tree
        first_slot = 1;
        bad >>= 1;
      }
    }
    if (go) {
      const int ns = ns_all - first_slot;
      const int n_add = bad ? (__ffs(bad) - 1) : ns;  // valid prefix of the connect walk
      calls += n_add + (bad ? 1 : 0);
      const int n = r.size[w * 2 + T];
      if (n + n_add > r.cap) { rrt_finish(r, w, q, MPB_RRT_CAPACITY, -1, -1, lane); return; }
      float *node = r.nodes + (((size_t)w * 2 + T) * r.cap + n) * D;
      for (int j = lane; j < n_add * D; j += 32) node[j] = eb[first_slot * D + j];
      const int near_c = phase == 3 ? r.nearest2[w] : r.nearest[w];
      for (int k = lane; k < n_add; k += 32) r.parent[((size_t)w * 2 + T) * r.cap + n + k] = k == 0 ? near_c : n + k - 1;
      if (lane == 0 && n_add) r.size[w * 2 + T] = n + n_add;
      __syncwarp();
      if (n_add == ns && r.reach[w]) {  // REACHED: the trees meet
        const int m_first = r.added[w], m_other = n + n_add - 1;  // first tree = 1 - T
        int is = T == 0 ? m_other : m_first, ig = T == 0 ? m_first : m_other;
        const int ps = r.parent[((size_t)w * 2 + 0) * r.cap + is];
        if (ps >= 0) is = ps;  // one step back so the shared state is not duplicated on the path
        else ig = r.parent[((size_t)w * 2 + 1) * r.cap + ig];
        if (lane == 0) r.iters[w] += calls;
        __syncwarp();
        rrt_finish(r, w, q, MPB_RRT_SOLVED, is, ig, lane);
        return;
      }
      if (n_add < ns) new_iter = true;  // ADVANCED n_add times, then TRAPPED
    }
    if (lane == 0) r.iters[w] += calls;
    __syncwarp();  // the sampling counter below reads it
  }
  if (new_iter) {
    // ---- head of the solve loop: pick the tree, maybe admit another goal, sample
    const int fs = r.flag_start[w];
    T = fs ? 0 : 1;
    const int gsz = r.size[w * 2 + 1], used = r.goals_used[w];
    if ((gsz == 0 || used < gsz / 2) && used < r.n_goals[q]) {
      if (gsz >= r.cap) { rrt_finish(r, w, q, MPB_RRT_CAPACITY, -1, -1, lane); return; }
      float *node = r.nodes + (((size_t)w * 2 + 1) * r.cap + gsz) * D;
      const float *g = r.goals + ((size_t)q * r.max_goals + used) * D;
      for (int j = lane; j < D; j += 32) node[j] = g[j];
      if (lane == 0) { r.parent[((size_t)w * 2 + 1) * r.cap + gsz] = -1; r.size[w * 2 + 1] = gsz + 1; r.goals_used[w] = used + 1; }
    } else if (gsz == 0) { rrt_finish(r, w, q, MPB_RRT_INVALID_GOAL, -1, -1, lane); return; }
    const unsigned ctr = (unsigned)r.iters[w];
    for (int j = lane; j < D; j += 32) tg[j] = fmaf(sp.hi[j] - sp.lo[j], rrt_uniform(r.seed, (unsigned)q, ctr, (unsigned)j), sp.lo[j]);
    __syncwarp();
    // ---- growTree(tree, rstate): the extend edge into slot 0 ...
    const float *base = r.nodes + ((size_t)w * 2 + T) * r.cap * D;
    float best;
    const int bi = rrt_nearest(base, r.size[w * 2 + T], D, tg, sp, lane, &best);
    bool reach0;
    rrt_emit_walk(ea, eb, 0, 1, 1, base + (size_t)bi * D, best, tg, range, D, sp, lane, &reach0);
    __syncwarp();
    for (int j = lane; j < D; j += 32) tg[j] = eb[j];  // q_new: the target of the connect loop
    __syncwarp();
    // ---- ... and the first steps of the connect walk of the other tree towards q_new into slots 1..
    const float *base2 = r.nodes + ((size_t)w * 2 + (1 - T)) * r.cap * D;
    float best2;
    const int bi2 = rrt_nearest(base2, r.size[w * 2 + (1 - T)], D, tg, sp, lane, &best2);
    bool reach2;
    const int ns2 = rrt_emit_walk(ea, eb, 1, S, S - 1, base2 + (size_t)bi2 * D, best2, tg, range, D, sp, lane, &reach2);
    if (lane == 0) {
      r.flag_start[w] = !fs; r.tsel[w] = T; r.phase[w] = 3;
      r.nearest[w] = bi; r.nearest2[w] = bi2; r.reach[w] = reach2 ? 1 : 0; r.n_steps[w] = 1 + ns2;
    }
    return;
  }
  // ---- the connect walk continues (phase 2): nearest neighbour of the connecting tree and the next S steps
  const float *base = r.nodes + ((size_t)w * 2 + T) * r.cap * D;
  float best;
  const int bi = rrt_nearest(base, r.size[w * 2 + T], D, tg, sp, lane, &best);
  bool reach;
  const int ns = rrt_emit_walk(ea, eb, 0, S, S, base + (size_t)bi * D, best, tg, range, D, sp, lane, &reach);
  if (lane == 0) { r.tsel[w] = T; r.phase[w] = 2; r.nearest[w] = bi; r.reach[w] = reach ? 1 : 0; r.n_steps[w] = ns; }
}

// End of the call: queries still in a slot keep status MPB_RRT_ACTIVE ("Timeout"); their growTree count is recorded.
__global__ void rrt_flush_kernel(RrtState r) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= r.n_slots) return;
  const int q = r.slot_query[w];
  if (q >= 0) r.q_iters[q] = r.iters[w];
}

// [Q][max_path][D] -> packed rows: query q owns rows off[q] .. off[q+1] of `packed`
__global__ void pack_paths_kernel(const float *__restrict__ paths, const int *__restrict__ off, int D, int max_path,
                                  float *__restrict__ packed) {
  const int q = blockIdx.x;
  const int o = off[q], n = (off[q + 1] - o) * D;
  const float *src = paths + (size_t)q * max_path * D;
  for (int k = threadIdx.x; k < n; k += blockDim.x) packed[(size_t)o * D + k] = src[k];
}
