// mpb.cu -- host side of libmpb.so: scene image compiler (double -> fp32 device tables, bounding
// volumes, bin classes), work-queue management, and the extern "C" entry points of include/mpb.h.
// There is no CPU compute path here: every batch entry point launches the sm_100a kernels of
// mpb_kernels.cuh or fails with MPB_ERR_CUDA.
#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "mpb_kernels.cuh"
#include "mpb_plan.cuh"

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) {
  g_err = msg;
  return code;
}
#define CU(call)                                                                                      \
  do {                                                                                                \
    cudaError_t e_ = (call);                                                                          \
    if (e_ != cudaSuccess)                                                                            \
      return fail(MPB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));                  \
  } while (0)

#define MPB_ASYNC_SLOTS 16
static const long long kMaxHitPasses = 4096;
struct AsyncSlot {
  unsigned *h = nullptr;          // pinned: totals[8] | counters[4] (64-bit)
  cudaEvent_t done = nullptr;
  std::vector<cudaEvent_t> ev;    // timing events, 5 per pass (only with mpb_scene_set_async_timing)
  size_t n_ev = 0;
  bool pending = false;
  long long n = 0, launches = 0;
  float host_ms = 0.f;
};
struct SmallGraph {
  long long n = 0;
  int mode = 0;                   // 0 verdicts, 1 + dense pair mask, 2 + hit list
  unsigned hit_cap = 0;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  long long launches = 0;
  // what the captured nodes point at; a graph is dropped when any of these moved
  const void *d_q = nullptr, *d_v = nullptr, *d_p = nullptr, *d_h = nullptr, *h_base = nullptr;
  long long chunk_at = 0;
  unsigned long long img_version = 0;
};

// cudaFuncAttributeMaxDynamicSharedMemorySize is per-function state shared by every scene of the process: only ever
// raise it (a second scene with a smaller hull pool must not lower the limit under a scene that needs more).
static cudaError_t raise_dyn_smem(const void *func, int device, int slot, size_t bytes) {
  static int limit[4][64] = {{0}};
  if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
  if ((int)bytes <= limit[slot][device]) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e == cudaSuccess) limit[slot][device] = (int)bytes;
  return e;
}

struct mpb_scene {
  int device = 0;
  int num_sms = 0;
  cudaStream_t stream = nullptr;
  // host copies (double) kept for set_q_base / set_pairs
  std::vector<DevJoint> h_joints;
  std::vector<DevObject> h_objs;
  std::vector<DevPair> h_pairs;
  std::vector<int> joint_qidx;
  // double-precision copy of the kinematic model for the IK entry point
  std::vector<double> h_jP64, h_jax64, h_qbase64, h_linkL64;
  std::vector<int> h_jtype, h_jparent, h_linkJoint;
  std::vector<int> obj_key, h_dynlist, h_objLink;
  std::vector<float4> h_cobb;
  std::vector<int32_t> h_pair_a, h_pair_b;  // the admitted pair list as given (build_pairs re-orders BVH pairs)
  size_t work_P = 0;                        // pair count the work queues were sized for
  int nq_full = 0;
  // device image
  DevJoint *d_joints = nullptr;
  DevObject *d_objs = nullptr;
  DevPair *d_pairs = nullptr;
  float4 *d_verts = nullptr;
  float *d_linkL = nullptr;
  int *d_linkJoint = nullptr;
  int *d_dynlist = nullptr;
  DevBvhNode *d_nodes = nullptr;
  float4 *d_prims = nullptr;
  int *d_pair_bin2 = nullptr;
  size_t pair_bin2_cap = 0;
  std::vector<int> h_pair_bin2;
  // cull table (sphere_kernel): pairs re-ordered into runs that share their moving object
  std::vector<float4> h_cullc, h_cullq;
  std::vector<int2> h_cullm;
  std::vector<int4> h_cullp;  // per cull entry: {pair, object a, object b, (slot a + 1) | (slot b + 1) << 16}: one load for the dense cull
  std::vector<int4> h_runs;
  float4 *d_cobb = nullptr;
  float4 *d_cullc = nullptr, *d_cullq = nullptr;
  int2 *d_cullm = nullptr;
  int4 *d_cullp = nullptr;
  int4 *d_runs = nullptr;
  size_t cull_cap = 0;
  SceneImg img{};
  // sphere_kernel's cull table in parameter space when it has at most 256 entries (4 KB; a 28 KB table for the C3 scene
  // thrashed the constant cache: cull leg 12.7 -> 15.5 ms); tier 0 = global tables
  typedef CullTabC<256, 64> CullSmall;
  CullSmall ct_small{};
  int cull_tier = 0;
  FkTab fk_tab{};            // joint / moving-object records of fk_table_kernel, passed by value (constant bank)
  float *d_pose = nullptr;   // sphere table [n_dyn * 3 + n_jslots * 12][pose_ld]
  float4 *d_obb = nullptr;   // OBB table [n_dyn][pose_ld][2]: unit quaternion xyzw, translation
  size_t pose_ld = 0;
  uint2 *d_sq = nullptr;     // survivor queue of the sphere phase: (configuration, cull-table entry)
  unsigned sq_cap = 0;
  size_t pairs_cap = 0;
  // work buffers
  long long chunk = 0;
  WorkQ wq{};
  // host-buffer calls: double-buffered device copies so H2D / kernels / D2H of consecutive chunks overlap
  float *d_q[2] = {nullptr, nullptr};
  unsigned int *d_verdict[2] = {nullptr, nullptr};
  unsigned int *d_pairbits[2] = {nullptr, nullptr};
  float *d_poses = nullptr;
  size_t d_q_cap[2] = {0, 0}, d_verdict_cap[2] = {0, 0}, d_pairbits_cap[2] = {0, 0}, d_poses_cap = 0, bins_cap = 0;
  cudaStream_t stream_h2d = nullptr, stream_d2h = nullptr;
  cudaStream_t stream_hits = nullptr;  // hit-list shares: not ordered behind stream_d2h's wait for the NEXT pass's kernels
  std::vector<cudaEvent_t> sync_events;
  long long host_chunk = 0;
  bool smem_verts = true;
  int narrow_blocks = 0;
  int sphere_blocks[2] = {0, 0}, dense_blocks[2] = {0, 0}, sphere_blocks_ct[1] = {0}, sphere_blocks_blk[2] = {0, 0};
  int expand_lane_blocks = 0, dense2_blocks = 0;
  // fk_sphere_kernel (FK + sphere phase fused, OBB rows on demand): branch-point joint slots kept on chip
  unsigned jneed = 0;
  size_t fk_sphere_smem = 0;
  int fk_sphere_blocks = 0;
  // broadphase grid over world-fixed partners (sphere_grid_kernel), built lazily before the first large pass
  bool grid_dirty = true, grid_on = false;
  int *d_grid_start = nullptr;
  unsigned short *d_grid_list = nullptr;
  int2 *d_grid_run = nullptr;
  size_t grid_start_cap = 0, grid_list_cap = 0, grid_run_cap = 0;
  int sphere_blocks_grid = 0;
  size_t narrow2_smem = 0;
  int narrow2_blocks = 0;
  bool has_primitives = false;  // some object is a sphere / box / capsule / cylinder or an octree (sufficient-hit test worthwhile)
  int sphere_cb = 4;         // tiles of 32 configurations per warp of sphere_blk_kernel (large cull tables)  // one resident wave of the static grid-stride kernels, per MODE
  std::vector<cudaEvent_t> events;
  // scratch blocks of the planning / distance entry points, kept between calls (cudaMalloc + cudaFree of the tree
  // buffers cost more than a thousand planning rounds)
  std::vector<std::pair<void *, size_t>> pool;
  // blocking calls read their queue counters through this pinned block
  unsigned *h_status = nullptr;  // [32]
  unsigned *h_hitcnt = nullptr;  // [kMaxHitPasses] hit-list length after each pass of a host-buffer call (same pinned block)
  // sparse hit list of host-buffer calls: [0] = counter, [1..] = (configuration, pair) entries
  uint2 *d_hits = nullptr;
  size_t d_hits_cap = 0;
  // device-pointer calls in flight (see collide_enqueue)
  AsyncSlot slots[MPB_ASYNC_SLOTS];
  unsigned *h_slots = nullptr;  // pinned, 32 words per slot
  int slot_head = 0;
  mpb_stats async_acc{};
  long long async_completed = 0, async_failed = 0;
  bool calibrated = false, async_timing = false;
  long long good_chunk = 0, cap_chunk = 1ll << 40;  // largest pass size that fitted / smallest that is still allowed
  // small host batches replayed as CUDA graphs (see collide_small)
  std::vector<SmallGraph> small_graphs;
  unsigned char *h_small = nullptr, *d_small = nullptr;  // pinned staging / device output block
  size_t h_small_cap = 0, d_small_cap = 0;
  unsigned long long img_version = 0;  // bumped whenever a by-value kernel parameter of the image changes
  float *d_qconst = nullptr;           // [n_joints] value of the joints the batch state does not write (q_base)
};

static void drop_small_graphs(mpb_scene *s);
static int harvest(mpb_scene *s, bool wait);
// Wait for everything this scene has in flight (its own streams and the device-pointer calls on caller streams).  Never
// cudaDeviceSynchronize: another host thread may be capturing a graph on another scene's stream.
static cudaError_t sync_scene(mpb_scene *s);

static cudaError_t sync_scene(mpb_scene *s) {
  cudaError_t e = cudaSuccess;
  for (auto &sl : s->slots)
    if (sl.pending && sl.done && (e = cudaEventSynchronize(sl.done)) != cudaSuccess) return e;
  if (s->stream && (e = cudaStreamSynchronize(s->stream)) != cudaSuccess) return e;
  if (s->stream_h2d && (e = cudaStreamSynchronize(s->stream_h2d)) != cudaSuccess) return e;
  if (s->stream_d2h && (e = cudaStreamSynchronize(s->stream_d2h)) != cudaSuccess) return e;
  if (s->stream_hits && (e = cudaStreamSynchronize(s->stream_hits)) != cudaSuccess) return e;
  return e;
}

extern "C" const char *mpb_last_error(void) { return g_err.c_str(); }

extern "C" int mpb_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

// ------------------------------------------------------------------------------------------ host geometry helpers
static void mul12(const double *a, const double *b, double *o) {  // 3x4 * 3x4 (implicit last row)
  double t[12];
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) t[4 * r + c] = a[4 * r] * b[c] + a[4 * r + 1] * b[4 + c] + a[4 * r + 2] * b[8 + c];
    t[4 * r + 3] = a[4 * r] * b[3] + a[4 * r + 1] * b[7] + a[4 * r + 2] * b[11] + a[4 * r + 3];
  }
  memcpy(o, t, sizeof(t));
}

// Jacobi eigen-decomposition of a symmetric 3x3; columns of V are eigenvectors.
static void jacobi3(double A[3][3], double V[3][3]) {
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) V[i][j] = i == j;
  for (int sweep = 0; sweep < 32; sweep++) {
    double off = fabs(A[0][1]) + fabs(A[0][2]) + fabs(A[1][2]);
    if (off < 1e-18) break;
    for (int p = 0; p < 2; p++)
      for (int q = p + 1; q < 3; q++) {
        if (fabs(A[p][q]) < 1e-300) continue;
        double th = (A[q][q] - A[p][p]) / (2 * A[p][q]);
        double t = (th >= 0 ? 1.0 : -1.0) / (fabs(th) + sqrt(th * th + 1));
        double c = 1 / sqrt(t * t + 1), s = t * c;
        for (int k = 0; k < 3; k++) {
          double akp = A[k][p], akq = A[k][q];
          A[k][p] = c * akp - s * akq;
          A[k][q] = s * akp + c * akq;
        }
        for (int k = 0; k < 3; k++) {
          double apk = A[p][k], aqk = A[q][k];
          A[p][k] = c * apk - s * aqk;
          A[q][k] = s * apk + c * aqk;
        }
        for (int k = 0; k < 3; k++) {
          double vkp = V[k][p], vkq = V[k][q];
          V[k][p] = c * vkp - s * vkq;
          V[k][q] = s * vkp + c * vkq;
        }
      }
  }
}

static void fit_box(const double *v, int n, const double R[3][3], double *c, double *h) {
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (int i = 0; i < n; i++)
    for (int a = 0; a < 3; a++) {
      double t = v[3 * i] * R[0][a] + v[3 * i + 1] * R[1][a] + v[3 * i + 2] * R[2][a];
      lo[a] = std::min(lo[a], t);
      hi[a] = std::max(hi[a], t);
    }
  double m[3];
  for (int a = 0; a < 3; a++) {
    m[a] = 0.5 * (lo[a] + hi[a]);
    h[a] = 0.5 * (hi[a] - lo[a]);
  }
  for (int r = 0; r < 3; r++) c[r] = R[r][0] * m[0] + R[r][1] * m[1] + R[r][2] * m[2];
}

// bounding sphere + OBB of a vertex cloud (object frame)
static void bound_points(const double *v, int n, DevObject &O) {
  double mean[3] = {0, 0, 0};
  for (int i = 0; i < n; i++)
    for (int a = 0; a < 3; a++) mean[a] += v[3 * i + a];
  for (int a = 0; a < 3; a++) mean[a] /= std::max(n, 1);
  double C[3][3] = {{0}};
  for (int i = 0; i < n; i++) {
    double d[3] = {v[3 * i] - mean[0], v[3 * i + 1] - mean[1], v[3 * i + 2] - mean[2]};
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++) C[a][b] += d[a] * d[b];
  }
  double V[3][3];
  jacobi3(C, V);
  // make V a proper rotation
  double det = V[0][0] * (V[1][1] * V[2][2] - V[1][2] * V[2][1]) - V[0][1] * (V[1][0] * V[2][2] - V[1][2] * V[2][0]) +
               V[0][2] * (V[1][0] * V[2][1] - V[1][1] * V[2][0]);
  if (det < 0)
    for (int r = 0; r < 3; r++) V[r][2] = -V[r][2];
  const double I[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  double c1[3], h1[3], c2[3], h2[3];
  fit_box(v, n, V, c1, h1);
  fit_box(v, n, I, c2, h2);
  const bool use_pca = h1[0] * h1[1] * h1[2] <= h2[0] * h2[1] * h2[2];
  const double(*R)[3] = use_pca ? V : I;
  const double *c = use_pca ? c1 : c2, *h = use_pca ? h1 : h2;
  const double pad = 1e-6;  // fp32 rounding of the stored box
  for (int a = 0; a < 3; a++) {
    O.oc[a] = (float)c[a];
    O.oh[a] = (float)(h[a] + pad);
    for (int b = 0; b < 3; b++) O.oR[3 * a + b] = (float)R[a][b];
  }
  // bounding sphere around the box centre
  double r2 = 0;
  for (int i = 0; i < n; i++) {
    double d[3] = {v[3 * i] - c[0], v[3 * i + 1] - c[1], v[3 * i + 2] - c[2]};
    r2 = std::max(r2, d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  }
  for (int a = 0; a < 3; a++) O.bc[a] = (float)c[a];
  O.br = (float)(sqrt(r2) + pad);
}

static void bound_box(double hx, double hy, double hz, DevObject &O) {
  const double pad = 1e-6;
  for (int a = 0; a < 3; a++) O.oc[a] = O.bc[a] = 0.f;
  O.oh[0] = (float)(hx + pad);
  O.oh[1] = (float)(hy + pad);
  O.oh[2] = (float)(hz + pad);
  for (int a = 0; a < 9; a++) O.oR[a] = (a % 4 == 0) ? 1.f : 0.f;
  O.br = (float)(sqrt(hx * hx + hy * hy + hz * hz) + pad);
}

// ------------------------------------------------------------------------------------------ BVH over leaf primitives
struct HostPrim {
  float4 a, b, c;      // TRI: vertices; VBOX: centre, half extents, unused
  double lo[3], hi[3];
};

static int build_bvh(std::vector<HostPrim> &prims, int first, int count, std::vector<DevBvhNode> &nodes) {
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  double clo[3] = {1e300, 1e300, 1e300}, chi[3] = {-1e300, -1e300, -1e300};
  for (int i = first; i < first + count; i++)
    for (int a = 0; a < 3; a++) {
      lo[a] = std::min(lo[a], prims[i].lo[a]);
      hi[a] = std::max(hi[a], prims[i].hi[a]);
      const double cc = 0.5 * (prims[i].lo[a] + prims[i].hi[a]);
      clo[a] = std::min(clo[a], cc);
      chi[a] = std::max(chi[a], cc);
    }
  const int me = (int)nodes.size();
  nodes.push_back(DevBvhNode{});
  for (int a = 0; a < 3; a++) {
    nodes[me].c[a] = (float)(0.5 * (lo[a] + hi[a]));
    nodes[me].h[a] = (float)(0.5 * (hi[a] - lo[a]) + 1e-6);
  }
  if (count <= 2) {
    nodes[me].left = ~first;
    nodes[me].right = count;
    return me;
  }
  int ax = 0;
  if (chi[1] - clo[1] > chi[ax] - clo[ax]) ax = 1;
  if (chi[2] - clo[2] > chi[ax] - clo[ax]) ax = 2;
  const int mid = first + count / 2;
  std::nth_element(prims.begin() + first, prims.begin() + mid, prims.begin() + first + count,
                   [ax](const HostPrim &x, const HostPrim &y) { return x.lo[ax] + x.hi[ax] < y.lo[ax] + y.hi[ax]; });
  const int l = build_bvh(prims, first, mid - first, nodes);
  const int r = build_bvh(prims, mid, first + count - mid, nodes);
  nodes[me].left = l;
  nodes[me].right = r;
  return me;
}

// Occupied cells -> exact cover by axis-aligned boxes: runs along x, equal runs merged along y, then z.
// The union of the boxes equals the union of the cells (what octomap's pruning also preserves).
struct CellBox { int lo[3], hi[3]; };  // inclusive cell ranges
static std::vector<CellBox> merge_cells(const int32_t *cells, int n) {
  std::vector<std::array<int, 3>> c(n);
  for (int i = 0; i < n; i++) c[i] = {cells[3 * i + 2], cells[3 * i + 1], cells[3 * i]};  // sort key (z, y, x)
  std::sort(c.begin(), c.end());
  c.erase(std::unique(c.begin(), c.end()), c.end());
  std::vector<CellBox> runs;
  for (size_t i = 0; i < c.size();) {
    size_t j = i + 1;
    while (j < c.size() && c[j][0] == c[i][0] && c[j][1] == c[i][1] && c[j][2] == c[j - 1][2] + 1) j++;
    runs.push_back(CellBox{{c[i][2], c[i][1], c[i][0]}, {c[j - 1][2], c[i][1], c[i][0]}});
    i = j;
  }
  for (int axis = 1; axis <= 2; axis++) {  // merge boxes identical in the other two axes and adjacent along `axis`
    const int o1 = axis == 1 ? 0 : 0, o2 = axis == 1 ? 2 : 1;
    std::sort(runs.begin(), runs.end(), [&](const CellBox &x, const CellBox &y) {
      if (x.lo[o1] != y.lo[o1]) return x.lo[o1] < y.lo[o1];
      if (x.hi[o1] != y.hi[o1]) return x.hi[o1] < y.hi[o1];
      if (x.lo[o2] != y.lo[o2]) return x.lo[o2] < y.lo[o2];
      if (x.hi[o2] != y.hi[o2]) return x.hi[o2] < y.hi[o2];
      return x.lo[axis] < y.lo[axis];
    });
    std::vector<CellBox> out;
    for (const CellBox &b : runs) {
      if (!out.empty()) {
        CellBox &l = out.back();
        if (l.lo[o1] == b.lo[o1] && l.hi[o1] == b.hi[o1] && l.lo[o2] == b.lo[o2] && l.hi[o2] == b.hi[o2] &&
            l.hi[axis] + 1 == b.lo[axis]) {
          l.hi[axis] = b.hi[axis];
          continue;
        }
      }
      out.push_back(b);
    }
    runs.swap(out);
  }
  return runs;
}

// Support cell table of one hull (layout in mpb_device.cuh).  Vertex i can be extreme for a direction of cell C only
// if, for every other vertex j, SOME direction of C prefers i: max over C of d . (v_i - v_j) >= 0.  The cell is a
// planar quad on the cube (directions need not be unit length), so that maximum sits at one of its four corners.
// Checking the condition against all j gives a superset of the exact set; the quad is widened by `margin` so that
// the device's float arithmetic may place a boundary direction in either neighbouring cell.
static void build_cell_table_uncached(const double *V, int nv, std::vector<unsigned char> &out);
// Process-wide cache keyed by the vertex bytes: re-compiling a scene after an object was added or removed (or compiling a
// second scene for the same robot) does not redo the O(cells x nv^2) table of geometry it has seen.
static void build_cell_table(const double *V, int nv, std::vector<unsigned char> &out) {
  static std::mutex mu;
  static std::map<std::string, std::vector<unsigned char>> cache;
  const std::string key(reinterpret_cast<const char *>(V), (size_t)nv * 3 * sizeof(double));
  {
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find(key);
    if (it != cache.end()) { out = it->second; return; }
  }
  build_cell_table_uncached(V, nv, out);
  std::lock_guard<std::mutex> lk(mu);
  if (cache.size() < 4096) cache[key] = out;
}
// Candidates of a cell = the vertices whose NORMAL CONE meets the cell: vertex i is the support point for a direction d iff
// d . (v_i - v_j) >= 0 for every j.  On a cube-map face d = (+-1, u, v) (permuted), so every j is a half-plane in (u, v)
// and the question is whether the cell's rectangle, clipped by all of them, stays non-empty (Sutherland-Hodgman on a
// polygon of a handful of vertices).  A cheap necessary condition runs first (for every j SOME corner of the cell prefers i
// to j; the first version stopped there): it already leaves ~3.1 candidates per cell for the Panda's hulls, the cones
// 2.9.  The half-planes are relaxed by 1e-6 (relative), so a vertex that loses to the winner by less than fp32 rounding
// is still listed and ties resolve as in the full scan.
static void build_cell_table_uncached(const double *V, int nv, std::vector<unsigned char> &out) {
  const int G = MPB_CELL_G;
  const double margin = 1e-3;
  std::vector<unsigned short> first(MPB_CELL_COUNT + 1, 0);
  std::vector<unsigned char> cand;
  std::vector<std::array<double, 2>> poly, next;
  for (int face = 0; face < 6; face++) {
    const double sgn = (face & 1) ? -1.0 : 1.0;
    const int am = face < 2 ? 0 : (face < 4 ? 1 : 2);          // dominant axis
    const int au = face < 2 ? 1 : 0, av = face < 4 ? 2 : 1;    // axes of u, v (mpb_cube_cell)
    for (int iv = 0; iv < G; iv++)
      for (int iu = 0; iu < G; iu++) {
        const double u0 = -1.0 + 2.0 * iu / G - margin, u1 = -1.0 + 2.0 * (iu + 1) / G + margin;
        const double v0 = -1.0 + 2.0 * iv / G - margin, v1 = -1.0 + 2.0 * (iv + 1) / G + margin;
        first[(face * G + iv) * G + iu] = (unsigned short)cand.size();
        const double cu[4] = {u0, u1, u0, u1}, cv[4] = {v0, v0, v1, v1};
        for (int i = 0; i < nv; i++) {
          bool maybe = true;
          for (int j = 0; j < nv && maybe; j++) {
            if (j == i) continue;
            const double w[3] = {V[3 * i] - V[3 * j], V[3 * i + 1] - V[3 * j + 1], V[3 * i + 2] - V[3 * j + 2]};
            const double a = w[au], b = w[av], c = sgn * w[am];
            double best = -1e300;
            for (int k = 0; k < 4; k++) best = std::max(best, a * cu[k] + b * cv[k] + c);
            maybe = best >= -(1e-6 * (std::fabs(a) + std::fabs(b) + std::fabs(c)) + 1e-300);
          }
          if (!maybe) continue;
          poly = {{u0, v0}, {u1, v0}, {u1, v1}, {u0, v1}};
          for (int j = 0; j < nv && !poly.empty(); j++) {
            if (j == i) continue;
            const double w[3] = {V[3 * i] - V[3 * j], V[3 * i + 1] - V[3 * j + 1], V[3 * i + 2] - V[3 * j + 2]};
            const double a = w[au], b = w[av], c = sgn * w[am];
            const double tol = 1e-6 * (std::fabs(a) + std::fabs(b) + std::fabs(c)) + 1e-300;
            // keep the part of the polygon with a u + b v + c >= -tol
            next.clear();
            const size_t n = poly.size();
            for (size_t k = 0; k < n; k++) {
              const auto &P = poly[k], &Q = poly[(k + 1) % n];
              const double fp = a * P[0] + b * P[1] + c + tol, fq = a * Q[0] + b * Q[1] + c + tol;
              if (fp >= 0) next.push_back(P);
              if ((fp >= 0) != (fq >= 0)) {
                const double t = fp / (fp - fq);
                next.push_back({P[0] + t * (Q[0] - P[0]), P[1] + t * (Q[1] - P[1])});
              }
            }
            poly.swap(next);
          }
          if (!poly.empty()) cand.push_back((unsigned char)i);
        }
        // some vertex is extreme for every direction, so a cell is never empty -- unless the input is degenerate (duplicated
        // vertices exclude each other only through the tolerance): then the cell falls back to the full scan
        if (cand.size() == first[(face * G + iv) * G + iu])
          for (int i = 0; i < nv; i++) cand.push_back((unsigned char)i);
      }
  }
  first[MPB_CELL_COUNT] = (unsigned short)cand.size();
  out.resize(2 * (MPB_CELL_COUNT + 1) + cand.size());
  memcpy(out.data(), first.data(), 2 * (MPB_CELL_COUNT + 1));
  memcpy(out.data() + 2 * (MPB_CELL_COUNT + 1), cand.data(), cand.size());
}

// ---- inner balls: up to MPB_INNER_BALLS balls that lie inside a shape (object frame).  If a ball of A intersects B
// (exactly, for a primitive B; or a ball of B otherwise) the two shapes intersect: scatter_kernel settles such items
// without the narrowphase.  Radii are shrunk by 2 % so that fp32 rounding of the poses cannot turn "inside" into "outside".
static void set_ball(DevObject &O, int k, double x, double y, double z, double r) {
  O.ball[k] = make_float4((float)x, (float)y, (float)z, r > 0 ? (float)(0.98 * r) : 0.f);
}
static void primitive_balls(DevObject &O) {
  for (int k = 0; k < MPB_INNER_BALLS; k++) O.ball[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  switch (O.type) {
    case MPB_GEOM_SPHERE: set_ball(O, 0, 0, 0, 0, O.prm[0]); break;
    case MPB_GEOM_BOX: {  // balls of the smallest half side along the longest axis
      const double h[3] = {O.prm[0], O.prm[1], O.prm[2]};
      int ax = 0;
      if (h[1] > h[ax]) ax = 1;
      if (h[2] > h[ax]) ax = 2;
      const double r = std::min(h[0], std::min(h[1], h[2])), span = h[ax] - r;
      for (int k = 0; k < MPB_INNER_BALLS; k++) {
        double c[3] = {0, 0, 0};
        c[ax] = MPB_INNER_BALLS > 1 ? -span + 2.0 * span * k / (MPB_INNER_BALLS - 1) : 0.0;
        set_ball(O, k, c[0], c[1], c[2], r);
      }
    } break;
    case MPB_GEOM_CAPSULE:
      for (int k = 0; k < MPB_INNER_BALLS; k++) set_ball(O, k, 0, 0, -O.prm[1] + 2.0 * O.prm[1] * k / (MPB_INNER_BALLS - 1), O.prm[0]);
      break;
    case MPB_GEOM_CYLINDER: {
      const double r = std::min(O.prm[0], O.prm[1]), span = O.prm[1] - r;
      for (int k = 0; k < MPB_INNER_BALLS; k++) set_ball(O, k, 0, 0, -span + 2.0 * span * k / (MPB_INNER_BALLS - 1), r);
    } break;
    default: break;
  }
}
// convex hull (vertices V, triangle faces F): the signed distance to the nearest face plane on a grid over the bounding box;
// greedy choice of the largest ball, then of the largest balls whose centres lie outside the balls chosen so far
static void hull_balls(const double *V, int nv, const int32_t *F, int nf, float4 *out) {
  for (int k = 0; k < MPB_INNER_BALLS; k++) out[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (nv < 4 || nf < 4) return;
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300}, ctr[3] = {0, 0, 0};
  for (int i = 0; i < nv; i++)
    for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], V[3 * i + a]); hi[a] = std::max(hi[a], V[3 * i + a]); ctr[a] += V[3 * i + a] / nv; }
  std::vector<std::array<double, 4>> planes;  // outward unit normal, offset: n.x <= d inside
  for (int f = 0; f < nf; f++) {
    const double *A0 = V + 3 * F[3 * f], *A1 = V + 3 * F[3 * f + 1], *A2 = V + 3 * F[3 * f + 2];
    double e1[3], e2[3], n[3];
    for (int a = 0; a < 3; a++) { e1[a] = A1[a] - A0[a]; e2[a] = A2[a] - A0[a]; }
    n[0] = e1[1] * e2[2] - e1[2] * e2[1]; n[1] = e1[2] * e2[0] - e1[0] * e2[2]; n[2] = e1[0] * e2[1] - e1[1] * e2[0];
    const double l = sqrt(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
    if (l < 1e-14) continue;
    for (int a = 0; a < 3; a++) n[a] /= l;
    double d = n[0] * A0[0] + n[1] * A0[1] + n[2] * A0[2];
    if (n[0] * ctr[0] + n[1] * ctr[1] + n[2] * ctr[2] > d) { for (int a = 0; a < 3; a++) n[a] = -n[a]; d = -d; }  // orient away from the interior point
    planes.push_back({n[0], n[1], n[2], d});
  }
  const int G = 14;
  std::vector<std::array<double, 4>> cand;
  for (int ix = 0; ix < G; ix++)
    for (int iy = 0; iy < G; iy++)
      for (int iz = 0; iz < G; iz++) {
        const double x[3] = {lo[0] + (hi[0] - lo[0]) * (ix + 0.5) / G, lo[1] + (hi[1] - lo[1]) * (iy + 0.5) / G, lo[2] + (hi[2] - lo[2]) * (iz + 0.5) / G};
        double r = 1e300;
        for (const auto &pl : planes) r = std::min(r, pl[3] - (pl[0] * x[0] + pl[1] * x[1] + pl[2] * x[2]));
        if (r > 0) cand.push_back({x[0], x[1], x[2], r});
      }
  std::sort(cand.begin(), cand.end(), [](const std::array<double, 4> &a, const std::array<double, 4> &b) { return a[3] > b[3]; });
  int n_out = 0;
  for (const auto &c : cand) {
    if (n_out == MPB_INNER_BALLS) break;
    if (n_out > 0 && c[3] < 0.35 * out[0].w) break;  // small balls settle nothing the large ones do not
    bool ok = true;
    for (int k = 0; k < n_out && ok; k++) {
      const double dx = c[0] - out[k].x, dy = c[1] - out[k].y, dz = c[2] - out[k].z;
      ok = sqrt(dx * dx + dy * dy + dz * dz) >= 0.9 * out[k].w;
    }
    if (ok) out[n_out++] = make_float4((float)c[0], (float)c[1], (float)c[2], (float)(0.98 * c[3]));
  }
}

// parameters, OBB, bounding sphere and inscribed ball of a primitive (object frame); p = obj_param of include/mpb.h
static void primitive_bounds(DevObject &O, const double *p) {
  switch (O.type) {
    case MPB_GEOM_SPHERE:
      O.prm[0] = (float)p[0];
      bound_box(p[0], p[0], p[0], O);
      O.br = (float)(p[0] + 1e-6);
      O.rin = O.prm[0];
      break;
    case MPB_GEOM_BOX:
      O.prm[0] = (float)(0.5 * p[0]); O.prm[1] = (float)(0.5 * p[1]); O.prm[2] = (float)(0.5 * p[2]);
      bound_box(0.5 * p[0], 0.5 * p[1], 0.5 * p[2], O);
      O.rin = std::min(O.prm[0], std::min(O.prm[1], O.prm[2]));
      break;
    case MPB_GEOM_CAPSULE:
      O.prm[0] = (float)p[0]; O.prm[1] = (float)(0.5 * p[1]);
      bound_box(p[0], p[0], 0.5 * p[1] + p[0], O);
      O.br = (float)(p[0] + 0.5 * p[1] + 1e-6);
      O.rin = O.prm[0];
      break;
    case MPB_GEOM_CYLINDER:
      O.prm[0] = (float)p[0]; O.prm[1] = (float)(0.5 * p[1]);
      bound_box(p[0], p[0], 0.5 * p[1], O);
      O.rin = std::min(O.prm[0], O.prm[1]);
      break;
    default: break;
  }
  primitive_balls(O);
}

// Everything of an object that depends on its pose L: sphere centres in the world / owner-joint frame, the cull
// record (MPB_COBB_WORDS float4, see dense_kernel) and, for a moving object, its record in fk_table_kernel's table.
static void object_derived(mpb_scene *s, int o) {
  DevObject &O = s->h_objs[o];
  for (int r = 0; r < 3; r++)
    O.bcw[r] = O.bcj[r] = O.L[4 * r] * O.bc[0] + O.L[4 * r + 1] * O.bc[1] + O.L[4 * r + 2] * O.bc[2] + O.L[4 * r + 3];
  double Lm[12], E[12], Ei[12], M[12];
  for (int k = 0; k < 12; k++) Lm[k] = O.L[k];
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) { E[4 * r + c] = O.oR[3 * r + c]; Ei[4 * r + c] = O.oR[3 * c + r]; }
    E[4 * r + 3] = O.oc[r];
  }
  for (int r = 0; r < 3; r++) Ei[4 * r + 3] = -(Ei[4 * r] * O.oc[0] + Ei[4 * r + 1] * O.oc[1] + Ei[4 * r + 2] * O.oc[2]);
  mul12(Lm, E, M);
  float4 *rec = s->h_cobb.data() + (size_t)MPB_COBB_WORDS * o;
  for (int r = 0; r < 3; r++) {
    rec[r] = make_float4((float)M[4 * r], (float)M[4 * r + 1], (float)M[4 * r + 2], (float)M[4 * r + 3]);
    rec[5 + r] = make_float4((float)E[4 * r], (float)E[4 * r + 1], (float)E[4 * r + 2], (float)E[4 * r + 3]);
    rec[8 + r] = make_float4((float)Ei[4 * r], (float)Ei[4 * r + 1], (float)Ei[4 * r + 2], (float)Ei[4 * r + 3]);
  }
  rec[3] = make_float4(O.oh[0], O.oh[1], O.oh[2], O.type >= MPB_GEOM_MESH ? 0.f : O.rin);
  float co[3];  // inscribed-ball centre in the OBB frame
  for (int r = 0; r < 3; r++) co[r] = (float)(Ei[4 * r] * O.ctr[0] + Ei[4 * r + 1] * O.ctr[1] + Ei[4 * r + 2] * O.ctr[2] + Ei[4 * r + 3]);
  int dyn = O.dyn;
  float df;
  memcpy(&df, &dyn, sizeof(float));
  rec[4] = make_float4(co[0], co[1], co[2], df);
  if (O.dyn >= 0)
    for (size_t e = 0; e < s->h_dynlist.size(); e++)
      if (s->h_dynlist[e] == o) {
        FkObjC &F = s->fk_tab.o[e];
        for (int r = 0; r < 3; r++) { F.rec[r] = rec[r]; F.bcj[r] = O.bcj[r]; }
        F.dyn = O.dyn;
      }
}

static inline float __int_as_float_host(int v) { float f; memcpy(&f, &v, sizeof f); return f; }

static int geom_key(int type, int geom) { return type == MPB_GEOM_CONVEX ? 16 + geom : type; }

static int build_pairs(mpb_scene *s, int n_pairs, const int32_t *pa, const int32_t *pb) {
  const int no = (int)s->h_objs.size();
  std::map<std::pair<int, int>, int> bins, bins2;
  if (pa != s->h_pair_a.data()) { s->h_pair_a.assign(pa, pa + n_pairs); s->h_pair_b.assign(pb, pb + n_pairs); }
  s->h_pairs.resize(n_pairs);
  s->h_pair_bin2.assign(n_pairs, -1);
  int expand_bin = -1, expand2_bin = -1;
  for (int p = 0; p < n_pairs; p++) {
    int a = pa[p], b = pb[p];
    if (a < 0 || a >= no || b < 0 || b >= no) return fail(MPB_ERR_INVALID, "pair index out of range");
    const bool bvh_a = s->h_objs[a].type >= MPB_GEOM_MESH, bvh_b = s->h_objs[b].type >= MPB_GEOM_MESH;
    int bin;
    if (bvh_a && bvh_b) {
      // mesh x mesh, mesh x octree: device convention A = mesh (triangle leaves), B = mesh or octree (pair index unchanged)
      if (s->h_objs[a].vcnt != MPB_GEOM_TRI) std::swap(a, b);
      if (s->h_objs[a].vcnt != MPB_GEOM_TRI)
        return fail(MPB_ERR_UNSUPPORTED, "octree x octree pairs are not produced by the reference path and are not supported");
      if (expand2_bin < 0) { expand2_bin = (int)bins.size(); bins[std::make_pair(-2, -2)] = expand2_bin; }
      bin = expand2_bin;
    } else if (bvh_a || bvh_b) {
      if (bvh_a) std::swap(a, b);  // device convention: A = shape, B = BVH object (pair index unchanged)
      if (expand_bin < 0) { expand_bin = (int)bins.size(); bins[std::make_pair(-1, -1)] = expand_bin; }
      bin = expand_bin;
      auto key2 = std::make_pair(s->obj_key[a], (int)s->h_objs[b].vcnt);  // (shape class, leaf type)
      auto it2 = bins2.find(key2);
      s->h_pair_bin2[p] = it2 == bins2.end() ? (bins2[key2] = (int)bins2.size()) : it2->second;
    } else {
      auto key = std::make_pair(s->obj_key[a], s->obj_key[b]);
      auto it = bins.find(key);
      bin = it == bins.end() ? (bins[key] = (int)bins.size()) : it->second;
    }
    if (bin > 0xffff) return fail(MPB_ERR_CAPACITY, "too many geometry-class bins");
    const int da = s->h_objs[a].dyn + 1, db = s->h_objs[b].dyn + 1;
    s->h_pairs[p] = DevPair{a, b, bin | (da << 16) | (db << 24), s->h_objs[a].br + s->h_objs[b].br + 1e-5f};
  }
  // cull table: entries sorted by (moving object, partner kind, pair index); see sphere_kernel
  {
    struct Ent { int m1, kind, p, m2; float c[3]; float rr2; int fixed_verdict; };
    std::vector<Ent> ents((size_t)n_pairs);
    for (int p = 0; p < n_pairs; p++) {
      const DevPair &pr = s->h_pairs[p];
      const DevObject &A = s->h_objs[pr.a], &B = s->h_objs[pr.b];
      Ent e{};
      e.p = p;
      e.rr2 = pr.rr * pr.rr;
      if (A.dyn >= 0 && B.dyn >= 0) { e.m1 = A.dyn; e.m2 = B.dyn; e.kind = 1; }
      else if (A.dyn >= 0) { e.m1 = A.dyn; e.m2 = -1; e.kind = 0; for (int k = 0; k < 3; k++) e.c[k] = B.bcw[k]; }
      else if (B.dyn >= 0) { e.m1 = B.dyn; e.m2 = -1; e.kind = 0; for (int k = 0; k < 3; k++) e.c[k] = A.bcw[k]; }
      else {  // both world-fixed: the moving centre reads as the origin, so the "partner centre" is the difference of the two
        // centres.  Distance mode gets the true bounds from it; the collide mode keeps a constant verdict (expanded form below)
        e.m1 = -1; e.m2 = -1; e.kind = 0;
        for (int k = 0; k < 3; k++) e.c[k] = B.bcw[k] - A.bcw[k];
        const float d2 = e.c[0] * e.c[0] + e.c[1] * e.c[1] + e.c[2] * e.c[2];
        e.fixed_verdict = d2 <= pr.rr * pr.rr ? 1 : -1;
      }
      ents[p] = e;
    }
    std::stable_sort(ents.begin(), ents.end(), [](const Ent &x, const Ent &y) {
      if (x.m1 != y.m1) return x.m1 < y.m1;
      if (x.kind != y.kind) return x.kind < y.kind;
      return x.p < y.p;
    });
    s->h_cullc.resize(n_pairs);
    s->h_cullq.resize(n_pairs);
    s->h_cullm.resize(n_pairs);
    s->h_cullp.resize(n_pairs);
    s->h_runs.clear();
    for (int k = 0; k < n_pairs; k++) {
      const Ent &e = ents[k];
      s->h_cullc[k] = make_float4(e.c[0], e.c[1], e.c[2], e.rr2);
      {  // expanded form of the fixed-partner test (sphere_kernel): slack = a few ulp of the largest term, (|a| + |c|)^2
        // with |a| within R1 + R2 of |c| whenever the verdict matters
        const double c2 = (double)e.c[0] * e.c[0] + (double)e.c[1] * e.c[1] + (double)e.c[2] * e.c[2];
        const double rr2 = e.rr2;
        double w;
        if (e.fixed_verdict > 0) w = -FLT_MAX;      // with the moving centre at the origin the test reads  w <= 0
        else if (e.fixed_verdict < 0) w = 1.0;
        else w = c2 - rr2 - 2e-6 * std::max(c2, rr2) - 1e-9;
        s->h_cullq[k] = make_float4(-2.f * e.c[0], -2.f * e.c[1], -2.f * e.c[2], (float)w);
      }
      s->h_cullm[k] = make_int2(e.p, (e.m1 + 1) | ((e.m2 + 1) << 16));
      {
        const DevPair &pr = s->h_pairs[e.p];
        s->h_cullp[k] = make_int4(e.p, pr.a, pr.b, (s->h_objs[pr.a].dyn + 1) | ((s->h_objs[pr.b].dyn + 1) << 16));
      }
      if (s->h_runs.empty() || s->h_runs.back().x != e.m1 || s->h_runs.back().w != e.kind)
        s->h_runs.push_back(make_int4(e.m1, k, 1, e.kind));
      else
        s->h_runs.back().z++;
    }
    s->h_runs.push_back(make_int4(-1, n_pairs, 0, 0));  // sentinel: the run cursor may step one past the last run
    {  // parameter-space copy for sphere_kernel<0>
      static const bool ct_on = !(getenv("MPB_CULL_PARAM") && atoi(getenv("MPB_CULL_PARAM")) == 0);
      auto fill = [&](float4 *q, int4 *runs) {
        for (int k = 0; k < n_pairs; k++) {
          const int m2 = (s->h_cullm[k].y >> 16) - 1;
          q[k] = m2 >= 0 ? make_float4(__int_as_float_host(m2), 0.f, 0.f, s->h_cullc[k].w) : s->h_cullq[k];
        }
        for (size_t r = 0; r < s->h_runs.size(); r++) runs[r] = s->h_runs[r];
      };
      s->cull_tier = 0;
      // 30 moving objects: 46 KB of staged sphere centres per block
      if (ct_on && n_pairs <= 256 && s->h_runs.size() <= 64 && s->img.n_dyn <= 30) { fill(s->ct_small.q, s->ct_small.runs); s->cull_tier = 1; }
    }
  }
  if (bins.size() >= 0xffffu || bins2.size() >= 0xffffu) return fail(MPB_ERR_CAPACITY, "more than 65534 geometry classes among the pairs");  // 16-bit bin keys
  s->grid_dirty = true;
  s->img.n_pairs = n_pairs;
  s->img.n_bins = (int)bins.size();
  s->img.n_bins2 = (int)bins2.size();
  s->img.expand_bin = expand_bin;
  s->img.expand2_bin = expand2_bin;
  s->img.pair_words = (n_pairs + 31) / 32;
  return MPB_OK;
}

static int upload_pairs(mpb_scene *s) {
  const size_t n = s->h_pairs.size();
  if (n > s->pairs_cap) {
    if (s->d_pairs) cudaFree(s->d_pairs);
    s->pairs_cap = std::max<size_t>(n, 64);
    CU(cudaMalloc(&s->d_pairs, s->pairs_cap * sizeof(DevPair)));
  }
  if (n) CU(cudaMemcpy(s->d_pairs, s->h_pairs.data(), n * sizeof(DevPair), cudaMemcpyHostToDevice));
  s->img.pairs = s->d_pairs;
  if (n > s->pair_bin2_cap) {
    if (s->d_pair_bin2) cudaFree(s->d_pair_bin2);
    s->pair_bin2_cap = std::max<size_t>(n, 64);
    CU(cudaMalloc(&s->d_pair_bin2, s->pair_bin2_cap * sizeof(int)));
  }
  if (n) CU(cudaMemcpy(s->d_pair_bin2, s->h_pair_bin2.data(), n * sizeof(int), cudaMemcpyHostToDevice));
  s->img.pair_bin2 = s->d_pair_bin2;
  if (n + 1 > s->cull_cap) {
    if (s->d_cullc) cudaFree(s->d_cullc);
    if (s->d_cullq) cudaFree(s->d_cullq);
    if (s->d_cullm) cudaFree(s->d_cullm);
    if (s->d_cullp) cudaFree(s->d_cullp);
    if (s->d_runs) cudaFree(s->d_runs);
    s->cull_cap = std::max<size_t>(n + 1, 64);
    CU(cudaMalloc(&s->d_cullc, s->cull_cap * sizeof(float4)));
    CU(cudaMalloc(&s->d_cullq, s->cull_cap * sizeof(float4)));
    CU(cudaMalloc(&s->d_cullm, s->cull_cap * sizeof(int2)));
    CU(cudaMalloc(&s->d_cullp, s->cull_cap * sizeof(int4)));
    CU(cudaMalloc(&s->d_runs, s->cull_cap * sizeof(int4)));
  }
  if (n) CU(cudaMemcpy(s->d_cullc, s->h_cullc.data(), n * sizeof(float4), cudaMemcpyHostToDevice));
  if (n) CU(cudaMemcpy(s->d_cullq, s->h_cullq.data(), n * sizeof(float4), cudaMemcpyHostToDevice));
  if (n) CU(cudaMemcpy(s->d_cullm, s->h_cullm.data(), n * sizeof(int2), cudaMemcpyHostToDevice));
  if (n) CU(cudaMemcpy(s->d_cullp, s->h_cullp.data(), n * sizeof(int4), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(s->d_runs, s->h_runs.data(), s->h_runs.size() * sizeof(int4), cudaMemcpyHostToDevice));
  s->img.cullc = s->d_cullc;
  s->img.cullq = s->d_cullq;
  s->img.cullm = s->d_cullm;
  s->img.cullp = s->d_cullp;
  s->img.runs = s->d_runs;
  return MPB_OK;
}

extern "C" int mpb_scene_create(const mpb_scene_desc *d, int device, mpb_scene **out) {
  if (!d || !out) return fail(MPB_ERR_INVALID, "null argument");
  *out = nullptr;
  if (d->n_joints > MPB_MAX_JOINTS) return fail(MPB_ERR_CAPACITY, "too many movable joints (MPB_MAX_JOINTS)");
  if (d->n_joints < 0 || d->n_objects < 0 || d->n_pairs < 0 || d->state_dim < 0 || d->n_links < 0)
    return fail(MPB_ERR_INVALID, "negative count in scene description");
  int ndev = mpb_device_count();
  if (ndev == 0) return fail(MPB_ERR_CUDA, "no CUDA device visible: libmpb has no CPU path");
  if (device < 0 || device >= ndev) return fail(MPB_ERR_INVALID, "device index out of range");
  CU(cudaSetDevice(device));
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  mpb_scene *s = new mpb_scene();
  s->device = device;
  s->num_sms = prop.multiProcessorCount;
  s->nq_full = d->nq_full;

  // joints
  std::vector<int> slot_of_q(std::max(d->nq_full, 1), -1);
  for (int k = 0; k < d->state_dim; k++) {
    int qi = d->state_map[k];
    if (qi < 0 || qi >= d->nq_full) { delete s; return fail(MPB_ERR_INVALID, "state_map entry out of range"); }
    slot_of_q[qi] = k;
  }
  s->h_joints.resize(d->n_joints);
  s->joint_qidx.assign(d->joint_qidx, d->joint_qidx + d->n_joints);
  s->h_jP64.assign(d->joint_placement, d->joint_placement + (size_t)12 * d->n_joints);
  s->h_jax64.assign(d->joint_axis, d->joint_axis + (size_t)3 * d->n_joints);
  s->h_jtype.assign(d->joint_type, d->joint_type + d->n_joints);
  s->h_jparent.assign(d->joint_parent, d->joint_parent + d->n_joints);
  s->h_qbase64.assign(d->q_base, d->q_base + d->nq_full);
  s->h_linkL64.assign(d->link_placement, d->link_placement + (size_t)12 * d->n_links);
  s->h_linkJoint.assign(d->link_joint, d->link_joint + d->n_links);
  for (int j = 0; j < d->n_joints; j++) {
    DevJoint &J = s->h_joints[j];
    for (int k = 0; k < 12; k++) J.P[k] = (float)d->joint_placement[12 * j + k];
    for (int k = 0; k < 3; k++) J.ax[k] = (float)d->joint_axis[3 * j + k];
    J.type = d->joint_type[j];
    J.parent = d->joint_parent[j];
    if (J.parent < 0 || J.parent > j) { delete s; return fail(MPB_ERR_INVALID, "joint table is not topologically ordered"); }
    int qi = d->joint_qidx[j];
    if (qi < 0 || qi >= d->nq_full) { delete s; return fail(MPB_ERR_INVALID, "joint_qidx out of range"); }
    J.slot = slot_of_q[qi];
    J.qconst = (float)d->q_base[qi];
    J.jslot = -1;
    J.pjslot = -1;
    J.dyn_first = 0;
    J.dyn_count = 0;
    J.kind = MPB_JK_GENERIC;
    if (J.type == MPB_JOINT_PRISMATIC) J.kind = MPB_JK_PRISMATIC;
    else
      for (int a = 0; a < 3; a++)
        if (fabs(fabs(d->joint_axis[3 * j + a]) - 1.0) < 1e-12 && fabs(d->joint_axis[3 * j + (a + 1) % 3]) < 1e-12 &&
            fabs(d->joint_axis[3 * j + (a + 2) % 3]) < 1e-12) {
          J.kind = MPB_JK_RX + a;
          J.ax[a] = d->joint_axis[3 * j + a] > 0 ? 1.f : -1.f;
        }
  }
  // objects
  s->h_objs.resize(d->n_objects);
  s->obj_key.resize(d->n_objects);
  int ndyn = 0;
  std::vector<float4> verts;
  std::vector<int> hull_off(d->n_hulls + 1, 0);
  for (int h = 0; h < d->n_hulls; h++) {
    hull_off[h] = (int)verts.size();
    for (int i = d->hull_vert_off[h]; i < d->hull_vert_off[h + 1]; i++)
      verts.push_back(make_float4((float)d->hull_verts[3 * i], (float)d->hull_verts[3 * i + 1], (float)d->hull_verts[3 * i + 2], 0.f));
  }
  hull_off[d->n_hulls] = (int)verts.size();
  const int n_hull_verts = (int)verts.size();
  // support cell tables behind the vertices (same pool: one TMA stage brings both into shared memory)
  std::vector<int> hull_cell(std::max(d->n_hulls, 0), -1);
  {
    std::vector<unsigned char> pool, one;
    for (int h = 0; h < d->n_hulls; h++) {
      const int v0 = d->hull_vert_off[h], nv = d->hull_vert_off[h + 1] - v0;
      if (nv <= 0 || nv > 256) continue;  // byte-sized candidate indices; larger hulls keep the full scan
      build_cell_table(d->hull_verts + 3 * (size_t)v0, nv, one);
      if (one.size() - 2 * (MPB_CELL_COUNT + 1) > 65535) continue;
      hull_cell[h] = (int)pool.size();
      pool.insert(pool.end(), one.begin(), one.end());
      pool.resize((pool.size() + 15) / 16 * 16, 0);
    }
    const size_t at = verts.size();
    verts.resize(at + pool.size() / 16);
    if (!pool.empty()) memcpy(verts.data() + at, pool.data(), pool.size());
  }
  // BVH objects: one tree per mesh / octree in the pools
  std::vector<DevBvhNode> nodes;
  std::vector<float4> prims;
  std::vector<int> mesh_root(std::max(d->n_meshes, 0), -1), oct_root(std::max(d->n_octrees, 0), -1);
  auto finish_tree = [&](std::vector<HostPrim> &hp) -> int {
    if (hp.empty()) return -1;
    std::vector<DevBvhNode> built;
    build_bvh(hp, 0, (int)hp.size(), built);
    // Breadth-first re-layout with siblings adjacent at an even index (right == left + 1, 64-byte aligned pair): a
    // traversal that keeps "first child of a passing inner node" on its stack reads both children with one access
    // (expand_lane_kernel).  Slot 1 pads the root to a pair; nothing points at it.
    std::vector<DevBvhNode> local(2);
    local[0] = built[0];
    local[1] = DevBvhNode{};
    local[1].left = ~0; local[1].right = 0;
    for (int a = 0; a < 3; a++) { local[1].c[a] = 0.f; local[1].h[a] = -1.f; }
    {
      std::vector<std::pair<int, int>> fifo{{0, 0}};
      for (size_t qi = 0; qi < fifo.size(); qi++) {
        const int was = fifo[qi].first, now = fifo[qi].second;
        if (built[was].left < 0) continue;
        const int at = (int)local.size();
        local.push_back(built[built[was].left]);
        local.push_back(built[built[was].right]);
        fifo.push_back({built[was].left, at});
        fifo.push_back({built[was].right, at + 1});
        local[now].left = at;
        local[now].right = at + 1;
      }
    }
    if (nodes.size() & 1) nodes.push_back(local[1]);
    const int node0 = (int)nodes.size(), prim0 = (int)(prims.size() / 3);
    for (DevBvhNode nd : local) {
      if (nd.left >= 0) { nd.left += node0; nd.right += node0; }
      else nd.left = ~(~nd.left + prim0);
      nodes.push_back(nd);
    }
    for (const HostPrim &h : hp) { prims.push_back(h.a); prims.push_back(h.b); prims.push_back(h.c); }
    return node0;
  };
  for (int m = 0; m < d->n_meshes; m++) {
    std::vector<HostPrim> hp;
    const double *V = d->mesh_verts + 3 * (size_t)d->mesh_vert_off[m];
    for (int t = d->mesh_tri_off[m]; t < d->mesh_tri_off[m + 1]; t++) {
      HostPrim h;
      const double *P[3] = {V + 3 * d->mesh_tris[3 * t], V + 3 * d->mesh_tris[3 * t + 1], V + 3 * d->mesh_tris[3 * t + 2]};
      h.a = make_float4((float)P[0][0], (float)P[0][1], (float)P[0][2], 0.f);
      h.b = make_float4((float)P[1][0], (float)P[1][1], (float)P[1][2], 0.f);
      h.c = make_float4((float)P[2][0], (float)P[2][1], (float)P[2][2], 0.f);
      for (int a = 0; a < 3; a++) {
        h.lo[a] = std::min(P[0][a], std::min(P[1][a], P[2][a]));
        h.hi[a] = std::max(P[0][a], std::max(P[1][a], P[2][a]));
      }
      hp.push_back(h);
    }
    mesh_root[m] = finish_tree(hp);
  }
  std::vector<std::array<double, 6>> oct_bounds(std::max(d->n_octrees, 0));
  for (int t = 0; t < d->n_octrees; t++) {
    const int c0 = d->oct_cell_off[t], nc = d->oct_cell_off[t + 1] - c0;
    const double res = d->oct_res[t];
    std::vector<CellBox> boxes = merge_cells(d->oct_cells + 3 * (size_t)c0, nc);
    std::vector<HostPrim> hp;
    std::array<double, 6> bb = {1e300, 1e300, 1e300, -1e300, -1e300, -1e300};
    for (const CellBox &b : boxes) {
      HostPrim h;
      double c[3], hh[3];
      for (int a = 0; a < 3; a++) {
        h.lo[a] = b.lo[a] * res;
        h.hi[a] = (b.hi[a] + 1) * res;
        c[a] = 0.5 * (h.lo[a] + h.hi[a]);
        hh[a] = 0.5 * (h.hi[a] - h.lo[a]);
        bb[a] = std::min(bb[a], h.lo[a]);
        bb[3 + a] = std::max(bb[3 + a], h.hi[a]);
      }
      h.a = make_float4((float)c[0], (float)c[1], (float)c[2], 0.f);
      h.b = make_float4((float)hh[0], (float)hh[1], (float)hh[2], 0.f);
      h.c = make_float4(0.f, 0.f, 0.f, 0.f);
      hp.push_back(h);
    }
    oct_bounds[t] = bb;
    oct_root[t] = finish_tree(hp);
  }
  for (int o = 0; o < d->n_objects; o++) {
    DevObject &O = s->h_objs[o];
    memset(&O, 0, sizeof(O));
    O.cell = -1;
    O.type = d->obj_type[o];
    O.geom = d->obj_geom[o];
    double L[12];
    int link = d->obj_link[o];
    if (link >= d->n_links) { delete s; return fail(MPB_ERR_INVALID, "obj_link out of range"); }
    if (link >= 0) {
      mul12(d->link_placement + 12 * link, d->obj_local + 12 * o, L);
      O.joint = d->link_joint[link];
    } else {
      memcpy(L, d->obj_local + 12 * o, sizeof(L));
      O.joint = 0;
    }
    for (int k = 0; k < 12; k++) O.L[k] = (float)L[k];
    O.dyn = O.joint > 0 ? ndyn++ : -1;
    const double *p = d->obj_param + 4 * o;
    switch (O.type) {
      case MPB_GEOM_SPHERE:
      case MPB_GEOM_BOX:
      case MPB_GEOM_CAPSULE:
      case MPB_GEOM_CYLINDER:
        primitive_bounds(O, p);
        break;
      case MPB_GEOM_CONVEX: {
        int h = O.geom;
        if (h < 0 || h >= d->n_hulls) { delete s; return fail(MPB_ERR_INVALID, "obj_geom (hull) out of range"); }
        int v0 = d->hull_vert_off[h], nv = d->hull_vert_off[h + 1] - v0;
        if (nv <= 0) { delete s; return fail(MPB_ERR_INVALID, "empty convex hull"); }
        O.voff = hull_off[h];
        O.vcnt = nv;
        O.cell = hull_cell[h];
        bound_points(d->hull_verts + 3 * v0, nv, O);
        double c[3] = {0, 0, 0};
        for (int i = 0; i < nv; i++)
          for (int a = 0; a < 3; a++) c[a] += d->hull_verts[3 * (v0 + i) + a];
        for (int a = 0; a < 3; a++) O.ctr[a] = (float)(c[a] / nv);
      } break;
      case MPB_GEOM_MESH: {
        const int m = O.geom;
        if (m < 0 || m >= d->n_meshes || mesh_root[m] < 0) { delete s; return fail(MPB_ERR_INVALID, "obj_geom (mesh) out of range or empty mesh"); }
        bound_points(d->mesh_verts + 3 * (size_t)d->mesh_vert_off[m], d->mesh_vert_off[m + 1] - d->mesh_vert_off[m], O);
        O.voff = mesh_root[m];
        O.vcnt = MPB_GEOM_TRI;
      } break;
      case MPB_GEOM_OCTREE: {
        const int t = O.geom;
        if (t < 0 || t >= d->n_octrees) { delete s; return fail(MPB_ERR_INVALID, "obj_geom (octree) out of range"); }
        if (oct_root[t] < 0) {  // empty tree: a far-away zero box that never survives the cull
          bound_box(0, 0, 0, O);
          O.oc[0] = O.bc[0] = 1e6f;
          O.voff = -1; O.vcnt = MPB_GEOM_VBOX;  // no tree: the traversal kernels skip the object
          break;
        }
        const auto &bb = oct_bounds[t];
        bound_box(0.5 * (bb[3] - bb[0]), 0.5 * (bb[4] - bb[1]), 0.5 * (bb[5] - bb[2]), O);
        for (int a = 0; a < 3; a++) O.oc[a] = O.bc[a] = (float)(0.5 * (bb[a] + bb[3 + a]));
        O.voff = oct_root[t];
        O.vcnt = MPB_GEOM_VBOX;
      } break;
      default:
        delete s;
        return fail(MPB_ERR_INVALID, "unknown geometry type");
    }
    switch (O.type) {  // ball around ctr contained in the shape (0 disables the sufficient-hit test)
      case MPB_GEOM_SPHERE: case MPB_GEOM_BOX: case MPB_GEOM_CAPSULE: case MPB_GEOM_CYLINDER: break;  // primitive_bounds
      case MPB_GEOM_CONVEX: {
        const int h = O.geom;
        const double *V = d->hull_verts + 3 * (size_t)d->hull_vert_off[h];
        double rmin = 1e300;
        for (int f = d->hull_face_off[h]; f < d->hull_face_off[h + 1]; f++) {
          const double *A0 = V + 3 * d->hull_faces[3 * f], *A1 = V + 3 * d->hull_faces[3 * f + 1], *A2 = V + 3 * d->hull_faces[3 * f + 2];
          double e1[3], e2[3], nrm[3];
          for (int a = 0; a < 3; a++) { e1[a] = A1[a] - A0[a]; e2[a] = A2[a] - A0[a]; }
          nrm[0] = e1[1] * e2[2] - e1[2] * e2[1]; nrm[1] = e1[2] * e2[0] - e1[0] * e2[2]; nrm[2] = e1[0] * e2[1] - e1[1] * e2[0];
          const double l = sqrt(nrm[0] * nrm[0] + nrm[1] * nrm[1] + nrm[2] * nrm[2]);
          if (l < 1e-14) continue;
          const double dist = fabs(((double)O.ctr[0] - A0[0]) * nrm[0] + ((double)O.ctr[1] - A0[1]) * nrm[1] + ((double)O.ctr[2] - A0[2]) * nrm[2]) / l;
          rmin = std::min(rmin, dist);
        }
        O.rin = rmin < 1e299 ? (float)(0.98 * rmin) : 0.f;  // valid for a convex hull with the centre inside
        hull_balls(V, d->hull_vert_off[h + 1] - d->hull_vert_off[h], d->hull_faces + 3 * (size_t)d->hull_face_off[h],
                   d->hull_face_off[h + 1] - d->hull_face_off[h], O.ball);
      } break;
      default: O.rin = 0.f; break;
    }
    s->obj_key[o] = geom_key(O.type, O.geom);
    if ((O.type >= MPB_GEOM_SPHERE && O.type <= MPB_GEOM_CYLINDER) || O.type == MPB_GEOM_OCTREE) s->has_primitives = true;
  }
  if (ndyn > MPB_MAX_DYN) { delete s; return fail(MPB_ERR_CAPACITY, "too many moving collision objects (MPB_MAX_DYN = 128)"); }
  std::vector<int> dynlist;
  for (int j = 0; j < d->n_joints; j++) {
    s->h_joints[j].dyn_first = (int)dynlist.size();
    for (int o = 0; o < d->n_objects; o++)
      if (s->h_objs[o].joint == j + 1) dynlist.push_back(o);
    s->h_joints[j].dyn_count = (int)dynlist.size() - s->h_joints[j].dyn_first;
  }
  // joint-table slots: joints that own moving objects, and parents of non-adjacent children (branch points)
  int n_jslots = 0;
  for (int j = 0; j < d->n_joints; j++) {
    const int par = s->h_joints[j].parent;  // model index; the previous joint has model index j
    if (par == 0 || par == j) continue;
    DevJoint &P = s->h_joints[par - 1];
    if (P.jslot < 0) P.jslot = n_jslots++;
    s->h_joints[j].pjslot = P.jslot;
  }
  for (int o = 0; o < d->n_objects; o++) s->h_objs[o].jslot = s->h_objs[o].joint > 0 ? s->h_joints[s->h_objs[o].joint - 1].jslot : -1;
  s->img.n_dyn = ndyn;
  s->img.n_jslots = n_jslots;
  s->img.w_sphere = 0;
  s->img.w_joint = ndyn * 3;
  // cull records, sphere centres and fk_table_kernel's parameter-space tables (joints; moving objects grouped by owner joint)
  s->h_dynlist = dynlist;
  s->h_cobb.assign((size_t)d->n_objects * MPB_COBB_WORDS, make_float4(0.f, 0.f, 0.f, 0.f));
  for (int j = 0; j < d->n_joints; j++) s->fk_tab.j[j] = s->h_joints[j];
  for (int o = 0; o < d->n_objects; o++) object_derived(s, o);
  s->h_objLink.assign(d->obj_link, d->obj_link + d->n_objects);
  std::vector<float4> &cobb = s->h_cobb;

  s->img.n_joints = d->n_joints;
  s->img.n_objects = d->n_objects;
  s->img.n_links = d->n_links;
  s->img.n_verts = n_hull_verts;
  s->img.pool_bytes = (int)(verts.size() * sizeof(float4));
  s->img.state_dim = d->state_dim;
  int rc = build_pairs(s, d->n_pairs, d->pair_a, d->pair_b);
  if (rc != MPB_OK) { delete s; return rc; }

  // upload
  CU(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
  auto up = [&](void **dst, const void *src, size_t bytes) -> cudaError_t {
    cudaError_t e = cudaMalloc(dst, std::max<size_t>(bytes, 16));
    if (e != cudaSuccess) return e;
    return bytes ? cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice) : cudaSuccess;
  };
  CU(up((void **)&s->d_joints, s->h_joints.data(), s->h_joints.size() * sizeof(DevJoint)));
  CU(up((void **)&s->d_objs, s->h_objs.data(), s->h_objs.size() * sizeof(DevObject)));
  CU(up((void **)&s->d_verts, verts.data(), verts.size() * sizeof(float4)));
  CU(up((void **)&s->d_cobb, cobb.data(), cobb.size() * sizeof(float4)));
  s->img.cobb = s->d_cobb;
  std::vector<float> linkL((size_t)d->n_links * 12);
  for (size_t k = 0; k < linkL.size(); k++) linkL[k] = (float)d->link_placement[k];
  CU(up((void **)&s->d_linkL, linkL.data(), linkL.size() * sizeof(float)));
  CU(up((void **)&s->d_linkJoint, d->link_joint, (size_t)d->n_links * sizeof(int)));
  CU(up((void **)&s->d_dynlist, dynlist.data(), dynlist.size() * sizeof(int)));
  s->img.dynlist = s->d_dynlist;
  CU(up((void **)&s->d_nodes, nodes.data(), nodes.size() * sizeof(DevBvhNode)));
  CU(up((void **)&s->d_prims, prims.data(), prims.size() * sizeof(float4)));
  s->img.nodes = s->d_nodes;
  s->img.prims = s->d_prims;
  s->img.joints = s->d_joints;
  s->img.objs = s->d_objs;
  s->img.verts = s->d_verts;
  s->img.linkL = s->d_linkL;
  s->img.linkJoint = s->d_linkJoint;
  rc = upload_pairs(s);
  if (rc != MPB_OK) { mpb_scene_destroy(s); return rc; }
  CU(cudaHostAlloc((void **)&s->h_status, (32 + kMaxHitPasses) * sizeof(unsigned), cudaHostAllocDefault));
  s->h_hitcnt = s->h_status + 32;
  {
    std::vector<float> qc(std::max(d->n_joints, 1), 0.f);
    for (int j = 0; j < d->n_joints; j++) qc[j] = s->h_joints[j].qconst;
    CU(up((void **)&s->d_qconst, qc.data(), qc.size() * sizeof(float)));
  }

  // narrowphase launch shape: hull pool in shared memory when it fits
  const size_t vbytes = verts.size() * sizeof(float4);
  s->smem_verts = vbytes <= 200 * 1024;
  int per_sm = 0;
  if (s->smem_verts) {
    CU(raise_dyn_smem((const void *)narrow_kernel<true>, device, 0, std::max<size_t>(vbytes, 16)));
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, narrow_kernel<true>, 256, vbytes));
  } else {
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, narrow_kernel<false>, 256, 0));
  }
  {  // phase-split variant: hull pool (padded to 128 B) + one survivor buffer per warp
    s->narrow2_smem = (s->smem_verts ? (vbytes + 127) / 128 * 128 : 0) + (size_t)8 * (MPB_SURV_F + 1) * MPB_SURV_CAP * sizeof(float);
    int b2 = 0;
    if (s->smem_verts) {
      CU(raise_dyn_smem((const void *)narrow2_kernel<true>, device, 2, s->narrow2_smem));
      CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b2, narrow2_kernel<true>, 256, s->narrow2_smem));
    } else {
      CU(raise_dyn_smem((const void *)narrow2_kernel<false>, device, 3, s->narrow2_smem));
      CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b2, narrow2_kernel<false>, 256, s->narrow2_smem));
    }
    s->narrow2_blocks = s->num_sms * std::max(b2, 1);
  }
  s->narrow_blocks = s->num_sms * std::max(per_sm, 1);
  {  // sphere_kernel / dense_kernel split their range statically over the grid: exactly one resident wave, or the
     // blocks beyond it run as a second, mostly empty wave
    int b = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, sphere_kernel<0, 0, 0>, 128, 0)); s->sphere_blocks[0] = s->num_sms * std::max(b, 1);
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, sphere_kernel<0, 256, 64>, 128, (size_t)s->img.n_dyn * 3 * 128 * sizeof(float))); s->sphere_blocks_ct[0] = s->num_sms * std::max(b, 1);
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, sphere_blk_kernel<2>, 128, 0)); s->sphere_blocks_blk[0] = s->num_sms * std::max(b, 1);
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, sphere_blk_kernel<4>, 128, 0)); s->sphere_blocks_blk[1] = s->num_sms * std::max(b, 1);
    if (getenv("MPB_SPHERE_CB")) s->sphere_cb = atoi(getenv("MPB_SPHERE_CB"));
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, sphere_kernel<1, 0, 0>, 128, 0)); s->sphere_blocks[1] = s->num_sms * std::max(b, 1);
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, dense_kernel<0>, 128, 0)); s->dense_blocks[0] = s->num_sms * std::max(b, 1);
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, dense_kernel<1>, 128, 0)); s->dense_blocks[1] = s->num_sms * std::max(b, 1);
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, dense2_kernel, 128, 0)); s->dense2_blocks = s->num_sms * std::max(b, 1);
    {  // fk_sphere_kernel: the joint-table slots some joint names as its parent stay in shared memory
      s->jneed = 0;
      for (const DevJoint &J : s->h_joints)
        if (J.pjslot >= 0) s->jneed |= 1u << J.pjslot;
      s->fk_sphere_smem = ((size_t)s->img.n_dyn * 3 + 12 * (size_t)__builtin_popcount(s->jneed)) * 128 * sizeof(float);
      if (s->fk_sphere_smem <= 48 * 1024) {
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, fk_sphere_kernel, 128, s->fk_sphere_smem));
        s->fk_sphere_blocks = s->num_sms * std::max(b, 1);
      }
    }
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, expand_lane_kernel, 128, 0)); s->expand_lane_blocks = s->num_sms * std::max(b, 1);
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, sphere_grid_kernel, 128, 0)); s->sphere_blocks_grid = s->num_sms * std::max(b, 1);
  }
  *out = s;
  return MPB_OK;
}

extern "C" void mpb_scene_destroy(mpb_scene *s) {
  if (!s) return;
  cudaSetDevice(s->device);
  if (s->stream) cudaStreamSynchronize(s->stream);
  void *ptrs[] = {s->d_joints, s->d_objs, s->d_pairs, s->d_verts, s->d_linkL, s->d_linkJoint, s->d_q[0], s->d_q[1], s->d_verdict[0],
                  s->d_verdict[1], s->d_pairbits[0], s->d_pairbits[1], s->d_poses, s->d_dynlist, s->d_nodes, s->d_prims, s->d_pair_bin2, s->d_pose, s->d_obb, s->d_sq, s->d_cobb, s->d_cullc, s->d_cullq, s->d_cullm, s->d_cullp, s->d_runs, s->wq.rec[0], s->wq.bin_count, s->wq.totals};
  for (void *p : ptrs)
    if (p) cudaFree(p);
  harvest(s, true);  // device-pointer calls may still be in flight on caller streams
  sync_scene(s);
  drop_small_graphs(s);
  if (s->h_small) cudaFreeHost(s->h_small);
  if (s->h_status) cudaFreeHost(s->h_status);
  if (s->h_slots) cudaFreeHost(s->h_slots);
  if (s->d_hits) cudaFree(s->d_hits);
  if (s->d_small) cudaFree(s->d_small);
  if (s->d_grid_start) cudaFree(s->d_grid_start);
  if (s->d_grid_list) cudaFree(s->d_grid_list);
  if (s->d_grid_run) cudaFree(s->d_grid_run);
  if (s->d_qconst) cudaFree(s->d_qconst);
  for (auto &sl : s->slots) {
    if (sl.done) cudaEventDestroy(sl.done);
    for (auto e : sl.ev) cudaEventDestroy(e);
  }
  for (auto &b : s->pool) cudaFree(b.first);
  for (auto e : s->events) cudaEventDestroy(e);
  for (auto e : s->sync_events) cudaEventDestroy(e);
  if (s->stream_h2d) cudaStreamDestroy(s->stream_h2d);
  if (s->stream_d2h) cudaStreamDestroy(s->stream_d2h);
  if (s->stream_hits) cudaStreamDestroy(s->stream_hits);
  if (s->stream) cudaStreamDestroy(s->stream);
  delete s;
}

extern "C" int mpb_scene_info(const mpb_scene *s, int32_t *n_pairs, int32_t *state_dim, int32_t *n_links, int32_t *n_objects,
                              int32_t *device) {
  if (!s) return fail(MPB_ERR_INVALID, "null scene");
  if (n_pairs) *n_pairs = s->img.n_pairs;
  if (state_dim) *state_dim = s->img.state_dim;
  if (n_links) *n_links = s->img.n_links;
  if (n_objects) *n_objects = s->img.n_objects;
  if (device) *device = s->device;
  return MPB_OK;
}

extern "C" int mpb_scene_set_q_base(mpb_scene *s, const double *q_base, int32_t nq_full) {
  if (!s || !q_base) return fail(MPB_ERR_INVALID, "null argument");
  if (nq_full != s->nq_full) return fail(MPB_ERR_INVALID, "q_base length does not match the scene");
  CU(cudaSetDevice(s->device));
  s->h_qbase64.assign(q_base, q_base + nq_full);
  // only joints the batch state does not write read q_base on the device: the legacy single-state calls change the
  // articulation's qpos before every query, but rarely the joints outside the move group
  bool changed = false;
  for (size_t j = 0; j < s->h_joints.size(); j++) {
    const float v = (float)q_base[s->joint_qidx[j]];
    if (s->h_joints[j].slot < 0 && v != s->h_joints[j].qconst) changed = true;
    s->h_joints[j].qconst = v;
    s->fk_tab.j[j].qconst = v;
  }
  if (!changed || s->h_joints.empty()) return MPB_OK;
  int rc = harvest(s, true);
  if (rc != MPB_OK) return rc;
  CU(cudaStreamSynchronize(s->stream));
  std::vector<float> qc(s->h_joints.size());
  for (size_t j = 0; j < qc.size(); j++) qc[j] = s->h_joints[j].qconst;
  CU(cudaMemcpy(s->d_qconst, qc.data(), qc.size() * sizeof(float), cudaMemcpyHostToDevice));  // kernels read it through this table
  CU(cudaMemcpy(s->d_joints, s->h_joints.data(), s->h_joints.size() * sizeof(DevJoint), cudaMemcpyHostToDevice));
  return MPB_OK;
}

extern "C" int mpb_scene_set_pairs(mpb_scene *s, int32_t n_pairs, const int32_t *pa, const int32_t *pb) {
  if (!s || n_pairs < 0 || (n_pairs && (!pa || !pb))) return fail(MPB_ERR_INVALID, "bad pair list");
  CU(cudaSetDevice(s->device));
  int rc = harvest(s, true);
  if (rc != MPB_OK) return rc;
  CU(sync_scene(s));
  rc = build_pairs(s, n_pairs, pa, pb);
  if (rc != MPB_OK) return rc;
  s->img_version++;
  s->calibrated = false;
  s->good_chunk = 0;
  s->cap_chunk = 1ll << 40;
  return upload_pairs(s);
}

// In-place update of objects of a compiled scene.  Geometry (type, hull / mesh / octree) stays; what may change is the
// pose of a world-fixed object (CollisionObject::setTransformation, python/pybind_fcl.hpp:254-259, used in place by
// mplib/sapien_utils/conversion.py:176-178), the pose of an attached body on its link (AttachedBodyTpl::setPose,
// src/attached_body.h) and the parameters of a primitive.  Only the records derived from those are recomputed.
extern "C" int mpb_scene_update_objects(mpb_scene *s, int32_t n, const int32_t *obj_idx, const double *obj_local, const double *obj_param) {
  if (!s || n < 0 || (n > 0 && (!obj_idx || (!obj_local && !obj_param)))) return fail(MPB_ERR_INVALID, "bad object update");
  if (n == 0) return MPB_OK;
  const int no = (int)s->h_objs.size();
  for (int k = 0; k < n; k++)
    if (obj_idx[k] < 0 || obj_idx[k] >= no) return fail(MPB_ERR_INVALID, "object index out of range");
  CU(cudaSetDevice(s->device));
  int rc = harvest(s, true);
  if (rc != MPB_OK) return rc;
  CU(sync_scene(s));
  for (int k = 0; k < n; k++) {
    const int o = obj_idx[k];
    DevObject &O = s->h_objs[o];
    if (obj_param && O.type >= MPB_GEOM_SPHERE && O.type <= MPB_GEOM_CYLINDER) primitive_bounds(O, obj_param + 4 * (size_t)k);
    if (obj_local) {
      double L[12];
      const int link = s->h_objLink[o];
      if (link >= 0) mul12(s->h_linkL64.data() + 12 * (size_t)link, obj_local + 12 * (size_t)k, L);
      else memcpy(L, obj_local + 12 * (size_t)k, sizeof(L));
      for (int i = 0; i < 12; i++) O.L[i] = (float)L[i];
    }
    object_derived(s, o);
  }
  // the cull table holds sphere centres and radius sums of the pairs: rebuild it (O(pairs), no geometry work)
  rc = build_pairs(s, (int)s->h_pair_a.size(), s->h_pair_a.data(), s->h_pair_b.data());
  if (rc != MPB_OK) return rc;
  CU(cudaMemcpy(s->d_objs, s->h_objs.data(), s->h_objs.size() * sizeof(DevObject), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(s->d_cobb, s->h_cobb.data(), s->h_cobb.size() * sizeof(float4), cudaMemcpyHostToDevice));
  s->img_version++;
  s->calibrated = false;
  s->good_chunk = 0;
  s->cap_chunk = 1ll << 40;
  return upload_pairs(s);
}

// ------------------------------------------------------------------------------------------ work buffers
template <typename T>
static cudaError_t ensure(T **p, size_t *cap, size_t need) {
  if (need <= *cap && *p) return cudaSuccess;
  if (*p) cudaFree(*p);
  *p = nullptr;
  *cap = 0;
  cudaError_t e = cudaMalloc((void **)p, std::max<size_t>(need, 1) * sizeof(T));
  if (e == cudaSuccess) *cap = need;
  return e;
}

static const long long kMaxChunk = 1ll << 20;       // configurations per pass
static const long long kSmallBatch = 4096;          // host batches up to this size replay a CUDA graph (collide_small)
static const size_t kMaxItems = (size_t)48 << 20;   // work-queue capacity bound (x56 B = 2.6 GiB)
static const size_t kMaxSurvivors = (size_t)192 << 20;  // survivor-queue capacity bound (x8 B = 1.5 GiB)

// Largest pass whose queue counters (32-bit: survivor slots, staged items) cannot wrap even if every admitted pair of
// every configuration survives.
static long long max_chunk_for(const mpb_scene *s) {
  const long long P = std::max(s->img.n_pairs, 1);
  const long long lim = ((1ll << 31) / P) / 32 * 32;
  return std::max<long long>(32, std::min<long long>(kMaxChunk, lim));
}

// Broadphase grid of sphere_grid_kernel: per run of world-fixed partners, every partner's sphere (radius = sum of the two
// bounding radii, as in the cull table) rasterised into a uniform grid; cell -> list of cull-table entries.  Built lazily
// (a few ms for the C3 scene) so that a scene edit followed by small queries does not pay for it.
static int ensure_grid(mpb_scene *s) {
  if (!s->grid_dirty) return MPB_OK;
  s->grid_dirty = false;
  s->grid_on = false;
  static const bool enabled = !(getenv("MPB_NO_GRID") && atoi(getenv("MPB_NO_GRID")));
  static const double grid_div = getenv("MPB_GRID_DIV") ? atof(getenv("MPB_GRID_DIV")) : 32.0;
  const size_t n_runs = s->h_runs.size();  // incl. the sentinel
  std::vector<int2> grun(n_runs, make_int2(-1, 0));
  size_t fixed_entries = 0;
  for (const int4 &r : s->h_runs)
    if (r.w == 0) fixed_entries += (size_t)r.z;
  if (!enabled || s->cull_tier != 0 || fixed_entries < 256) return MPB_OK;  // small tables: the parameter-space / direct kernels
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  for (const int4 &r : s->h_runs) {
    if (r.w != 0) continue;
    for (int k = r.y; k < r.y + r.z; k++) {
      const float4 c = s->h_cullc[k];
      const double rr = std::sqrt((double)c.w);
      const double cc[3] = {c.x, c.y, c.z};
      for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], cc[a] - rr); hi[a] = std::max(hi[a], cc[a] + rr); }
    }
  }
  double ext = 0;
  for (int a = 0; a < 3; a++) ext = std::max(ext, hi[a] - lo[a]);
  if (!(ext > 0) || !(ext < 1e6)) return MPB_OK;
  const double h = std::max(ext / grid_div, 0.02);
  int gn[3];
  for (int a = 0; a < 3; a++) gn[a] = std::max(1, std::min((int)(grid_div * 1.25), (int)std::ceil((hi[a] - lo[a]) / h)));
  const size_t cells = (size_t)gn[0] * gn[1] * gn[2];
  std::vector<int> start;
  std::vector<unsigned short> list;
  const double pad = 1e-4;  // fp32 rounding of the moving centre near a cell face
  static const bool obb_filter = !(getenv("MPB_GRID_NO_OBB") && atoi(getenv("MPB_GRID_NO_OBB")));
  for (size_t ri = 0; ri < n_runs; ri++) {
    const int4 r = s->h_runs[ri];
    if (r.w != 0 || r.z == 0 || r.z > 65535) continue;
    std::vector<int> cnt(cells + 1, 0);
    std::vector<std::pair<int, unsigned short>> ins;  // (cell, entry)
    for (int e = 0; e < r.z; e++) {
      const float4 c = s->h_cullc[r.y + e];
      const double rr = std::sqrt((double)c.w) + pad;
      const double cc[3] = {c.x, c.y, c.z};
      // the partner's oriented box (world frame): the moving object lies inside the ball (its sphere centre, rm), so the pair
      // can only touch when that centre is within rm of the box -- for a long capsule or a flat box a much smaller region than
      // the ball around the partner's bounding sphere.  Tested at the cell centre with the cell's half diagonal as slack.
      const DevPair &pr = s->h_pairs[s->h_cullm[r.y + e].x];
      const int fo = s->h_objs[pr.a].dyn < 0 ? pr.a : pr.b;
      const float4 *box = s->h_cobb.data() + (size_t)MPB_COBB_WORDS * fo;
      const double rm = rr - (double)s->h_objs[fo].br + 0.5 * 1.7320508075688772 * h + 1e-4;
      int c0[3], c1[3];
      for (int a = 0; a < 3; a++) {
        c0[a] = std::max(0, (int)std::floor((cc[a] - rr - lo[a]) / h));
        c1[a] = std::min(gn[a] - 1, (int)std::floor((cc[a] + rr - lo[a]) / h));
      }
      for (int z = c0[2]; z <= c1[2]; z++)
        for (int y = c0[1]; y <= c1[1]; y++)
          for (int x = c0[0]; x <= c1[0]; x++) {
            const int idx[3] = {x, y, z};
            double d2 = 0;
            for (int a = 0; a < 3; a++) {
              const double b0 = lo[a] + idx[a] * h, b1 = b0 + h;
              const double d = cc[a] < b0 ? b0 - cc[a] : (cc[a] > b1 ? cc[a] - b1 : 0.0);
              d2 += d * d;
            }
            if (d2 > rr * rr) continue;
            if (obb_filter && r.x >= 0) {  // (a run of fixed x fixed pairs lives in difference coordinates: no box filter there)
              const double g[3] = {lo[0] + (x + 0.5) * h, lo[1] + (y + 0.5) * h, lo[2] + (z + 0.5) * h};
              const double d0 = g[0] - box[0].w, d1 = g[1] - box[1].w, d2g = g[2] - box[2].w;
              const double l0 = box[0].x * d0 + box[1].x * d1 + box[2].x * d2g, l1 = box[0].y * d0 + box[1].y * d1 + box[2].y * d2g,
                           l2 = box[0].z * d0 + box[1].z * d1 + box[2].z * d2g;
              const double e0 = std::max(std::fabs(l0) - box[3].x, 0.0), e1 = std::max(std::fabs(l1) - box[3].y, 0.0), e2 = std::max(std::fabs(l2) - box[3].z, 0.0);
              if (e0 * e0 + e1 * e1 + e2 * e2 > rm * rm) continue;
            }
            ins.push_back(std::make_pair((z * gn[1] + y) * gn[0] + x, (unsigned short)e));
          }
    }
    for (const auto &p : ins) cnt[(size_t)p.first + 1]++;
    for (size_t c = 0; c < cells; c++) cnt[c + 1] = cnt[c] + (cnt[c + 1] + 3) / 4 * 4;  // lists padded to multiples of four entries
    grun[ri] = make_int2((int)start.size(), (int)list.size());
    const size_t l0 = list.size();
    list.resize(l0 + (size_t)cnt[cells], (unsigned short)0xffff);
    std::vector<int> cur(cnt.begin(), cnt.end() - 1);
    for (const auto &p : ins) list[l0 + (size_t)cur[(size_t)p.first]++] = p.second;  // entries ascending within a cell (insertion order)
    start.insert(start.end(), cnt.begin(), cnt.end());
  }
  if (start.empty()) return MPB_OK;
  CU(ensure(&s->d_grid_start, &s->grid_start_cap, start.size()));
  CU(ensure(&s->d_grid_list, &s->grid_list_cap, std::max<size_t>(list.size(), 1)));
  CU(ensure(&s->d_grid_run, &s->grid_run_cap, grun.size()));
  CU(cudaMemcpy(s->d_grid_start, start.data(), start.size() * sizeof(int), cudaMemcpyHostToDevice));
  if (!list.empty()) CU(cudaMemcpy(s->d_grid_list, list.data(), list.size() * sizeof(unsigned short), cudaMemcpyHostToDevice));
  CU(cudaMemcpy(s->d_grid_run, grun.data(), grun.size() * sizeof(int2), cudaMemcpyHostToDevice));
  s->img.grid_start = s->d_grid_start;
  s->img.grid_list = s->d_grid_list;
  s->img.grid_run = s->d_grid_run;
  for (int a = 0; a < 3; a++) { s->img.grid_o[a] = (float)lo[a]; s->img.grid_n[a] = gn[a]; }
  s->img.grid_inv_h = (float)(1.0 / h);
  s->grid_on = true;
  return MPB_OK;
}

static int prepare_work(mpb_scene *s, long long chunk) {
  // grow-only: buffers sized for a larger pass serve a smaller one (collide calls of different sizes and planning
  // rounds alternate on one scene; re-allocating GBs of queues between them would dominate small calls)
  if (int grc = ensure_grid(s)) return grc;
  const size_t P = (size_t)std::max(s->img.n_pairs, 1);
  if (chunk <= s->chunk && s->wq.rec[0] && P <= s->work_P && (size_t)std::max(std::max(s->img.n_bins, s->img.n_bins2), 1) <= s->bins_cap) return MPB_OK;
  if (int hrc = harvest(s, true)) return hrc;  // calls in flight (device-pointer mode) still use the old buffers
  CU(sync_scene(s));
  chunk = std::max(chunk, s->chunk);
  s->work_P = P;
  // staging may hold up to one unused half-chunk per reservation: worst case 2 x items + one chunk per warp
  // plus one partly used reservation per warp of dense_kernel / expand_kernel
  size_t cap = std::min((size_t)chunk * P * 2 + (size_t)s->num_sms * 8 * 4 * MPB_STAGE_CHUNK, kMaxItems);
  cap = std::max<size_t>(cap, 4096);
  if (s->wq.rec[0]) { cudaFree(s->wq.rec[0]); s->wq.rec[0] = s->wq.rec[1] = nullptr; }
  if (s->d_pose) { cudaFree(s->d_pose); s->d_pose = nullptr; }
  if (s->d_sq) { cudaFree(s->d_sq); s->d_sq = nullptr; }
  s->pose_ld = (size_t)(chunk + 31) / 32 * 32;
  const size_t pose_words = (size_t)s->img.n_jslots * 12 + (size_t)s->img.n_dyn * 3;
  CU(cudaMalloc((void **)&s->d_pose, std::max<size_t>(pose_words * s->pose_ld, 4) * sizeof(float)));
  if (s->d_obb) { cudaFree(s->d_obb); s->d_obb = nullptr; }
  CU(cudaMalloc((void **)&s->d_obb, std::max<size_t>((size_t)s->img.n_dyn * 2 * s->pose_ld, 1) * sizeof(float4)));
  // survivors of the sphere phase: every admitted pair of every configuration at worst, bounded like the work queue
  const size_t sq_cap = std::max<size_t>(std::min((size_t)chunk * P + (size_t)s->num_sms * 64 * MPB_SQ_CHUNK_GRID, kMaxSurvivors), 4096);
  CU(cudaMalloc((void **)&s->d_sq, sq_cap * sizeof(uint2)));
  s->sq_cap = (unsigned)sq_cap;
  // one slab: item records of the two rounds (64 bytes each), binned order, bin keys
  CU(cudaMalloc((void **)&s->wq.rec[0], cap * (2 * 64 + sizeof(unsigned) + sizeof(unsigned short))));
  s->wq.rec[1] = s->wq.rec[0] + cap * 4;
  s->wq.order = reinterpret_cast<unsigned *>(s->wq.rec[1] + cap * 4);
  s->wq.skey = reinterpret_cast<unsigned short *>(s->wq.order + cap);
  s->wq.cap = (unsigned)cap;
  const size_t nb = (size_t)std::max(std::max(s->img.n_bins, s->img.n_bins2), 1);
  if (nb > s->bins_cap || !s->wq.bin_count) {
    if (s->wq.bin_count) cudaFree(s->wq.bin_count);
    CU(cudaMalloc((void **)&s->wq.bin_count, (4 * nb + 8) * sizeof(unsigned)));
    s->bins_cap = nb;
  }
  s->wq.bin_cursor = s->wq.bin_count + nb;
  s->wq.bin_start = s->wq.bin_cursor + nb;
  s->wq.unit_start = s->wq.bin_start + nb + 1;
  if (!s->wq.totals) {  // totals[8] and counters[4] share one 64-byte block: one D2H copy brings both
    CU(cudaMalloc((void **)&s->wq.totals, 8 * sizeof(unsigned) + 4 * sizeof(unsigned long long)));
    s->wq.counters = reinterpret_cast<unsigned long long *>(s->wq.totals + 8);
  }
  s->chunk = chunk;
  return MPB_OK;
}

static cudaEvent_t get_event(mpb_scene *s, size_t i) {
  while (s->events.size() <= i) {
    cudaEvent_t e;
    cudaEventCreate(&e);
    s->events.push_back(e);
  }
  return s->events[i];
}

static cudaEvent_t get_sync_event(mpb_scene *s, size_t i) {
  while (s->sync_events.size() <= i) {
    cudaEvent_t e;
    cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
    s->sync_events.push_back(e);
  }
  return s->sync_events[i];
}

// Result sink of one pass (see HitOut): `hits` / `n_hits` are call-wide, `cfg0` is the pass's offset in the call.
struct PassOut {
  uint32_t *dv = nullptr, *dp = nullptr;
  uint2 *hits = nullptr;
  unsigned *n_hits = nullptr;
  unsigned hit_cap = 0, cfg0 = 0;
  bool whole_pass = false;  // host input: copy q of the pass in one piece (multi-pass calls)
  bool reset_call = false;  // small passes: the fused reset also zeroes the call's counters (first pass of a graph-captured call)
};

// Timing events of one pass: [0] start, [1] FK + cull done, [2] narrowphase done, [3], [4] around fk_table_kernel.
struct PassEvents {
  cudaEvent_t e[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  bool on = false;
};

// Kernels of one pass over m configurations.  hq == nullptr: q already sits at dq.  Otherwise the pass's q is
// copied from the host pointer hq in sub-ranges on stream_h2d and FK + cull runs per sub-range as it lands, all
// appending to the same staging queue; binning and narrowphase then run once over the whole pass.
static int launch_pass(mpb_scene *s, float *dq_w, const float *dq, long long m, const PassOut &po, cudaStream_t st,
                       const PassEvents &pe, long long *launches, const float *hq, size_t *sync_ev) {
  const int D = s->img.state_dim, W = s->img.pair_words, nb = std::max(std::max(s->img.n_bins, s->img.n_bins2), 1);
  const size_t vwords = (size_t)(m + 31) / 32;
  uint32_t *dv = po.dv, *dp = po.dp;
  HitOut out;
  out.verdict = dv; out.pair_bits = dp;
  out.pair_words = W;
  const bool small = m <= kSmallBatch;
  // upper bound of the items any queue of this pass can hold: sizes the grids of small passes (every kernel below is
  // grid-stride or ticket-driven, so a smaller grid is only less parallel, never wrong)
  const long long work = m * (long long)std::max(s->img.n_pairs, 1);
  auto grid_for = [&](int full, long long per_block) { return small ? (int)std::max<long long>(1, std::min<long long>(full, (work + per_block - 1) / per_block)) : full; };
  if (small) {
    const unsigned pw = dp ? (unsigned)(m * W) : 0u;
    const int rb = (int)std::max<size_t>(1, std::min<size_t>(64, ((size_t)pw + vwords + 4095) / 4096));
    small_reset_kernel<<<rb, 256, 0, st>>>(s->wq, nb, dv, (unsigned)vwords, dp, pw, po.n_hits, po.reset_call ? 1 : 0);
  } else {
    CU(cudaMemsetAsync(dv, 0, vwords * sizeof(uint32_t), st));
    if (dp) CU(cudaMemsetAsync(dp, 0, (size_t)m * W * sizeof(uint32_t), st));
    pass_reset_kernel<<<1, 256, 0, st>>>(s->wq, nb);
  }
  const bool ev = pe.on;
  cudaEvent_t e0 = pe.e[0], e1 = pe.e[1], e2 = pe.e[2], f0 = pe.e[3], f1 = pe.e[4];
  if (ev) CU(cudaEventRecord(e0, st));
  if (ev && hq) { CU(cudaEventRecord(f0, st)); CU(cudaEventRecord(f1, st)); }
  // host input: q arrives in sub-ranges so that FK + cull of one overlaps the copy of the next.  A call of several passes
  // already overlaps the copy of pass c+1 with the kernels of pass c (double buffering), so its passes take q whole:
  // the thread-per-configuration cull kernels need ~10^6 configurations to fill the GPU
  // fk_sphere_kernel cuts the DRAM traffic of a C2 pass from ~790 MB to ~290 MB but FK is issue-bound (fk_table_kernel runs at
  // 74 % issue-active), and the second walk of the chain costs more than the writes it saves: C2 0.365 -> 0.377 ms, C5 3.92 ->
  // 4.57 ms per step.  Off unless asked for.
  static const int lazy_env = getenv("MPB_LAZY_FK") && atoi(getenv("MPB_LAZY_FK"));
  const bool lazy = lazy_env && !small && s->cull_tier == 1 && s->fk_sphere_blocks > 0;
  const long long sub = hq ? (po.whole_pass ? m : std::max<long long>(s->host_chunk, 64)) : m;
  for (long long off = 0; off < m; off += sub) {
    const long long mm = std::min(sub, m - off);
    if (hq) {
      if (D > 0) CU(cudaMemcpyAsync(dq_w + off * D, hq + off * D, (size_t)mm * D * sizeof(float), cudaMemcpyHostToDevice, s->stream_h2d));
      cudaEvent_t landed = get_sync_event(s, (*sync_ev)++);
      CU(cudaEventRecord(landed, s->stream_h2d));
      CU(cudaStreamWaitEvent(st, landed, 0));
    }
    if (s->img.n_pairs > 0 && lazy) {
      // parameter-space tier: FK + sphere phase fused, no sphere table, OBB rows only for objects that survivors name
      const int cblocks = (int)((mm + 127) / 128);
      if (ev && !hq) { CU(cudaEventRecord(f0, st)); CU(cudaEventRecord(f1, st)); }
      fk_sphere_kernel<<<std::min(cblocks, s->fk_sphere_blocks), 128, s->fk_sphere_smem, st>>>(s->img, s->fk_tab, s->ct_small, s->d_qconst, dq + off * D, mm, off, s->jneed,
                                                                                              s->d_obb, s->pose_ld, s->d_sq, s->sq_cap, s->wq.totals);
      *launches += 1;
    } else if (s->img.n_pairs > 0) {
      const int cblocks = (int)((mm + 127) / 128);
      if (ev && !hq) CU(cudaEventRecord(f0, st));
      fk_table_kernel<<<cblocks, 128, 0, st>>>(s->img, s->fk_tab, s->d_qconst, dq + off * D, mm, s->d_pose + off, s->pose_ld, s->d_obb + (size_t)off * 2);
      if (ev && !hq) CU(cudaEventRecord(f1, st));
      const int sblocks = std::min(cblocks, s->sphere_blocks[0]);  // persistent warps: survivor-queue chunks carry over between tiles
      if (small)
        sphere_pairs_kernel<<<(int)std::max<long long>(1, std::min<long long>((mm * (long long)s->img.n_pairs + 255) / 256, (long long)s->num_sms * 16)), 256, 0, st>>>(
            s->img, s->d_pose + off, s->pose_ld, mm, off, s->d_sq, s->sq_cap, s->wq.totals);
      else if (s->grid_on)
        sphere_grid_kernel<<<std::min(cblocks, s->sphere_blocks_grid), 128, 0, st>>>(s->img, s->d_pose + off, s->pose_ld, mm, off, s->d_sq, s->sq_cap, s->wq.totals);
      else if (s->cull_tier == 1)
        sphere_kernel<0, 256, 64><<<std::min(cblocks, s->sphere_blocks_ct[0]), 128, (size_t)s->img.n_dyn * 3 * 128 * sizeof(float), st>>>(s->img, s->ct_small, s->d_pose + off, s->pose_ld, mm, off, s->d_sq,
                                                                                              s->sq_cap, s->wq.totals);
      else if (s->sphere_cb == 2)
        sphere_blk_kernel<2><<<std::min((cblocks + 1) / 2, s->sphere_blocks_blk[0]), 128, 0, st>>>(s->img, s->d_pose + off, s->pose_ld, mm, off, s->d_sq, s->sq_cap, s->wq.totals);
      else if (s->sphere_cb == 4)
        sphere_blk_kernel<4><<<std::min((cblocks + 3) / 4, s->sphere_blocks_blk[1]), 128, 0, st>>>(s->img, s->d_pose + off, s->pose_ld, mm, off, s->d_sq, s->sq_cap, s->wq.totals);
      else
        sphere_kernel<0, 0, 0><<<sblocks, 128, 0, st>>>(s->img, CullTabC<0, 0>{}, s->d_pose + off, s->pose_ld, mm, off, s->d_sq, s->sq_cap, s->wq.totals);
      *launches += 2;
    }
  }
  *launches += 1;
  if (s->img.n_pairs == 0) {
    if (ev) {
      if (!hq) { CU(cudaEventRecord(f0, st)); CU(cudaEventRecord(f1, st)); }
      CU(cudaEventRecord(e1, st));
      CU(cudaEventRecord(e2, st));
    }
    return MPB_OK;
  }
  // the sufficient-hit test pays where collisions are common and deep: scenes with primitive obstacles / octrees; hull x hull
  // pairs of a robot against itself (C2, C5) seldom overlap by more than a few millimetres
  static const int presolve_env = !(getenv("MPB_NO_PRESOLVE") && atoi(getenv("MPB_NO_PRESOLVE")));
  const int presolve = presolve_env && s->has_primitives;
  static const int dense_v2 = !(getenv("MPB_DENSE_V1") && atoi(getenv("MPB_DENSE_V1")));
  if (dense_v2) dense2_kernel<<<grid_for(s->dense2_blocks, 128), 128, 0, st>>>(s->img, s->d_obb, s->pose_ld, s->d_sq, s->sq_cap, s->wq, out, presolve);
  else dense_kernel<0><<<grid_for(s->dense_blocks[0], 128), 128, 0, st>>>(s->img, s->d_obb, s->pose_ld, s->d_sq, s->sq_cap, s->wq, out, presolve);
  *launches += 1;
  static const int refill_min = getenv("MPB_REFILL_MIN") ? atoi(getenv("MPB_REFILL_MIN")) : 8;
  static const double unit_factor = getenv("MPB_UNIT_FACTOR") ? atof(getenv("MPB_UNIT_FACTOR")) : 1.0;
  const int nblocks = grid_for(s->narrow_blocks, 64);  // narrowphase: a block of 8 warps per 64 possible items of a small pass
  const unsigned n_warps = (unsigned)std::max(1.0, nblocks * 8.0 * unit_factor);
  scan_bins_kernel<<<1, 32, 0, st>>>(s->wq, s->img.n_bins, s->img.expand_bin, s->img.expand2_bin, n_warps, s->sq_cap);
  scatter_kernel<false><<<grid_for(s->num_sms * 2, 1024), 1024, 0, st>>>(s->img, s->wq, 0, out);  // round 0: dense_kernel ran the sufficient test
  if (ev) CU(cudaEventRecord(e1, st));
  const size_t vsm = s->smem_verts ? (size_t)s->img.pool_bytes : 0;
  static const int narrow_v2 = !(getenv("MPB_NARROW_V1") && atoi(getenv("MPB_NARROW_V1")));
  const int nblocks2 = grid_for(s->narrow2_blocks, 64);
  if (narrow_v2) {
    if (s->smem_verts) narrow2_kernel<true><<<nblocks2, 256, s->narrow2_smem, st>>>(s->img, s->wq, out, s->img.n_bins, 0, refill_min, 0);
    else narrow2_kernel<false><<<nblocks2, 256, s->narrow2_smem, st>>>(s->img, s->wq, out, s->img.n_bins, 0, refill_min, 0);
  } else if (s->smem_verts) narrow_kernel<true><<<nblocks, 256, vsm, st>>>(s->img, s->wq, out, s->img.n_bins, 0, refill_min);
  else narrow_kernel<false><<<nblocks, 256, 0, st>>>(s->img, s->wq, out, s->img.n_bins, 0, refill_min);
  *launches += 3;
  if (s->img.expand2_bin >= 0) {  // mesh x mesh / mesh x octree pairs: exact leaf tests inside the descent
    expand2_kernel<<<grid_for(s->num_sms * 8, 4), 128, 0, st>>>(s->img, s->wq, out);
    *launches += 1;
  }
  if (s->img.expand_bin >= 0) {  // shape x mesh / octree pairs: expand to leaf items, second narrowphase round
    pass_reset_kernel<<<1, 256, 0, st>>>(s->wq, nb);
    static const int lane_on = !(getenv("MPB_EXPAND_WARP") && atoi(getenv("MPB_EXPAND_WARP")));
    expand_kernel<<<grid_for(s->num_sms * 8, 16), 128, 0, st>>>(s->img, s->wq, out, small ? 0 : lane_on);
    if (!small && lane_on) { expand_lane_kernel<<<s->expand_lane_blocks, 128, 0, st>>>(s->img, s->wq, out); *launches += 1; }
    scan_bins_kernel<<<1, 32, 0, st>>>(s->wq, s->img.n_bins2, -1, -1, (unsigned)std::max(1.0, s->narrow_blocks * 8.0 * unit_factor), s->sq_cap);
    if (presolve) scatter_kernel<true><<<s->num_sms * 4, 512, 0, st>>>(s->img, s->wq, 1, out);  // leaf items: inner balls vs voxel box
    else scatter_kernel<false><<<s->num_sms * 2, 1024, 0, st>>>(s->img, s->wq, 1, out);
    if (narrow_v2) {
      if (s->smem_verts) narrow2_kernel<true><<<s->narrow2_blocks, 256, s->narrow2_smem, st>>>(s->img, s->wq, out, s->img.n_bins2, 1, refill_min, 0);
      else narrow2_kernel<false><<<s->narrow2_blocks, 256, s->narrow2_smem, st>>>(s->img, s->wq, out, s->img.n_bins2, 1, refill_min, 0);
    } else if (s->smem_verts) narrow_kernel<true><<<s->narrow_blocks, 256, vsm, st>>>(s->img, s->wq, out, s->img.n_bins2, 1, refill_min);
    else narrow_kernel<false><<<s->narrow_blocks, 256, 0, st>>>(s->img, s->wq, out, s->img.n_bins2, 1, refill_min);
    *launches += 5;
  }
  if (po.hits && dp && W > 0) {  // sparse hit list: compacted out of the pass's dense mask
    const long long n_words = m * (long long)W;
    const int blocks = (int)std::max<long long>(1, std::min<long long>((n_words + 1023) / 1024, (long long)s->num_sms * 8));
    hits_compact_kernel<<<blocks, 256, 0, st>>>(dp, n_words, W, po.cfg0, po.hits, po.n_hits, po.hit_cap);
    *launches += 1;
  }
  if (ev) CU(cudaEventRecord(e2, st));
  CU(cudaGetLastError());
  return MPB_OK;
}

// What a caller wants back per (configuration, pair): nothing, the dense mask, or the sparse hit list.
struct CollideOut {
  uint32_t *verdict = nullptr;
  uint32_t *pair_bits = nullptr;
  uint32_t *hits = nullptr;      // [hit_cap][2]
  int64_t hit_cap = 0;
  uint32_t *n_hits = nullptr;    // [1], same space as hits
};

static void fill_stats(const mpb_scene *s, mpb_stats *stats, long long n, const unsigned long long *counters, long long launches) {
  stats->n_states = n;
  stats->pairs_admitted = n * (long long)s->img.n_pairs;
  stats->pairs_after_cull = (long long)counters[0];
  stats->pairs_after_spheres = (long long)counters[3];
  stats->narrow_iterations = (long long)counters[1];
  stats->narrow_capped = (long long)counters[2];
  stats->kernel_launches = launches;
  stats->fk_bytes_per_state = (int32_t)(4 * s->img.state_dim + 4 * (s->img.n_jslots * 12 + s->img.n_dyn * 3) + 32 * s->img.n_dyn);
}

static void add_pass_times(const cudaEvent_t *e, float *cull, float *narrow, float *fk) {
  float a = 0.f, b = 0.f, c = 0.f;
  cudaEventElapsedTime(&a, e[0], e[1]);
  cudaEventElapsedTime(&b, e[1], e[2]);
  cudaEventElapsedTime(&c, e[3], e[4]);
  *cull += a;
  *narrow += b;
  *fk += c;
}

// One blocking attempt over the whole batch with a given chunk size.  Returns MPB_OK and sets *overflow.
// Host buffers: chunk c uses device buffer set c & 1; its H2D runs on stream_h2d, its kernels on `st`, its D2H on
// stream_d2h, so the copies of neighbouring chunks overlap the kernels (PCIe is full duplex).
static int collide_attempt(mpb_scene *s, const float *q, long long n, const CollideOut &co, int space,
                           cudaStream_t st, long long chunk, bool *overflow, mpb_stats *stats) {
  const auto t_begin = std::chrono::steady_clock::now();
  int rc = prepare_work(s, chunk);
  if (rc != MPB_OK) return rc;
  const int D = s->img.state_dim, W = s->img.pair_words;
  const bool host = space == MPB_HOST;
  const bool want_hits = co.hits != nullptr && co.n_hits != nullptr;
  const bool want_mask = co.pair_bits != nullptr || want_hits;
  if (host) {
    for (int b = 0; b < 2; b++) {
      CU(ensure(&s->d_q[b], &s->d_q_cap[b], (size_t)chunk * std::max(D, 1)));
      CU(ensure(&s->d_verdict[b], &s->d_verdict_cap[b], (size_t)(chunk + 31) / 32));
      if (want_mask) CU(ensure(&s->d_pairbits[b], &s->d_pairbits_cap[b], (size_t)chunk * std::max(W, 1)));
    }
    if (!s->stream_h2d) CU(cudaStreamCreateWithFlags(&s->stream_h2d, cudaStreamNonBlocking));
    if (!s->stream_d2h) CU(cudaStreamCreateWithFlags(&s->stream_d2h, cudaStreamNonBlocking));
    if (!s->stream_hits) CU(cudaStreamCreateWithFlags(&s->stream_hits, cudaStreamNonBlocking));
  } else if (want_hits) {
    CU(ensure(&s->d_pairbits[0], &s->d_pairbits_cap[0], (size_t)chunk * std::max(W, 1)));  // the dense mask stays on the device
  }
  uint2 *d_hits = nullptr;
  unsigned *d_nhits = nullptr;
  unsigned hit_cap = 0;
  if (want_hits) {
    hit_cap = (unsigned)std::min<int64_t>(std::max<int64_t>(co.hit_cap, 0), (int64_t)1 << 30);
    if (host) {
      CU(ensure(&s->d_hits, &s->d_hits_cap, (size_t)hit_cap + 2));
      d_hits = s->d_hits + 1;                                  // [0] holds the counter
      d_nhits = reinterpret_cast<unsigned *>(s->d_hits);
    } else {
      d_hits = reinterpret_cast<uint2 *>(co.hits);
      d_nhits = co.n_hits;
    }
    CU(cudaMemsetAsync(d_nhits, 0, sizeof(unsigned), st));
  }
  CU(cudaMemsetAsync(s->wq.totals, 0, 8 * sizeof(unsigned), st));
  CU(cudaMemsetAsync(s->wq.counters, 0, 4 * sizeof(unsigned long long), st));
  size_t ev = 0, sync_ev = 0;
  long long launches = 0, c = 0;
  if (host) {  // the copy stream starts after whatever the caller already queued on `st`
    cudaEvent_t start = get_sync_event(s, sync_ev++);
    CU(cudaEventRecord(start, st));
    CU(cudaStreamWaitEvent(s->stream_h2d, start, 0));
  }
  std::vector<cudaEvent_t> ev_k, ev_out, ev_cnt;  // per pass: kernels done, results copied out, hit count known
  unsigned hits_copied = 0;
  for (long long off = 0; off < n; off += chunk, c++) {
    const long long m = std::min(chunk, n - off);
    const size_t vwords = (size_t)(m + 31) / 32;
    PassEvents pe;
    pe.on = true;
    for (int k = 0; k < 5; k++) pe.e[k] = get_event(s, ev++);
    PassOut po;
    po.hits = d_hits; po.n_hits = d_nhits; po.hit_cap = hit_cap; po.cfg0 = (unsigned)off;
    if (!host) {
      po.dv = co.verdict + off / 32;
      po.dp = co.pair_bits ? co.pair_bits + (size_t)off * W : (want_hits ? s->d_pairbits[0] : nullptr);
      rc = launch_pass(s, nullptr, q + off * D, m, po, st, pe, &launches, nullptr, &sync_ev);
      if (rc != MPB_OK) return rc;
      continue;
    }
    const int b = (int)(c & 1);
    if (c >= 2) {
      CU(cudaStreamWaitEvent(s->stream_h2d, ev_k[c - 2], 0));  // kernels of pass c-2 read d_q[b]
      CU(cudaStreamWaitEvent(st, ev_out[c - 2], 0));           // D2H of pass c-2 read the result buffers
    }
    po.dv = s->d_verdict[b];
    po.dp = want_mask ? s->d_pairbits[b] : nullptr;
    po.whole_pass = n > chunk;
    rc = launch_pass(s, s->d_q[b], s->d_q[b], m, po, st, pe, &launches, q + off * D, &sync_ev);
    if (rc != MPB_OK) return rc;
    ev_k.push_back(get_sync_event(s, sync_ev++));
    ev_out.push_back(get_sync_event(s, sync_ev++));
    CU(cudaEventRecord(ev_k[c], st));
    CU(cudaStreamWaitEvent(s->stream_d2h, ev_k[c], 0));
    CU(cudaMemcpyAsync(co.verdict + off / 32, po.dv, vwords * sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream_d2h));
    if (co.pair_bits)
      CU(cudaMemcpyAsync(co.pair_bits + (size_t)off * W, po.dp, (size_t)m * W * sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream_d2h));
    CU(cudaEventRecord(ev_out[c], s->stream_d2h));
    if (want_hits && c < kMaxHitPasses) {
      // hit list: the counter after this pass lands in a pinned slot; once the PREVIOUS pass's count is known (this pass
      // is already queued behind it, so the GPU stays busy) its share of the list is copied out under this pass's kernels
      CU(cudaMemcpyAsync(s->h_hitcnt + c, d_nhits, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
      ev_cnt.push_back(get_sync_event(s, sync_ev++));
      CU(cudaEventRecord(ev_cnt[c], st));
      if (c >= 1) {
        CU(cudaEventSynchronize(ev_cnt[c - 1]));
        const unsigned upto = std::min(s->h_hitcnt[c - 1], hit_cap);
        if (upto > hits_copied) {
          // its own stream: stream_d2h already waits for THIS pass's kernels, and the share would be copied a pass late
          CU(cudaMemcpyAsync(co.hits + 2 * (size_t)hits_copied, d_hits + hits_copied, (size_t)(upto - hits_copied) * sizeof(uint2), cudaMemcpyDeviceToHost, s->stream_hits));
          hits_copied = upto;
        }
      }
    }
  }
  const auto t_enq = std::chrono::steady_clock::now();
  unsigned *totals = s->h_status;                                                        // pinned
  unsigned long long *counters = reinterpret_cast<unsigned long long *>(s->h_status + 8);
  CU(cudaMemcpyAsync(totals, s->wq.totals, 8 * sizeof(unsigned) + 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
  if (want_hits && host) CU(cudaMemcpyAsync(s->h_status + 16, d_nhits, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  *overflow = totals[2] != 0;
  if (want_hits && host && !*overflow) {
    const unsigned nh = s->h_status[16];
    *co.n_hits = nh;
    const unsigned upto = std::min(nh, hit_cap);
    if (upto > hits_copied)
      CU(cudaMemcpyAsync(co.hits + 2 * (size_t)hits_copied, d_hits + hits_copied, (size_t)(upto - hits_copied) * sizeof(uint2), cudaMemcpyDeviceToHost, s->stream_hits));
  }
  if (host) CU(cudaStreamSynchronize(s->stream_d2h));
  if (host && want_hits) CU(cudaStreamSynchronize(s->stream_hits));
  const auto t_end = std::chrono::steady_clock::now();
  if (!*overflow) {  // what the device-pointer fast path sizes its passes from: the largest pass size known to fit
    s->good_chunk = s->calibrated ? std::max(s->good_chunk, chunk) : chunk;
    s->calibrated = true;
  } else {
    s->cap_chunk = std::min(s->cap_chunk, std::max<long long>(32, (chunk / 4 + 31) / 32 * 32));  // passes of this size overflow
  }
  if (stats) {
    fill_stats(s, stats, n, counters, launches);
    float cull = 0.f, narrow = 0.f, fk = 0.f;
    for (size_t i = 0; i + 5 <= ev; i += 5) add_pass_times(&s->events[i], &cull, &narrow, &fk);
    stats->gpu_ms_fk = fk;
    stats->gpu_ms_cull = cull;
    stats->gpu_ms_narrow = narrow;
    stats->gpu_ms = cull + narrow;
    stats->host_ms_enqueue = std::chrono::duration<float, std::milli>(t_enq - t_begin).count();
    stats->host_ms_total = std::chrono::duration<float, std::milli>(t_end - t_begin).count();
  }
  return MPB_OK;
}

// ---- device-pointer calls that return after enqueueing -------------------------------------------------------------
// The queue counters of a call are snapshotted by an in-stream copy into a pinned slot of a small ring; a later call
// (or mpb_scene_poll) harvests the finished slots: statistics are accumulated, an overflowed call is recorded as
// failed.  The passes are sized from the last blocking call on the scene (the calibration), which is why the first
// device-pointer call after a scene / pair-list change runs through the blocking protocol.
static void harvest_slot(mpb_scene *s, AsyncSlot &sl) {
  const unsigned *totals = sl.h;
  const unsigned long long *counters = reinterpret_cast<const unsigned long long *>(sl.h + 8);
  mpb_stats st;
  memset(&st, 0, sizeof(st));
  fill_stats(s, &st, sl.n, counters, sl.launches);
  for (size_t i = 0; i + 5 <= sl.n_ev; i += 5) add_pass_times(&sl.ev[i], &st.gpu_ms_cull, &st.gpu_ms_narrow, &st.gpu_ms_fk);
  st.gpu_ms = st.gpu_ms_cull + st.gpu_ms_narrow;
  mpb_stats &a = s->async_acc;
  a.n_states += st.n_states; a.pairs_admitted += st.pairs_admitted; a.pairs_after_cull += st.pairs_after_cull;
  a.pairs_after_spheres += st.pairs_after_spheres; a.narrow_iterations += st.narrow_iterations; a.narrow_capped += st.narrow_capped;
  a.kernel_launches += st.kernel_launches; a.gpu_ms += st.gpu_ms; a.gpu_ms_cull += st.gpu_ms_cull; a.gpu_ms_narrow += st.gpu_ms_narrow;
  a.gpu_ms_fk += st.gpu_ms_fk; a.host_ms_enqueue += sl.host_ms; a.host_ms_total += sl.host_ms; a.fk_bytes_per_state = st.fk_bytes_per_state;
  s->async_completed++;
  if (totals[2]) { s->async_failed++; s->calibrated = false; s->good_chunk = 0; }  // the next call re-calibrates through the blocking protocol
  sl.pending = false;
}

static int harvest(mpb_scene *s, bool wait) {
  for (int k = 0; k < MPB_ASYNC_SLOTS; k++) {
    AsyncSlot &sl = s->slots[(s->slot_head + k) % MPB_ASYNC_SLOTS];  // oldest first
    if (!sl.pending) continue;
    if (wait) CU(cudaEventSynchronize(sl.done));
    else {
      cudaError_t e = cudaEventQuery(sl.done);
      if (e == cudaErrorNotReady) break;  // slots complete in order
      if (e != cudaSuccess) return fail(MPB_ERR_CUDA, std::string("cudaEventQuery: ") + cudaGetErrorString(e));
    }
    harvest_slot(s, sl);
  }
  return MPB_OK;
}

static int collide_enqueue(mpb_scene *s, const float *q, long long n, const CollideOut &co, cudaStream_t st, mpb_stats *stats) {
  const auto t_begin = std::chrono::steady_clock::now();
  const int D = s->img.state_dim, W = s->img.pair_words;
  const long long chunk = std::min<long long>(std::min<long long>(max_chunk_for(s), s->cap_chunk), (n + 31) / 32 * 32);
  int rc = prepare_work(s, chunk);
  if (rc != MPB_OK) return rc;
  const bool want_hits = co.hits != nullptr && co.n_hits != nullptr;
  if (want_hits && !co.pair_bits) CU(ensure(&s->d_pairbits[0], &s->d_pairbits_cap[0], (size_t)chunk * std::max(W, 1)));
  if (!s->h_slots) {
    CU(cudaHostAlloc((void **)&s->h_slots, (size_t)MPB_ASYNC_SLOTS * 32 * sizeof(unsigned), cudaHostAllocDefault));
    for (int k = 0; k < MPB_ASYNC_SLOTS; k++) {
      s->slots[k].h = s->h_slots + 32 * k;
      CU(cudaEventCreateWithFlags(&s->slots[k].done, cudaEventDisableTiming));
    }
  }
  AsyncSlot &sl = s->slots[s->slot_head];
  if (sl.pending) {  // ring full: wait for the oldest call
    CU(cudaEventSynchronize(sl.done));
    harvest_slot(s, sl);
  }
  s->slot_head = (s->slot_head + 1) % MPB_ASYNC_SLOTS;
  unsigned hit_cap = 0;
  if (want_hits) {
    hit_cap = (unsigned)std::min<int64_t>(std::max<int64_t>(co.hit_cap, 0), (int64_t)1 << 30);
    CU(cudaMemsetAsync(co.n_hits, 0, sizeof(unsigned), st));
  }
  CU(cudaMemsetAsync(s->wq.totals, 0, 8 * sizeof(unsigned), st));
  CU(cudaMemsetAsync(s->wq.counters, 0, 4 * sizeof(unsigned long long), st));
  const bool timing = s->async_timing;
  size_t ev = 0, sync_ev = 0;
  long long launches = 0;
  for (long long off = 0; off < n; off += chunk) {
    const long long m = std::min(chunk, n - off);
    PassEvents pe;
    pe.on = timing;
    if (timing) {
      while (sl.ev.size() < ev + 5) { cudaEvent_t e; CU(cudaEventCreate(&e)); sl.ev.push_back(e); }
      for (int k = 0; k < 5; k++) pe.e[k] = sl.ev[ev++];
    }
    PassOut po;
    po.dv = co.verdict + off / 32;
    po.dp = co.pair_bits ? co.pair_bits + (size_t)off * W : (want_hits ? s->d_pairbits[0] : nullptr);
    po.hits = reinterpret_cast<uint2 *>(co.hits); po.n_hits = co.n_hits; po.hit_cap = hit_cap; po.cfg0 = (unsigned)off;
    rc = launch_pass(s, nullptr, q + off * D, m, po, st, pe, &launches, nullptr, &sync_ev);
    if (rc != MPB_OK) return rc;
  }
  CU(cudaMemcpyAsync(sl.h, s->wq.totals, 8 * sizeof(unsigned), cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(sl.h + 8, s->wq.counters, 4 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
  CU(cudaEventRecord(sl.done, st));
  sl.pending = true;
  sl.n = n;
  sl.launches = launches;
  sl.n_ev = ev;
  sl.host_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
  if (stats) {
    stats->n_states = n;
    stats->pairs_admitted = n * (long long)s->img.n_pairs;
    stats->kernel_launches = launches;
    stats->host_ms_enqueue = stats->host_ms_total = sl.host_ms;
    stats->async_pending = 1;
  }
  return MPB_OK;
}

extern "C" int mpb_scene_poll(mpb_scene *s, int wait, mpb_stats *sum, int64_t *calls_completed, int64_t *calls_failed) {
  if (!s) return fail(MPB_ERR_INVALID, "null scene");
  CU(cudaSetDevice(s->device));
  int rc = harvest(s, wait != 0);
  if (rc != MPB_OK) return rc;
  if (sum) *sum = s->async_acc;
  if (calls_completed) *calls_completed = s->async_completed;
  if (calls_failed) *calls_failed = s->async_failed;
  const long long failed = s->async_failed;
  memset(&s->async_acc, 0, sizeof(s->async_acc));
  s->async_completed = 0;
  s->async_failed = 0;
  if (failed) return fail(MPB_ERR_CAPACITY, std::to_string(failed) + " device-pointer call(s) overflowed the work queues: their outputs are incomplete, repeat them");
  return MPB_OK;
}

extern "C" int mpb_scene_set_async_timing(mpb_scene *s, int on) {
  if (!s) return fail(MPB_ERR_INVALID, "null scene");
  s->async_timing = on != 0;
  return MPB_OK;
}

// ---- small host batches: the whole call (H2D, kernels, D2H) replayed as one CUDA graph -----------------------------
// A legacy single-state query (collide(), Planner.check_for_*_collision, the isValid stub of INTEGRATION.md) is a batch
// of one: a dozen launches and copies whose enqueue cost, not their execution, sets the latency.  Calls of at most
// kSmallBatch states go through pinned staging buffers and a graph cached per (n, outputs) shape.

static void drop_small_graphs(mpb_scene *s) {
  for (auto &g : s->small_graphs) {
    if (g.exec) cudaGraphExecDestroy(g.exec);
    if (g.graph) cudaGraphDestroy(g.graph);
  }
  s->small_graphs.clear();
}

static int collide_small(mpb_scene *s, const float *q, long long n, const CollideOut &co, bool *overflow, mpb_stats *stats) {
  const auto t_begin = std::chrono::steady_clock::now();
  const int D = s->img.state_dim, W = s->img.pair_words;
  const bool want_hits = co.hits != nullptr && co.n_hits != nullptr;
  const int mode = want_hits ? 2 : (co.pair_bits ? 1 : 0);
  const long long chunk = (n + 31) / 32 * 32;
  cudaStream_t st = s->stream;
  const size_t vwords = (size_t)(n + 31) / 32;
  // the graph copies a bounded hit list; a call that produces more goes through the chunked path
  const unsigned user_cap = want_hits ? (unsigned)std::min<int64_t>(std::max<int64_t>(co.hit_cap, 0), (int64_t)1 << 30) : 0;
  const unsigned hit_cap = std::min(user_cap, 65536u);
  // output block, same layout on the device and in the pinned staging buffer:
  //   [n_hits, pad to 16 B | verdict words, padded to 16 B | dense mask (mode 1) or hit list (mode 2)]
  const size_t q_bytes = ((size_t)n * std::max(D, 1) * sizeof(float) + 15) / 16 * 16;
  const size_t v_off = 16, v_bytes = (vwords * 4 + 15) / 16 * 16;
  const size_t p_off = v_off + v_bytes;
  const size_t p_bytes = mode == 1 ? (size_t)n * std::max(W, 1) * 4 : (mode == 2 ? (size_t)hit_cap * sizeof(uint2) : 0);
  const size_t out_bytes = (p_off + p_bytes + 15) / 16 * 16;
  SmallGraph *g = nullptr;
  for (auto &x : s->small_graphs)
    if (x.n == n && x.mode == mode && x.hit_cap == hit_cap) g = &x;
  if (!g) {
    if (s->small_graphs.size() >= 8) {
      CU(cudaStreamSynchronize(st));
      SmallGraph &old = s->small_graphs.front();
      if (old.exec) cudaGraphExecDestroy(old.exec);
      if (old.graph) cudaGraphDestroy(old.graph);
      s->small_graphs.erase(s->small_graphs.begin());
    }
    const size_t need = q_bytes + out_bytes + 64;
    if (need > s->h_small_cap || out_bytes > s->d_small_cap) {
      CU(cudaStreamSynchronize(st));
      drop_small_graphs(s);  // they point into the old staging blocks
      if (need > s->h_small_cap) {
        if (s->h_small) cudaFreeHost(s->h_small);
        s->h_small = nullptr;
        s->h_small_cap = 0;
        CU(cudaHostAlloc((void **)&s->h_small, need * 2, cudaHostAllocDefault));
        s->h_small_cap = need * 2;
      }
      CU(ensure(&s->d_small, &s->d_small_cap, out_bytes * 2));
    }
    CU(ensure(&s->d_q[0], &s->d_q_cap[0], (size_t)chunk * std::max(D, 1)));
    if (mode == 2) CU(ensure(&s->d_pairbits[0], &s->d_pairbits_cap[0], (size_t)chunk * std::max(W, 1)));  // dense scratch of the compaction
    s->small_graphs.push_back(SmallGraph{});
    g = &s->small_graphs.back();
    g->n = n; g->mode = mode; g->hit_cap = hit_cap;
    g->d_q = s->d_q[0]; g->d_v = s->d_small; g->d_p = mode == 2 ? s->d_pairbits[0] : nullptr; g->h_base = s->h_small; g->chunk_at = s->chunk;
    g->img_version = s->img_version;
    unsigned char *hb = s->h_small, *db = s->d_small;
    CU(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    int rc = MPB_OK;
    cudaError_t ce = cudaSuccess;
    do {
      if (D > 0 && (ce = cudaMemcpyAsync(s->d_q[0], hb, (size_t)n * D * sizeof(float), cudaMemcpyHostToDevice, st)) != cudaSuccess) break;
      PassOut po;
      po.dv = reinterpret_cast<uint32_t *>(db + v_off);
      po.dp = mode == 1 ? reinterpret_cast<uint32_t *>(db + p_off) : (mode == 2 ? s->d_pairbits[0] : nullptr);
      po.reset_call = true;
      po.n_hits = reinterpret_cast<unsigned *>(db);
      if (mode == 2) { po.hits = reinterpret_cast<uint2 *>(db + p_off); po.hit_cap = hit_cap; }
      size_t sync_ev = 0;
      long long launches = 0;
      rc = launch_pass(s, nullptr, s->d_q[0], n, po, st, PassEvents{}, &launches, nullptr, &sync_ev);
      g->launches = launches;
      if (rc != MPB_OK) break;
      if ((ce = cudaMemcpyAsync(hb + q_bytes, db, out_bytes, cudaMemcpyDeviceToHost, st)) != cudaSuccess) break;
      ce = cudaMemcpyAsync(hb + q_bytes + out_bytes, s->wq.totals, 64, cudaMemcpyDeviceToHost, st);
    } while (0);
    cudaError_t ee = cudaStreamEndCapture(st, &g->graph);
    if (rc == MPB_OK && ce == cudaSuccess && ee == cudaSuccess) ee = cudaGraphInstantiate(&g->exec, g->graph, 0);
    if (rc != MPB_OK || ce != cudaSuccess || ee != cudaSuccess) {
      if (g->graph) cudaGraphDestroy(g->graph);
      s->small_graphs.pop_back();
      cudaGetLastError();
      if (rc != MPB_OK) return rc;
      return fail(MPB_ERR_CUDA, std::string("graph capture of the small-batch call: ") + cudaGetErrorString(ce != cudaSuccess ? ce : ee));
    }
  }
  unsigned char *hb = s->h_small;
  const unsigned char *ho = hb + q_bytes;
  const unsigned *h_st = reinterpret_cast<const unsigned *>(ho + out_bytes);
  if (D > 0) memcpy(hb, q, (size_t)n * D * sizeof(float));
  CU(cudaGraphLaunch(g->exec, st));
  CU(cudaStreamSynchronize(st));
  *overflow = h_st[2] != 0;
  if (mode == 2 && !*overflow) {
    const unsigned nh = *reinterpret_cast<const unsigned *>(ho);
    if (nh > hit_cap && user_cap > hit_cap) *overflow = true;  // more hits than the graph copies: chunked path
    else {
      *co.n_hits = nh;
      memcpy(co.hits, ho + p_off, (size_t)std::min(nh, hit_cap) * sizeof(uint2));
    }
  }
  if (!*overflow) {
    memcpy(co.verdict, ho + v_off, vwords * 4);
    if (mode == 1) memcpy(co.pair_bits, ho + p_off, (size_t)n * W * 4);
  }
  if (stats) {
    fill_stats(s, stats, n, reinterpret_cast<const unsigned long long *>(h_st + 8), g->launches);
    stats->host_ms_enqueue = stats->host_ms_total = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
  }
  return MPB_OK;
}

static bool small_graphs_valid(mpb_scene *s) {
  for (auto &g : s->small_graphs)
    if (g.d_q != s->d_q[0] || g.d_v != s->d_small || (g.mode == 2 && g.d_p != s->d_pairbits[0]) || g.h_base != s->h_small ||
        g.chunk_at != s->chunk || g.img_version != s->img_version)
      return false;
  return true;
}

static int collide_common(mpb_scene *s, const float *q, int64_t n, const CollideOut &co, mpb_stats *stats, int space, void *stream,
                          bool may_async) {
  if (!s) return fail(MPB_ERR_INVALID, "null scene");
  if (n < 0 || (n > 0 && (!co.verdict || (!q && s->img.state_dim > 0)))) return fail(MPB_ERR_INVALID, "null batch pointer");
  if (space != MPB_HOST && space != MPB_DEVICE) return fail(MPB_ERR_INVALID, "space must be MPB_HOST or MPB_DEVICE");
  if (co.hits && (!co.n_hits || co.hit_cap < 0)) return fail(MPB_ERR_INVALID, "hit list without a counter / negative capacity");
  if (stats) memset(stats, 0, sizeof(*stats));
  if (n == 0) {
    if (co.hits && space == MPB_HOST) *co.n_hits = 0;
    return MPB_OK;
  }
  CU(cudaSetDevice(s->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  if (s->host_chunk == 0) {  // host buffers: H2D sub-range size (FK + cull of one sub-range overlaps the next copy)
    const char *e = getenv("MPB_HOST_CHUNK");
    s->host_chunk = e ? std::max<long long>(64, atoll(e) / 64 * 64) : (1ll << 17);
  }
  int rc = harvest(s, false);
  if (rc != MPB_OK) return rc;
  // a pass size at most as large as one that already went through the blocking protocol without overflow
  const long long want = std::min<long long>(std::min<long long>(max_chunk_for(s), s->cap_chunk), (n + 31) / 32 * 32);
  if (space == MPB_DEVICE && may_async && s->calibrated && s->good_chunk >= want) return collide_enqueue(s, q, n, co, st, stats);
  // every blocking protocol below owns the work queues: calls still in flight finish first
  rc = harvest(s, true);
  if (rc != MPB_OK) return rc;
  static const bool small_on = !(getenv("MPB_NO_SMALL_GRAPH") && atoi(getenv("MPB_NO_SMALL_GRAPH")));
  if (space == MPB_HOST && n <= kSmallBatch && !stream && small_on) {
    rc = prepare_work(s, (n + 31) / 32 * 32);
    if (rc != MPB_OK) return rc;
    if (!small_graphs_valid(s)) { CU(cudaStreamSynchronize(s->stream)); drop_small_graphs(s); }
    bool overflow = false;
    rc = collide_small(s, q, n, co, &overflow, stats);
    if (rc != MPB_OK) return rc;
    if (!overflow) return MPB_OK;
  }
  long long chunk = want;
  for (int attempt = 0; attempt < 12; attempt++) {
    bool overflow = false;
    rc = collide_attempt(s, q, n, co, space, st, chunk, &overflow, stats);
    if (rc != MPB_OK) return rc;
    if (!overflow) return MPB_OK;
    chunk = std::max<long long>(32, (chunk / 4 + 31) / 32 * 32);  // queue overflowed: smaller passes, exact redo
  }
  return fail(MPB_ERR_CAPACITY, "work queue overflow persists at the minimum chunk size");
}

extern "C" int mpb_collide_batch(mpb_scene *s, const float *q, int64_t n, uint32_t *verdict_bits, uint32_t *pair_bits,
                                 mpb_stats *stats, int space, void *stream) {
  CollideOut co;
  co.verdict = verdict_bits;
  co.pair_bits = pair_bits;
  return collide_common(s, q, n, co, stats, space, stream, true);
}

extern "C" int mpb_collide_batch_hits(mpb_scene *s, const float *q, int64_t n, uint32_t *verdict_bits, uint32_t *hits, int64_t hit_cap,
                                      uint32_t *n_hits, mpb_stats *stats, int space, void *stream) {
  if (!hits || !n_hits) return fail(MPB_ERR_INVALID, "null hit list / counter");
  CollideOut co;
  co.verdict = verdict_bits;
  co.hits = hits;
  co.hit_cap = hit_cap;
  co.n_hits = n_hits;
  return collide_common(s, q, n, co, stats, space, stream, true);
}

extern "C" int mpb_fk_batch(mpb_scene *s, const float *q, int64_t n, float *poses, int space, void *stream) {
  if (!s) return fail(MPB_ERR_INVALID, "null scene");
  if (n < 0 || (n > 0 && (!poses || (!q && s->img.state_dim > 0)))) return fail(MPB_ERR_INVALID, "null batch pointer");
  if (space != MPB_HOST && space != MPB_DEVICE) return fail(MPB_ERR_INVALID, "space must be MPB_HOST or MPB_DEVICE");
  if (n == 0 || s->img.n_links == 0) return MPB_OK;
  CU(cudaSetDevice(s->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  const int D = s->img.state_dim, L = s->img.n_links;
  const long long chunk = std::min<long long>(kMaxChunk, n);
  if (space == MPB_HOST) {
    CU(ensure(&s->d_q[0], &s->d_q_cap[0], (size_t)chunk * std::max(D, 1)));
    CU(ensure(&s->d_poses, &s->d_poses_cap, (size_t)chunk * L * 7));
  }
  for (long long off = 0; off < n; off += chunk) {
    const long long m = std::min(chunk, n - off);
    const float *dq = q + off * D;
    float *dp = poses + (size_t)off * L * 7;
    if (space == MPB_HOST) {
      if (D > 0) CU(cudaMemcpyAsync(s->d_q[0], dq, (size_t)m * D * sizeof(float), cudaMemcpyHostToDevice, st));
      dq = s->d_q[0];
      dp = s->d_poses;
    }
    const int blocks = (int)std::min<long long>((m + 127) / 128, (long long)s->num_sms * 16);
    fk_pose_kernel<<<blocks, 128, 0, st>>>(s->img, dq, m, dp);
    CU(cudaGetLastError());
    if (space == MPB_HOST)
      CU(cudaMemcpyAsync(poses + (size_t)off * L * 7, dp, (size_t)m * L * 7 * sizeof(float), cudaMemcpyDeviceToHost, st));
  }
  if (space == MPB_HOST) CU(cudaStreamSynchronize(st));
  return MPB_OK;
}

// ------------------------------------------------------------------------------------------ diagnostics: support cell table
extern "C" int mpb_debug_cell_table(const double *verts, int32_t n_verts, uint8_t *out, int64_t cap, int64_t *n_bytes) {
  if (!verts || n_verts < 1 || n_verts > 255 || !n_bytes) return fail(MPB_ERR_INVALID, "mpb_debug_cell_table: 1..255 vertices");
  std::vector<unsigned char> t;
  build_cell_table_uncached(verts, n_verts, t);
  *n_bytes = (int64_t)t.size();
  if (out && cap >= (int64_t)t.size()) memcpy(out, t.data(), t.size());
  return MPB_OK;
}

// ------------------------------------------------------------------------------------------ FK without allocation
// mpb_fk_batch_noalloc: the FK-only call for callers that may not allocate (code under CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// fk_pose_kernel keeps the joint-frame table T[] of a configuration in per-thread local memory (the driver's own scratch
// allocation); this variant keeps it in a fixed array of the library image, word-major [joint * 12 + word][configuration]
// (coalesced), so the call needs nothing but the caller's input and output buffers.  Every GPU that loads the library
// holds MPB_NOALLOC_SCRATCH_BYTES until exit.  All calls on a device share the array: calls that overlap on the GPU
// (two streams) race.  Nothing in it has to start at zero -- a thread reads only words it wrote in the same launch.
static __device__ __align__(256) float g_fk_scratch[MPB_NOALLOC_SCRATCH_BYTES / sizeof(float)];

__global__ void __launch_bounds__(128) fk_pose_scratch_kernel(SceneImg img, const float *__restrict__ q, long long n, float *__restrict__ poses,
                                                              float *scratch, size_t ld) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float *T = scratch + i;  // word w of joint frame j: T[(j * 12 + w) * ld]
  auto put = [&](int j, const Xf &x) {
#pragma unroll
    for (int k = 0; k < 9; k++) T[(size_t)(j * 12 + k) * ld] = x.r[k];
#pragma unroll
    for (int k = 0; k < 3; k++) T[(size_t)(j * 12 + 9 + k) * ld] = x.p[k];
  };
  auto get = [&](int j) {
    Xf x;
#pragma unroll
    for (int k = 0; k < 9; k++) x.r[k] = T[(size_t)(j * 12 + k) * ld];
#pragma unroll
    for (int k = 0; k < 3; k++) x.p[k] = T[(size_t)(j * 12 + 9 + k) * ld];
    return x;
  };
  const float *qrow = q + i * img.state_dim;
  put(0, xf_identity());
  for (int j = 0; j < img.n_joints; j++) {  // the arithmetic of fk_joints, term by term
    const DevJoint &J = img.joints[j];
    const float qj = J.slot >= 0 ? qrow[J.slot] : J.qconst;
    const Xf li = xf_mul(load_xf12(J.P), joint_motion(J, qj));
    put(j + 1, xf_mul(get(J.parent), li));
  }
  for (int l = 0; l < img.n_links; l++) {
    const Xf w = xf_mul(get(img.linkJoint[l]), load_xf12(img.linkL + 12 * l));
    float *o = poses + ((size_t)i * img.n_links + l) * 7;
    float qt[4];
    mat_to_quat(w.r, qt);
    o[0] = w.p[0]; o[1] = w.p[1]; o[2] = w.p[2]; o[3] = qt[0]; o[4] = qt[1]; o[5] = qt[2]; o[6] = qt[3];
  }
}

// Device address of g_fk_scratch on `device` (which must be current): in host code the array's name is only a handle for
// the runtime, not a device pointer.  Resolved once per device.
static int fk_scratch_address(int device, float **ptr) {
  static std::mutex mu;
  static float *addr[64] = {};
  if (device < 0 || device >= 64) return fail(MPB_ERR_INVALID, "device index out of range");
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[device]) {
    void *p = nullptr;
    CU(cudaGetSymbolAddress(&p, g_fk_scratch));
    addr[device] = static_cast<float *>(p);
  }
  *ptr = addr[device];
  return MPB_OK;
}

extern "C" int mpb_fk_batch_noalloc(mpb_scene *s, const float *q_dev, int64_t n, float *poses_dev, void *stream) {
  if (!s) return fail(MPB_ERR_INVALID, "null scene");
  if (n < 0 || (n > 0 && (!poses_dev || (!q_dev && s->img.state_dim > 0)))) return fail(MPB_ERR_INVALID, "null batch pointer");
  if (n == 0 || s->img.n_links == 0) return MPB_OK;
  const size_t ld = ((size_t)n + 31) / 32 * 32;
  const size_t words = (size_t)(s->img.n_joints + 1) * 12;
  const size_t max_n = MPB_NOALLOC_SCRATCH_BYTES / sizeof(float) / words / 32 * 32;
  if (ld > max_n)
    return fail(MPB_ERR_CAPACITY, "mpb_fk_batch_noalloc: n = " + std::to_string(n) + " exceeds the fixed scratch array (" +
                                      std::to_string(MPB_NOALLOC_SCRATCH_BYTES) + " bytes = at most " + std::to_string(max_n) +
                                      " configurations of this scene per call); split the batch or use mpb_fk_batch");
  CU(cudaSetDevice(s->device));
  float *scratch = nullptr;
  if (int rc = fk_scratch_address(s->device, &scratch)) return rc;
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  fk_pose_scratch_kernel<<<(int)((n + 127) / 128), 128, 0, st>>>(s->img, q_dev, n, poses_dev, scratch, ld);
  CU(cudaGetLastError());
  return MPB_OK;
}

// ------------------------------------------------------------------------------------------ FP32 peak probe
__global__ void __launch_bounds__(256) fma_probe_kernel(float *out, int iters) {
  float a[8];
#pragma unroll
  for (int k = 0; k < 8; k++) a[k] = 1.0f + 1e-3f * (threadIdx.x + k);
  const float m = 1.000001f, c = 1e-7f;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int r = 0; r < 8; r++)
#pragma unroll
      for (int k = 0; k < 8; k++) a[k] = fmaf(a[k], m, c);
  }
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 8; k++) s += a[k];
  if (s == 12345.678f) out[0] = s;  // keep the chain alive
}

extern "C" int mpb_fp32_peak(int device, float *tflops) {
  if (!tflops) return fail(MPB_ERR_INVALID, "null argument");
  int ndev = mpb_device_count();
  if (device < 0 || device >= ndev) return fail(MPB_ERR_CUDA, "no such CUDA device");
  CU(cudaSetDevice(device));
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  float *d_out = nullptr;
  CU(cudaMalloc(&d_out, 16));
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0));
  CU(cudaEventCreate(&e1));
  const int blocks = prop.multiProcessorCount * 8, iters = 4096;
  float best = 0.f;
  for (int rep = 0; rep < 4; rep++) {
    CU(cudaEventRecord(e0));
    fma_probe_kernel<<<blocks, 256>>>(d_out, iters);
    CU(cudaEventRecord(e1));
    CU(cudaEventSynchronize(e1));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, e0, e1));
    const double flop = (double)blocks * 256.0 * iters * 64.0 * 2.0;
    if (rep > 0) best = std::max(best, (float)(flop / (ms * 1e-3) / 1e12));
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d_out);
  *tflops = best;
  return MPB_OK;
}

// ------------------------------------------------------------------------------------------ planning layer
struct DevSpace {
  SpaceDesc sp{};
  float *buf = nullptr;  // lo | hi | seg | so2 bytes
};

struct DevBufs;
static int upload_space(const mpb_scene *s, const mpb_space_desc *space, DevSpace *out, float *buf) {
  if (!space || !space->lo || !space->hi || !space->seg) return fail(MPB_ERR_INVALID, "null state-space description");
  const int D = space->dim;
  if (D != s->img.state_dim) return fail(MPB_ERR_INVALID, "state-space dimension does not match the scene's state_dim");
  if (D <= 0) return fail(MPB_ERR_INVALID, "empty state space");
  for (int j = 0; j < D; j++)
    if (!(space->seg[j] > 0.f) || !(space->hi[j] >= space->lo[j])) return fail(MPB_ERR_INVALID, "state-space bounds / segment lengths must be positive");
  std::vector<float> h((size_t)4 * D, 0.f);
  memcpy(h.data(), space->lo, D * sizeof(float));
  memcpy(h.data() + D, space->hi, D * sizeof(float));
  memcpy(h.data() + 2 * D, space->seg, D * sizeof(float));
  unsigned char *so2 = reinterpret_cast<unsigned char *>(h.data() + 3 * D);
  for (int j = 0; j < D; j++) so2[j] = space->so2 ? space->so2[j] : 0;
  out->buf = buf;  // caller-owned, >= 4 * dim floats
  CU(cudaMemcpy(out->buf, h.data(), h.size() * sizeof(float), cudaMemcpyHostToDevice));
  out->sp.dim = D;
  out->sp.lo = out->buf;
  out->sp.hi = out->buf + D;
  out->sp.seg = out->buf + 2 * D;
  out->sp.so2 = reinterpret_cast<const unsigned char *>(out->buf + 3 * D);
  return MPB_OK;
}

struct DevBufs {  // scratch blocks of one call: taken from the scene's pool, handed back on scope exit
  mpb_scene *s;
  std::vector<std::pair<void *, size_t>> held;
  explicit DevBufs(mpb_scene *scene) : s(scene) {}
  template <typename T>
  cudaError_t alloc(T **p, size_t n) {
    const size_t bytes = (std::max<size_t>(n, 1) * sizeof(T) + 255) / 256 * 256;
    int best = -1;  // smallest pooled block that fits without wasting more than half of itself
    for (size_t i = 0; i < s->pool.size(); i++)
      if (s->pool[i].second >= bytes && s->pool[i].second <= 2 * bytes + 4096 &&
          (best < 0 || s->pool[i].second < s->pool[(size_t)best].second))
        best = (int)i;
    if (best >= 0) {
      *p = (T *)s->pool[(size_t)best].first;
      held.push_back(s->pool[(size_t)best]);
      s->pool.erase(s->pool.begin() + best);
      return cudaSuccess;
    }
    cudaError_t e = cudaMalloc((void **)p, bytes);
    if (e != cudaSuccess) {  // pooled blocks of other shapes may be what exhausts the device: drop them and retry
      cudaGetLastError();
      for (auto &b : s->pool) cudaFree(b.first);
      s->pool.clear();
      e = cudaMalloc((void **)p, bytes);
    }
    if (e == cudaSuccess) held.push_back(std::make_pair((void *)*p, bytes));
    return e;
  }
  ~DevBufs() {  // every entry point synchronises its stream before returning, so the blocks are idle here
    for (auto &b : held) s->pool.push_back(b);
    while (s->pool.size() > 48) {
      cudaFree(s->pool.front().first);
      s->pool.erase(s->pool.begin());
    }
  }
};

extern "C" int mpb_check_motion_batch(mpb_scene *s, const mpb_space_desc *space, const float *a, const float *b, int64_t n_edges,
                                      uint32_t *valid_bits, mpb_stats *stats, int space_flag, void *stream) {
  if (!s) return fail(MPB_ERR_INVALID, "null scene");
  if (n_edges < 0 || (n_edges > 0 && (!a || !b || !valid_bits))) return fail(MPB_ERR_INVALID, "null batch pointer");
  if (space_flag != MPB_HOST && space_flag != MPB_DEVICE) return fail(MPB_ERR_INVALID, "space must be MPB_HOST or MPB_DEVICE");
  if (stats) memset(stats, 0, sizeof(*stats));
  if (n_edges == 0) return MPB_OK;
  if (n_edges > (1ll << 30)) return fail(MPB_ERR_CAPACITY, "too many edges in one call");
  CU(cudaSetDevice(s->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  DevBufs bufs(s);
  DevSpace dsp;
  float *d_space = nullptr;
  CU(bufs.alloc(&d_space, (size_t)4 * std::max(s->img.state_dim, 1)));
  int rc = upload_space(s, space, &dsp, d_space);
  if (rc != MPB_OK) return rc;
  const int D = s->img.state_dim;
  const bool host = space_flag == MPB_HOST;
  const float *da = a, *db = b;
  uint32_t *dvalid = valid_bits;
  const size_t vwords = (size_t)(n_edges + 31) / 32;
  if (host) {
    float *ta = nullptr, *tb = nullptr;
    CU(bufs.alloc(&ta, (size_t)n_edges * D));
    CU(bufs.alloc(&tb, (size_t)n_edges * D));
    CU(bufs.alloc(&dvalid, vwords));
    CU(cudaMemcpyAsync(ta, a, (size_t)n_edges * D * sizeof(float), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(tb, b, (size_t)n_edges * D * sizeof(float), cudaMemcpyHostToDevice, st));
    da = ta;
    db = tb;
  }
  int *d_nd = nullptr;
  unsigned *d_first = nullptr, *d_total = nullptr;
  CU(bufs.alloc(&d_nd, (size_t)n_edges));
  CU(bufs.alloc(&d_first, (size_t)n_edges));
  CU(bufs.alloc(&d_total, 2));
  CU(cudaMemsetAsync(d_total, 0, 2 * sizeof(unsigned), st));
  const int eblocks = (int)((n_edges + 255) / 256);
  motion_count_kernel<<<eblocks, 256, 0, st>>>(dsp.sp, da, db, n_edges, d_nd, d_first, d_total);
  unsigned total = 0;
  CU(cudaMemcpyAsync(&total, d_total, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  if (total == 0 || total > (1u << 30)) return fail(MPB_ERR_CAPACITY, "motion batch expands to too many states");
  float *d_states = nullptr;
  int *d_owner = nullptr;
  uint32_t *d_verdict = nullptr;
  CU(bufs.alloc(&d_states, (size_t)total * D));
  CU(bufs.alloc(&d_owner, (size_t)total));
  CU(bufs.alloc(&d_verdict, (size_t)(total + 31) / 32));
  motion_expand_kernel<<<eblocks, 256, 0, st>>>(dsp.sp, da, db, n_edges, d_nd, d_first, 0, d_states, d_owner);
  fill_valid_bits_kernel<<<(int)((vwords + 255) / 256), 256, 0, st>>>(dvalid, n_edges);
  CU(cudaGetLastError());
  mpb_stats inner;
  {
    CollideOut co;
    co.verdict = d_verdict;
    rc = collide_common(s, d_states, (int64_t)total, co, &inner, MPB_DEVICE, (void *)st, false);  // blocking: overflow is redone
  }
  if (rc != MPB_OK) return rc;
  motion_reduce_kernel<<<(int)((total + 255u) / 256u), 256, 0, st>>>(d_verdict, d_owner, total, dvalid);
  CU(cudaGetLastError());
  if (host) CU(cudaMemcpyAsync(valid_bits, dvalid, vwords * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));  // the scratch buffers are freed on return
  if (stats) {
    *stats = inner;
    stats->kernel_launches += 4;
  }
  return MPB_OK;
}

extern "C" int mpb_rrtconnect_batch(mpb_scene *s, const mpb_space_desc *space, const float *starts, const float *goals,
                                    const int32_t *n_goals, int32_t max_goals, int64_t n_queries, const mpb_rrt_params *params,
                                    int32_t *status, int32_t *path_len, float *paths, mpb_rrt_stats *stats) {
  if (!s) return fail(MPB_ERR_INVALID, "null scene");
  if (!params) return fail(MPB_ERR_INVALID, "null parameters");
  if (n_queries < 0 || max_goals < 1) return fail(MPB_ERR_INVALID, "bad query count / max_goals");
  if (n_queries > 0 && (!starts || !goals || !n_goals || !status || !path_len || !paths)) return fail(MPB_ERR_INVALID, "null batch pointer");
  if (stats) memset(stats, 0, sizeof(*stats));
  if (n_queries == 0) return MPB_OK;
  if (n_queries > (1 << 20)) return fail(MPB_ERR_CAPACITY, "too many queries in one call");
  const auto t_begin = std::chrono::steady_clock::now();
  CU(cudaSetDevice(s->device));
  if (int hrc = harvest(s, true)) return hrc;  // device-pointer collide calls in flight share the work queues
  cudaStream_t st = s->stream;
  DevBufs bufs(s);
  DevSpace dsp;
  float *d_space = nullptr;
  CU(bufs.alloc(&d_space, (size_t)4 * std::max(s->img.state_dim, 1)));
  int rc = upload_space(s, space, &dsp, d_space);
  if (rc != MPB_OK) return rc;
  const int D = s->img.state_dim, Q = (int)n_queries;
  const int W = params->n_slots > 0 ? std::min(params->n_slots, Q) : Q;  // concurrent searches
  const int cap = std::max(params->tree_capacity, 2), max_path = std::max(params->max_path, 2);
  const int max_rounds = std::max(params->max_rounds, 1), check_every = params->check_every > 0 ? params->check_every : 16;
  float range = params->range;
  if (!(range > 1e-6f)) {  // OMPL selfConfig: 20 % of the maximum extent (sum over the weighted subspaces)
    double ext = 0;
    for (int j = 0; j < D; j++) ext += (space->so2 && space->so2[j]) ? 3.14159265358979323846 : (double)space->hi[j] - space->lo[j];
    range = (float)(0.2 * ext);
  }
  int smax = 1;
  for (int j = 0; j < D; j++) smax = std::max(smax, (int)std::ceil((double)range / space->seg[j]) + 1);
  if (smax > 4096) return fail(MPB_ERR_CAPACITY, "range / validity resolution gives too many samples per edge");
  const int S = params->connect_steps > 0 ? std::min(std::max(params->connect_steps, 2), 32) : 8;
  const long long n_states = (long long)W * S * smax;
  rc = prepare_work(s, (n_states + 31) / 32 * 32);
  if (rc != MPB_OK) return rc;

  RrtState r{};
  r.n_queries = Q; r.n_slots = W; r.cap = cap; r.dim = D; r.smax = smax; r.max_goals = max_goals; r.max_path = max_path;
  r.S = S;
  r.max_rounds_q = params->max_rounds_per_query;
  r.seed = params->seed;
  float *d_start = nullptr, *d_goals = nullptr, *d_states = nullptr;
  int *d_ngoals = nullptr, *d_ints = nullptr, *d_qints = nullptr;
  uint32_t *d_verdict = nullptr;
  CU(bufs.alloc(&r.nodes, (size_t)W * 2 * cap * D));
  CU(bufs.alloc(&r.parent, (size_t)W * 2 * cap));
  CU(bufs.alloc(&d_start, (size_t)Q * D));
  CU(bufs.alloc(&d_goals, (size_t)Q * max_goals * D));
  CU(bufs.alloc(&d_ngoals, (size_t)Q));
  CU(bufs.alloc(&d_ints, (size_t)W * 15 + 4));   // per slot: size[2] slot_query rounds_used goals_used phase flag tsel nearest nearest2 reach n_steps added iters | pending[2]
  CU(bufs.alloc(&d_qints, (size_t)Q * 3));       // per query: status, path_len, iters
  CU(bufs.alloc(&r.target, (size_t)W * D));
  CU(bufs.alloc(&r.edge_a, (size_t)W * S * D));
  CU(bufs.alloc(&r.edge_b, (size_t)W * S * D));
  CU(bufs.alloc(&d_states, (size_t)n_states * D));
  CU(bufs.alloc(&d_verdict, (size_t)(n_states + 31) / 32 + 1));
  CU(bufs.alloc(&r.q_paths, (size_t)Q * max_path * D));
  r.size = d_ints;
  r.slot_query = d_ints + 2 * W; r.rounds_used = d_ints + 3 * W; r.goals_used = d_ints + 4 * W; r.phase = d_ints + 5 * W;
  r.flag_start = d_ints + 6 * W; r.tsel = d_ints + 7 * W; r.nearest = d_ints + 8 * W; r.nearest2 = d_ints + 9 * W;
  r.reach = d_ints + 10 * W; r.n_steps = d_ints + 11 * W; r.added = d_ints + 12 * W; r.iters = d_ints + 13 * W;
  r.pending = reinterpret_cast<unsigned *>(d_ints + 15 * W);
  r.q_status = d_qints; r.q_pathlen = d_qints + Q; r.q_iters = d_qints + 2 * Q;
  r.start = d_start; r.goals = d_goals; r.n_goals = d_ngoals;
  {
    std::vector<int> h((size_t)W * 15 + 4, 0);
    for (int w = 0; w < W; w++) h[2 * (size_t)W + w] = -1;  // every slot idle: the first round hands out the first W queries
    CU(cudaMemcpyAsync(d_ints, h.data(), h.size() * sizeof(int), cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(d_qints, 0, (size_t)Q * 3 * sizeof(int), st));  // status MPB_RRT_ACTIVE, no path
    CU(cudaMemcpyAsync(d_start, starts, (size_t)Q * D * sizeof(float), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d_goals, goals, (size_t)Q * max_goals * D * sizeof(float), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(d_ngoals, n_goals, (size_t)Q * sizeof(int), cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(d_verdict, 0, ((size_t)(n_states + 31) / 32 + 1) * sizeof(uint32_t), st));
    CU(cudaMemsetAsync(r.edge_a, 0xff, (size_t)W * S * D * sizeof(float), st));  // NaN: idle slots cost no narrowphase work
    CU(cudaMemsetAsync(r.edge_b, 0xff, (size_t)W * S * D * sizeof(float), st));
    CU(cudaMemsetAsync(s->wq.totals, 0, 8 * sizeof(unsigned), st));
    CU(cudaMemsetAsync(s->wq.counters, 0, 4 * sizeof(unsigned long long), st));
    CU(cudaStreamSynchronize(st));  // h goes out of scope
  }
  long long launches_per_round = 0;
  auto enqueue_round = [&](long long *launches) -> int {
    size_t sync_ev = 0;
    rrt_advance_kernel<<<(W * 32 + 255) / 256, 256, 0, st>>>(r, dsp.sp, range, d_verdict);
    motion_expand_kernel<<<(W * S + 255) / 256, 256, 0, st>>>(dsp.sp, r.edge_a, r.edge_b, (long long)W * S, nullptr, nullptr, smax, d_states, nullptr);
    *launches += 2;
    PassOut po;
    po.dv = d_verdict;
    return launch_pass(s, nullptr, d_states, n_states, po, st, PassEvents{}, launches, nullptr, &sync_ev);
  };
  // the round has fixed launch shapes: capture it once, replay it
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  const bool use_graph = getenv("MPB_NO_GRAPH") == nullptr;
  if (use_graph) {
    CU(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    rc = enqueue_round(&launches_per_round);
    cudaError_t ce = cudaStreamEndCapture(st, &graph);
    if (rc != MPB_OK || ce != cudaSuccess) {
      if (graph) cudaGraphDestroy(graph);
      return rc != MPB_OK ? rc : fail(MPB_ERR_CUDA, std::string("graph capture: ") + cudaGetErrorString(ce));
    }
    CU(cudaGraphInstantiate(&exec, graph, 0));
  }
  cudaEvent_t e0 = get_event(s, 0), e1 = get_event(s, 1);
  CU(cudaEventRecord(e0, st));
  long long rounds = 0, launches = 0;
  unsigned pending[2] = {0, 0};
  unsigned totals[8] = {0};
  while (rounds < max_rounds && pending[1] < (unsigned)Q) {
    const int k = (int)std::min<long long>(check_every, max_rounds - rounds);
    for (int i = 0; i < k; i++) {
      if (use_graph) { CU(cudaGraphLaunch(exec, st)); launches += launches_per_round; }
      else { rc = enqueue_round(&launches); if (rc != MPB_OK) return rc; }
    }
    rounds += k;
    CU(cudaMemcpyAsync(pending, r.pending, sizeof(pending), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(totals, s->wq.totals, sizeof(totals), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (totals[2]) { rc = fail(MPB_ERR_CAPACITY, "work queue overflow inside the planning round"); break; }
    if (params->time_limit_s > 0.f &&
        std::chrono::duration<float>(std::chrono::steady_clock::now() - t_begin).count() > params->time_limit_s) break;
  }
  if (exec) cudaGraphExecDestroy(exec);
  if (graph) cudaGraphDestroy(graph);
  if (rc != MPB_OK) return rc;
  // one more advance so that the verdict of the last round is consumed (it may complete a query)
  rrt_advance_kernel<<<(W * 32 + 255) / 256, 256, 0, st>>>(r, dsp.sp, range, d_verdict);
  rrt_flush_kernel<<<(W + 127) / 128, 128, 0, st>>>(r);
  launches += 2;
  CU(cudaEventRecord(e1, st));
  CU(cudaGetLastError());
  std::vector<int> h_iters(Q);
  CU(cudaMemcpyAsync(status, r.q_status, (size_t)Q * sizeof(int), cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(path_len, r.q_pathlen, (size_t)Q * sizeof(int), cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(h_iters.data(), r.q_iters, (size_t)Q * sizeof(int), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  {  // waypoints: only the used rows cross PCIe (packed on the device, scattered into the caller's matrix here)
    std::vector<int> off((size_t)Q + 1, 0);
    for (int q = 0; q < Q; q++) off[(size_t)q + 1] = off[(size_t)q] + path_len[q];
    const int total = off[(size_t)Q];
    if (total > 0) {
      int *d_off = nullptr;
      float *d_packed = nullptr;
      CU(bufs.alloc(&d_off, (size_t)Q + 1));
      CU(bufs.alloc(&d_packed, (size_t)total * D));
      CU(cudaMemcpyAsync(d_off, off.data(), ((size_t)Q + 1) * sizeof(int), cudaMemcpyHostToDevice, st));
      pack_paths_kernel<<<Q, 128, 0, st>>>(r.q_paths, d_off, D, max_path, d_packed);
      launches += 1;
      std::vector<float> h_packed((size_t)total * D);
      CU(cudaMemcpyAsync(h_packed.data(), d_packed, h_packed.size() * sizeof(float), cudaMemcpyDeviceToHost, st));
      CU(cudaStreamSynchronize(st));
      for (int q = 0; q < Q; q++)
        if (path_len[q] > 0)
          memcpy(paths + (size_t)q * max_path * D, h_packed.data() + (size_t)off[(size_t)q] * D, (size_t)path_len[q] * D * sizeof(float));
    }
  }
  if (stats) {
    stats->rounds = rounds;
    stats->states_checked = rounds * n_states;
    stats->kernel_launches = launches;
    for (int q = 0; q < Q; q++) {
      stats->grow_calls += h_iters[q];
      stats->n_solved += status[q] == MPB_RRT_SOLVED;
    }
    cudaEventElapsedTime(&stats->gpu_ms, e0, e1);
    stats->host_ms_total = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
  }
  return MPB_OK;
}

// ------------------------------------------------------------------------------------------ distance
#include "mpb_distance.cuh"

extern "C" int mpb_distance_batch(mpb_scene *s, const float *q, int64_t n, float *min_dist, int32_t *argmin, mpb_stats *stats,
                                  int space, void *stream) {
  if (!s) return fail(MPB_ERR_INVALID, "null scene");
  if (n < 0 || (n > 0 && (!min_dist || (!q && s->img.state_dim > 0)))) return fail(MPB_ERR_INVALID, "null batch pointer");
  if (space != MPB_HOST && space != MPB_DEVICE) return fail(MPB_ERR_INVALID, "space must be MPB_HOST or MPB_DEVICE");
  if (stats) memset(stats, 0, sizeof(*stats));
  if (n == 0) return MPB_OK;
  CU(cudaSetDevice(s->device));
  if (int hrc = harvest(s, true)) return hrc;
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  const auto t_begin = std::chrono::steady_clock::now();
  const int D = s->img.state_dim;
  const bool host = space == MPB_HOST;
  const size_t vsm = s->smem_verts ? (size_t)s->img.pool_bytes : 0;
  if (s->smem_verts) CU(raise_dyn_smem((const void *)distance_kernel<true>, s->device, 1, std::max<size_t>(vsm, 16)));
  int per_sm = 0;
  if (s->smem_verts) CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, distance_kernel<true>, 256, vsm));
  else CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, distance_kernel<false>, 256, 0));
  const int dblocks = s->num_sms * std::max(per_sm, 1);
  DevBufs bufs(s);
  long long chunk = std::min<long long>(kMaxChunk, (n + 31) / 32 * 32);
  // candidates can be a large fraction of n * n_pairs: start from a pass size the queue certainly holds
  while (chunk > 4096 && (size_t)chunk * std::max(s->img.n_pairs, 1) > kMaxItems / 2) chunk = (chunk / 2 + 31) / 32 * 32;
  float *dq_buf = nullptr, *dmin = nullptr;
  int *darg = nullptr;
  unsigned long long *dbest = nullptr;
  CU(bufs.alloc(&dbest, (size_t)chunk));
  if (host) {
    CU(bufs.alloc(&dq_buf, (size_t)chunk * std::max(D, 1)));
    CU(bufs.alloc(&dmin, (size_t)chunk));
    CU(bufs.alloc(&darg, (size_t)chunk));
  }
  long long launches = 0;
  unsigned long long counters_total[4] = {0, 0, 0, 0};
  float ms_cull = 0.f, ms_narrow = 0.f;
  for (long long off = 0; off < n;) {
    const long long m = std::min(chunk, n - off);
    int rc = prepare_work(s, chunk);
    if (rc != MPB_OK) return rc;
    const int nb = std::max(std::max(s->img.n_bins, s->img.n_bins2), 1);
    const float *dq = q + off * D;
    if (host) {
      if (D > 0) CU(cudaMemcpyAsync(dq_buf, q + off * D, (size_t)m * D * sizeof(float), cudaMemcpyHostToDevice, st));
      dq = dq_buf;
    }
    float *omin = host ? dmin : min_dist + off;
    int *oarg = host ? darg : (argmin ? argmin + off : nullptr);
    CU(cudaMemsetAsync(dbest, 0xff, (size_t)m * sizeof(unsigned long long), st));
    CU(cudaMemsetAsync(s->wq.bin_count, 0, (size_t)nb * sizeof(unsigned), st));
    CU(cudaMemsetAsync(s->wq.totals, 0, 8 * sizeof(unsigned), st));
    CU(cudaMemsetAsync(s->wq.counters, 0, 4 * sizeof(unsigned long long), st));
    cudaEvent_t e0 = get_event(s, 0), e1 = get_event(s, 1), e2 = get_event(s, 2);
    CU(cudaEventRecord(e0, st));
    if (s->img.n_pairs > 0) {
      const int cblocks = (int)((m + 127) / 128);
      fk_table_kernel<<<cblocks, 128, 0, st>>>(s->img, s->fk_tab, s->d_qconst, dq, m, s->d_pose, s->pose_ld, s->d_obb);
      sphere_kernel<1, 0, 0><<<std::min(cblocks, s->sphere_blocks[1]), 128, 0, st>>>(s->img, CullTabC<0, 0>{}, s->d_pose, s->pose_ld, m, 0, s->d_sq, s->sq_cap, s->wq.totals);
      dense_kernel<1><<<s->dense_blocks[1], 128, 0, st>>>(s->img, s->d_obb, s->pose_ld, s->d_sq, s->sq_cap, s->wq, HitOut{}, 0);
      scan_bins_kernel<<<1, 32, 0, st>>>(s->wq, s->img.n_bins, s->img.expand_bin, s->img.expand2_bin, (unsigned)dblocks * 8u, s->sq_cap);
      scatter_kernel<false><<<s->num_sms * 2, 1024, 0, st>>>(s->img, s->wq, 0, HitOut{});
      CU(cudaEventRecord(e1, st));
      if (s->smem_verts) distance_kernel<true><<<dblocks, 256, vsm, st>>>(s->img, s->wq, dbest, s->img.n_bins);
      else distance_kernel<false><<<dblocks, 256, 0, st>>>(s->img, s->wq, dbest, s->img.n_bins);
      launches += 4;
      // mesh / octree pairs last: branch and bound against the minimum the convex / primitive pairs left in dbest
      if (s->img.expand_bin >= 0) { distance_expand_kernel<<<s->num_sms * 8, 128, 0, st>>>(s->img, s->wq, dbest); launches += 1; }
      if (s->img.expand2_bin >= 0) { distance_expand2_kernel<<<s->num_sms * 8, 128, 0, st>>>(s->img, s->wq, dbest); launches += 1; }
    } else CU(cudaEventRecord(e1, st));
    distance_unpack_kernel<<<(int)((m + 255) / 256), 256, 0, st>>>(dbest, m, omin, oarg);
    launches += 1;
    CU(cudaEventRecord(e2, st));
    CU(cudaGetLastError());
    unsigned totals[8] = {0};
    unsigned long long counters[4] = {0, 0, 0, 0};
    CU(cudaMemcpyAsync(totals, s->wq.totals, sizeof(totals), cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(counters, s->wq.counters, sizeof(counters), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (totals[2]) {  // candidate queue overflow: redo this pass with a smaller chunk (exact redo)
      if (chunk <= 32) return fail(MPB_ERR_CAPACITY, "work queue overflow persists at the minimum chunk size");
      chunk = std::max<long long>(32, (chunk / 4 + 31) / 32 * 32);
      continue;
    }
    if (host) {
      CU(cudaMemcpyAsync(min_dist + off, dmin, (size_t)m * sizeof(float), cudaMemcpyDeviceToHost, st));
      if (argmin) CU(cudaMemcpyAsync(argmin + off, darg, (size_t)m * sizeof(int), cudaMemcpyDeviceToHost, st));
      CU(cudaStreamSynchronize(st));
    }
    float a = 0.f, b = 0.f;
    cudaEventElapsedTime(&a, e0, e1);
    cudaEventElapsedTime(&b, e1, e2);
    ms_cull += a;
    ms_narrow += b;
    for (int k = 0; k < 4; k++) counters_total[k] += counters[k];
    off += m;
  }
  if (stats) {
    stats->n_states = n;
    stats->pairs_admitted = n * (long long)s->img.n_pairs;
    stats->pairs_after_cull = (long long)counters_total[0];
    stats->narrow_iterations = (long long)counters_total[1];
    stats->narrow_capped = (long long)counters_total[2];
    stats->kernel_launches = launches;
    stats->gpu_ms_cull = ms_cull;
    stats->gpu_ms_narrow = ms_narrow;
    stats->gpu_ms = ms_cull + ms_narrow;
    stats->host_ms_total = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
  }
  return MPB_OK;
}

// ------------------------------------------------------------------------------------------ inverse kinematics
#include "mpb_ik.cuh"

// Chain of movable joints from the root to `link`'s parent joint, and the link placement (and its inverse) in that joint's frame.
static int build_chain(mpb_scene *s, int32_t link, const uint8_t *mask, std::vector<IkJoint> &hc, double *L12, double *Li) {
  if (link < 0 || link >= (int)s->h_linkJoint.size()) return fail(MPB_ERR_INVALID, "The link index is out of bound!");
  std::vector<int> chain;  // model indices, 0 = universe
  for (int j = s->h_linkJoint[link]; j != 0; j = s->h_jparent[j - 1]) chain.push_back(j);
  std::reverse(chain.begin(), chain.end());
  if ((int)chain.size() > MPB_IK_MAX_CHAIN) return fail(MPB_ERR_CAPACITY, "kinematic chain longer than MPB_IK_MAX_CHAIN");
  hc.resize(chain.size());
  for (size_t c = 0; c < chain.size(); c++) {
    const int j = chain[c] - 1;
    IkJoint &K = hc[c];
    memcpy(K.P, s->h_jP64.data() + 12 * (size_t)j, sizeof(K.P));
    memcpy(K.ax, s->h_jax64.data() + 3 * (size_t)j, sizeof(K.ax));
    K.qconst = s->h_qbase64[s->joint_qidx[j]];
    K.type = s->h_jtype[j];
    K.slot = s->h_joints[j].slot;
    K.masked = (K.slot >= 0 && mask && mask[K.slot]) ? 1 : 0;
    K.pad = 0;
  }
  const double *L = s->h_linkL64.data() + 12 * (size_t)link;
  if (L12) memcpy(L12, L, 12 * sizeof(double));
  for (int r = 0; r < 3; r++) {
    for (int c = 0; c < 3; c++) Li[4 * r + c] = L[4 * c + r];
    Li[4 * r + 3] = -(L[r] * L[3] + L[4 + r] * L[7] + L[8 + r] * L[11]);
  }
  return MPB_OK;
}

extern "C" int mpb_ik_clik_batch(mpb_scene *s, int32_t link, const double *poses, const double *q_init, const uint8_t *mask, int64_t n,
                                 double eps, int32_t max_iter, double dt, double damp, double *q_out, uint8_t *success, double *err) {
  if (!s) return fail(MPB_ERR_INVALID, "null scene");
  if (n < 0 || (n > 0 && (!poses || !q_init || !q_out || !success || !err))) return fail(MPB_ERR_INVALID, "null batch pointer");
  if (link < 0 || link >= (int)s->h_linkJoint.size()) return fail(MPB_ERR_INVALID, "The link index is out of bound!");
  if (n == 0) return MPB_OK;
  const int D = s->img.state_dim;
  std::vector<IkJoint> hc;
  double Li[12];
  if (int rc = build_chain(s, link, mask, hc, nullptr, Li)) return rc;
  if (hc.empty()) return fail(MPB_ERR_INVALID, "the link is fixed to the world: nothing to solve for");
  CU(cudaSetDevice(s->device));
  cudaStream_t st = s->stream;
  DevBufs bufs(s);
  IkJoint *d_chain = nullptr;
  double *d_Li = nullptr, *d_pose = nullptr, *d_q0 = nullptr, *d_q = nullptr, *d_err = nullptr;
  unsigned char *d_ok = nullptr;
  CU(bufs.alloc(&d_chain, hc.size()));
  CU(bufs.alloc(&d_Li, (size_t)12));
  CU(bufs.alloc(&d_pose, (size_t)n * 7));
  CU(bufs.alloc(&d_q0, (size_t)n * std::max(D, 1)));
  CU(bufs.alloc(&d_q, (size_t)n * std::max(D, 1)));
  CU(bufs.alloc(&d_err, (size_t)n * 6));
  CU(bufs.alloc(&d_ok, (size_t)n));
  CU(cudaMemcpyAsync(d_chain, hc.data(), hc.size() * sizeof(IkJoint), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_Li, Li, sizeof(Li), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_pose, poses, (size_t)n * 7 * sizeof(double), cudaMemcpyHostToDevice, st));
  if (D > 0) CU(cudaMemcpyAsync(d_q0, q_init, (size_t)n * D * sizeof(double), cudaMemcpyHostToDevice, st));
  ik_clik_kernel<<<(int)((n + 63) / 64), 64, 0, st>>>(d_chain, (int)hc.size(), d_Li, d_pose, d_q0, D, n, eps, max_iter, dt, damp, d_q, d_ok, d_err);
  CU(cudaGetLastError());
  if (D > 0) CU(cudaMemcpyAsync(q_out, d_q, (size_t)n * D * sizeof(double), cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(success, d_ok, (size_t)n, cudaMemcpyDeviceToHost, st));
  CU(cudaMemcpyAsync(err, d_err, (size_t)n * 6 * sizeof(double), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));  // hc / Li go out of scope
  return MPB_OK;
}

extern "C" int mpb_link_jacobian_batch(mpb_scene *s, int32_t link, const double *q, int64_t n, int32_t local, double *jac, double *link_pose) {
  if (!s) return fail(MPB_ERR_INVALID, "null scene");
  if (n < 0 || (n > 0 && (!q || !jac))) return fail(MPB_ERR_INVALID, "null batch pointer");
  if (link < 0 || link >= (int)s->h_linkJoint.size()) return fail(MPB_ERR_INVALID, "The link index is out of bound!");
  if (n == 0) return MPB_OK;
  const int D = s->img.state_dim;
  if (D <= 0) return fail(MPB_ERR_INVALID, "the scene has no batch state");
  std::vector<IkJoint> hc;
  double L[12], Li[12];
  if (int rc = build_chain(s, link, nullptr, hc, L, Li)) return rc;
  CU(cudaSetDevice(s->device));
  cudaStream_t st = s->stream;
  DevBufs bufs(s);
  IkJoint *d_chain = nullptr;
  double *d_L = nullptr, *d_q = nullptr, *d_J = nullptr, *d_T = nullptr;
  CU(bufs.alloc(&d_chain, std::max<size_t>(hc.size(), 1)));
  CU(bufs.alloc(&d_L, (size_t)12));
  CU(bufs.alloc(&d_q, (size_t)n * D));
  CU(bufs.alloc(&d_J, (size_t)n * 6 * D));
  CU(bufs.alloc(&d_T, (size_t)n * 12));
  if (!hc.empty()) CU(cudaMemcpyAsync(d_chain, hc.data(), hc.size() * sizeof(IkJoint), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_L, L, sizeof(L), cudaMemcpyHostToDevice, st));
  CU(cudaMemcpyAsync(d_q, q, (size_t)n * D * sizeof(double), cudaMemcpyHostToDevice, st));
  link_jacobian_kernel<<<(int)((n + 63) / 64), 64, 0, st>>>(d_chain, (int)hc.size(), d_L, d_q, D, n, local, d_J, d_T);
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(jac, d_J, (size_t)n * 6 * D * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (link_pose) CU(cudaMemcpyAsync(link_pose, d_T, (size_t)n * 12 * sizeof(double), cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return MPB_OK;
}
