/*
 * mpb.h -- C ABI of the B200 batched state-validity engine ("libmpb.so").
 *
 * The reference (KolinGuo/MPlib @ c647ab9) has no C ABI: its pybind11 module binds the C++
 * classes directly (python/pybind.cpp:16-27).  This header is the boundary the replacement
 * introduces: the host layer that mirrors the reference classes (mplib_b200/pymp) compiles a
 * PlanningWorld into the flat `mpb_scene_desc` below and calls these entry points; each entry
 * point names the reference function(s) whose per-state work it performs for a whole batch.
 *
 * Conventions: plain C structs, caller-owned memory, no exceptions across the boundary.  Every
 * function returns MPB_OK (0) or a negative status; `mpb_last_error()` gives a thread-local
 * message.  Matrices are row-major 3x4 ([R|p], 12 doubles).  Pose vectors are
 * [x, y, z, qw, qx, qy, qz] (src/math_utils.cpp:12-18).
 */
#ifndef MPB_H_
#define MPB_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPB_OK 0
#define MPB_ERR_INVALID (-1)   /* bad argument / malformed description (reference: std::invalid_argument / ASSERT) */
#define MPB_ERR_CUDA (-2)      /* CUDA runtime failure, or no usable device */
#define MPB_ERR_CAPACITY (-3)  /* a static capacity of this build was exceeded */
#define MPB_ERR_UNSUPPORTED (-4)

/* joint types: Pinocchio JointModelR* / JointModelP* / JointModelRUB* (src/pinocchio_model.cpp:69-81) */
#define MPB_JOINT_REVOLUTE 0
#define MPB_JOINT_PRISMATIC 1
#define MPB_JOINT_CONTINUOUS 2

/* collision geometry types (python/pybind_fcl.hpp:59-239, the subset the hot path meets) */
#define MPB_GEOM_CONVEX 0   /* fcl::Convex    : obj_geom = hull index                         */
#define MPB_GEOM_SPHERE 1   /* fcl::Sphere    : param = {radius}                              */
#define MPB_GEOM_BOX 2      /* fcl::Box       : param = {side x, side y, side z}              */
#define MPB_GEOM_CAPSULE 3  /* fcl::Capsule   : param = {radius, lz}                          */
#define MPB_GEOM_CYLINDER 4 /* fcl::Cylinder  : param = {radius, lz}                          */
#define MPB_GEOM_MESH 5     /* fcl::BVHModel<OBBRSS> : obj_geom = mesh index                  */
#define MPB_GEOM_OCTREE 6   /* fcl::OcTree    : obj_geom = octree index                       */

/* pointer-space flags for batch calls */
#define MPB_HOST 0    /* q / outputs are host pointers; copies happen inside the call */
#define MPB_DEVICE 1  /* q / outputs are device pointers on the scene's device       */

/*
 * Flat description of one PlanningWorld (src/planning_world.h:138-141 holds the same content as
 * maps of shared_ptr).  All articulations are merged into one kinematic forest; `q_full` is the
 * concatenation of every articulation's full qpos; a batch *state* (length `state_dim`) writes
 * the entries `state_map[k]` of `q_full` -- the move-group scatter of
 * ArticulatedModelTpl::setQpos (src/articulated_model.cpp:108-113) composed with the name-ordered
 * split of PlanningWorldTpl::setQposAll (src/planning_world.cpp:249-262); all other entries keep
 * `q_base` (the articulations' current_qpos_).
 */
typedef struct mpb_scene_desc {
  /* movable joints, topological order; model index 0 is the universe, joint i has model index i+1 */
  int32_t n_joints;
  const int32_t *joint_parent;    /* [n_joints] model index of the parent joint                */
  const double *joint_placement;  /* [n_joints][12]                                            */
  const int32_t *joint_type;      /* [n_joints] MPB_JOINT_*                                    */
  const double *joint_axis;       /* [n_joints][3] unit axis in the joint frame                */
  const int32_t *joint_qidx;      /* [n_joints] index into q_full                              */
  int32_t nq_full;
  const double *q_base;           /* [nq_full]                                                 */
  int32_t state_dim;
  const int32_t *state_map;       /* [state_dim]                                               */

  /* link frames (Pinocchio BODY frames): oMf = oMi[link_joint] * link_placement */
  int32_t n_links;
  const int32_t *link_joint;      /* [n_links] model joint index (0 = universe)                */
  const double *link_placement;   /* [n_links][12]                                             */

  /* collision objects: pose = linkpose(obj_link) * obj_local, or obj_local if obj_link < 0 */
  int32_t n_objects;
  const int32_t *obj_type;        /* [n_objects] MPB_GEOM_*                                    */
  const int32_t *obj_link;        /* [n_objects] link index or -1 (world-fixed)                */
  const double *obj_local;        /* [n_objects][12]                                           */
  const double *obj_param;        /* [n_objects][4]                                            */
  const int32_t *obj_geom;        /* [n_objects] hull / mesh / octree index, else -1           */

  /* convex hull pool */
  int32_t n_hulls;
  const int32_t *hull_vert_off;   /* [n_hulls+1] */
  const double *hull_verts;       /* [.][3]      */
  const int32_t *hull_face_off;   /* [n_hulls+1] */
  const int32_t *hull_faces;      /* [.][3] vertex indices local to the hull */

  /* triangle mesh pool */
  int32_t n_meshes;
  const int32_t *mesh_vert_off;   /* [n_meshes+1] */
  const double *mesh_verts;       /* [.][3]       */
  const int32_t *mesh_tri_off;    /* [n_meshes+1] */
  const int32_t *mesh_tris;       /* [.][3]       */

  /* octree pool: occupied finest-level voxels as integer cell coordinates floor(p/res) */
  int32_t n_octrees;
  const double *oct_res;          /* [n_octrees]   */
  const int32_t *oct_cell_off;    /* [n_octrees+1] */
  const int32_t *oct_cells;       /* [.][3]        */

  /* ACM-admitted object pairs, argument order of ::fcl::collide(a, b) preserved */
  int32_t n_pairs;
  const int32_t *pair_a;
  const int32_t *pair_b;
} mpb_scene_desc;

typedef struct mpb_scene mpb_scene; /* opaque: device image + work queues + streams */

/* Per-call counters (SURVEY 5 "metrics": new; the reference has none). */
typedef struct mpb_stats {
  int64_t n_states;
  int64_t pairs_admitted;    /* n_states * n_pairs                                   */
  int64_t pairs_after_cull;  /* items that reached an exact narrowphase              */
  int64_t pairs_after_spheres; /* survivor-queue slots after the bounding-sphere phase (incl. chunk padding) */
  int64_t narrow_iterations; /* support-mapping evaluations (MPR/GJK iterations)     */
  int64_t narrow_capped;     /* items that hit the iteration cap (reported, not hidden) */
  int64_t kernel_launches;   /* kernels launched by this call                        */
  float gpu_ms;              /* device time of the call's kernels (CUDA events)      */
  float gpu_ms_cull;         /* FK + broadphase kernels                              */
  float gpu_ms_narrow;       /* narrowphase kernel(s)                                */
  float host_ms_enqueue;     /* host time spent issuing the call's copies and launches */
  float host_ms_total;       /* host wall time of the call, including the final synchronisation */
  float gpu_ms_fk;           /* fk_table_kernel alone (device-resident input; 0 when the input is streamed from the host) */
  int32_t fk_bytes_per_state; /* algorithmic bytes fk_table_kernel moves per state: 4*state_dim read + sphere and OBB table rows written */
  int32_t async_pending;     /* 1: the call returned after enqueueing; its counters arrive through mpb_scene_poll */
} mpb_stats;

/* Thread-local message of the last failing call on this thread. */
const char *mpb_last_error(void);

/* Number of CUDA devices visible (0 if none); never fails. */
int mpb_device_count(void);

/* Compile `desc` into a device image on CUDA device `device`.
 * Replaces the per-state reads of PlanningWorldTpl's maps (src/planning_world.cpp:276-481). */
int mpb_scene_create(const mpb_scene_desc *desc, int device, mpb_scene **out);
void mpb_scene_destroy(mpb_scene *scene);

/* Replace q_base (ArticulatedModelTpl::current_qpos_, src/articulated_model.cpp:104-113). */
int mpb_scene_set_q_base(mpb_scene *scene, const double *q_base, int32_t nq_full);

/* Replace the admitted pair list (after an ACM edit, src/collision_matrix.cpp:34-138). */
int mpb_scene_set_pairs(mpb_scene *scene, int32_t n_pairs, const int32_t *pair_a, const int32_t *pair_b);

/* Update objects of the compiled scene in place: the pose of a world-fixed object (CollisionObject::setTransformation,
 * python/pybind_fcl.hpp:254-259), the pose of an attached body on its link (src/attached_body.h:48-54) and / or the
 * parameters of a primitive.  Geometry type and hull / mesh / octree stay what they were; nothing is re-compiled.
 *   obj_idx [n]   obj_local [n][12] nullable (as mpb_scene_desc::obj_local)   obj_param [n][4] nullable */
int mpb_scene_update_objects(mpb_scene *scene, int32_t n, const int32_t *obj_idx, const double *obj_local,
                             const double *obj_param);

/* Batched PinocchioModelTpl::computeForwardKinematics + getLinkPose for every link
 * (src/pinocchio_model.cpp:271-312).  q: [n][state_dim] float32; poses: [n][n_links][7] float32. */
int mpb_fk_batch(mpb_scene *scene, const float *q, int64_t n, float *poses, int space, void *stream);

/* mpb_fk_batch for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.
 * Calls no cudaMalloc / cudaFree / allocating routine: the joint-frame table of the batch lives in a fixed array of
 * MPB_NOALLOC_SCRATCH_BYTES inside the library image (held by every GPU that loads the library, until exit).
 *   q_dev [n][state_dim] float32 and poses_dev [n][n_links][7] float32 are DEVICE pointers; the call returns after the
 *   enqueue on `stream` (NULL: the scene's stream) and the caller synchronises.
 *   n x (joints + 1) x 48 bytes must fit the array (Panda: 34 944 configurations per call); more: MPB_ERR_CAPACITY.
 * All calls on one GPU share the array: calls that overlap on the GPU (different streams) race; calls on one stream do
 * not.  The array keeps the last call's contents; nothing in it needs clearing (every word is written before it is read).
 * Same arithmetic as mpb_fk_batch: the results are identical. */
#define MPB_NOALLOC_SCRATCH_BYTES (16u << 20)
int mpb_fk_batch_noalloc(mpb_scene *scene, const float *q_dev, int64_t n, float *poses_dev, void *stream);

/* Batched ValidityCheckerTpl::isValid = setQposAll + collideFull (src/ompl_planner.h:59-62,
 * src/planning_world.cpp:249-262, 483-490).
 *   q            [n][state_dim] float32
 *   verdict_bits [ceil(n/32)] uint32, bit (i&31) of word i>>5 set iff state i collides
 *   pair_bits    nullable, [n][ceil(n_pairs/32)] uint32, bit p set iff admitted pair p collides
 * `stream` is a cudaStream_t (NULL = the scene's own stream).
 * MPB_HOST: the call returns after the results are in the host buffers.  Batches of at most 4096 states on the
 *   scene's own stream are replayed as one CUDA graph (copies + kernels) through pinned staging buffers, so a legacy
 *   single-state query costs one graph launch.
 * MPB_DEVICE: the call returns after ENQUEUEING on `stream` (stats->async_pending = 1); nothing synchronises.  The
 *   passes are sized from the last blocking call on the scene, so the first device-pointer call after
 *   mpb_scene_create / mpb_scene_set_pairs / mpb_scene_update_objects runs the blocking protocol once (it returns
 *   with the stream idle and full stats).  If a later call still overflows the bounded work queues its outputs are
 *   incomplete; that is never silent: mpb_scene_poll (and through it the caller) learns it as MPB_ERR_CAPACITY,
 *   and the next call re-calibrates. */
int mpb_collide_batch(mpb_scene *scene, const float *q, int64_t n, uint32_t *verdict_bits,
                      uint32_t *pair_bits, mpb_stats *stats, int space, void *stream);

/* mpb_collide_batch with the per-pair result as a SPARSE list instead of the dense mask: entry k of `hits` is
 * (configuration index, admitted pair index) of a colliding pair -- the content of the WorldCollisionResult list of
 * PlanningWorldTpl::collideFull (src/planning_world.cpp:483-490), which also holds colliding pairs only.  Order is
 * unspecified.  hits [hit_cap][2] uint32, n_hits [1] uint32 (same space as the other buffers); *n_hits may exceed
 * hit_cap, in which case the list is truncated to hit_cap entries.  The dense mask stays in device memory. */
int mpb_collide_batch_hits(mpb_scene *scene, const float *q, int64_t n, uint32_t *verdict_bits, uint32_t *hits,
                           int64_t hit_cap, uint32_t *n_hits, mpb_stats *stats, int space, void *stream);

/* Harvest the device-pointer calls that have finished (wait != 0: wait for all of them first).
 *   sum              nullable; counters and times summed over the calls harvested since the previous poll
 *                    (gpu_ms* only when mpb_scene_set_async_timing is on)
 *   calls_completed  nullable; how many calls that is        calls_failed  nullable; how many of them overflowed
 * Returns MPB_ERR_CAPACITY when calls_failed > 0 (the outputs of those calls are incomplete). */
int mpb_scene_poll(mpb_scene *scene, int wait, mpb_stats *sum, int64_t *calls_completed, int64_t *calls_failed);

/* Record per-pass CUDA timing events for device-pointer calls too (off by default: five events per pass). */
int mpb_scene_set_async_timing(mpb_scene *scene, int on);

/* Batched ValidityCheckerTpl::clearance = setQposAll + distanceFull (src/ompl_planner.h:69-72,
 * src/planning_world.cpp:492-720) with the default DistanceRequest (python/pybind_fcl.hpp:311-315):
 *   min_dist [n] float32   minimum over the ACM-admitted pairs of ::fcl::distance: the separation distance, or -1 when
 *                          some admitted pair touches / intersects; FLT_MAX when no pair is admitted
 *   argmin   [n] int32     nullable; index of the pair attaining it (first such pair in list order), -1 if none
 * Pairs that involve a mesh (BVHModel) or an octree give the minimum over their leaf primitives, by branch and bound
 * over the object's tree (the role of the Mesh... / OcTree... distance traversal nodes of FCL); overlapping triangles of a
 * mesh x mesh pair give 0 (FCL's triDistance), every other touching / intersecting leaf -1. */
int mpb_distance_batch(mpb_scene *scene, const float *q, int64_t n, float *min_dist, int32_t *argmin, mpb_stats *stats,
                       int space, void *stream);

/*
 * Planning state space as OMPLPlannerTpl::build_state_space builds it (src/ompl_planner.cpp:245-296): one 1-D
 * RealVector subspace per bounded joint, one SO(2) subspace per continuous joint, all with weight 1, so the
 * space distance is the sum of the per-dimension distances.  `seg[j]` is the longest valid segment of subspace j:
 * 0.01 * (hi[j] - lo[j]) for RealVector, 0.01 * pi for SO(2) (OMPL's default 1 % validity resolution, which
 * the reference never changes).
 */
typedef struct mpb_space_desc {
  int32_t dim;           /* == state_dim of the scene */
  const float *lo;       /* [dim] lower_joint_limits_ (-pi for SO(2))  */
  const float *hi;       /* [dim] upper_joint_limits_ (+pi for SO(2))  */
  const float *seg;      /* [dim] longest valid segment per subspace   */
  const uint8_t *so2;    /* [dim] 1 = SO(2) subspace; nullable = none  */
} mpb_space_desc;

/* Batched ompl::base::DiscreteMotionValidator::checkMotion(a, b) with ValidityCheckerTpl::isValid as the state
 * test (the validator OMPL installs for the reference's SpaceInformation, src/ompl_planner.cpp:62-70):
 * valid iff b and every interior sample a + (j/nd)(b-a), j = 1..nd-1, is collision-free, where
 * nd = max_j ceil(|b_j - a_j| / seg[j]).  `a` is assumed valid, as in OMPL.
 *   a, b        [n_edges][dim] float32
 *   valid_bits  [ceil(n_edges/32)] uint32, bit e set iff motion e is valid
 * stats->n_states reports the number of sampled states. */
int mpb_check_motion_batch(mpb_scene *scene, const mpb_space_desc *space, const float *a, const float *b, int64_t n_edges,
                           uint32_t *valid_bits, mpb_stats *stats, int space_flag, void *stream);

#define MPB_RRT_ACTIVE 0         /* not finished within max_rounds / time_limit ("Timeout" in OMPL terms) */
#define MPB_RRT_SOLVED 1         /* "Exact solution" */
#define MPB_RRT_CAPACITY 2       /* a tree reached tree_capacity nodes */
#define MPB_RRT_INVALID_GOAL 3   /* no goal state given ("Invalid goal") */
#define MPB_RRT_PATH_TOO_LONG 4  /* solved, but the path has more than max_path waypoints */

typedef struct mpb_rrt_params {
  float range;            /* RRTConnect::setRange; <= 0: OMPL's default 0.2 * maximum extent of the space */
  int32_t max_rounds;     /* lock-step rounds at most (a round consumes 1 .. connect_steps growTree calls per query) */
  float time_limit_s;     /* wall-clock budget of the call (OMPL's `time`), <= 0: none                */
  int32_t tree_capacity;  /* nodes per tree per query                                                 */
  int32_t max_path;       /* waypoints per returned path at most                                      */
  uint64_t seed;
  int32_t check_every;    /* rounds between host checks of the active-query counter (<= 0: 16)        */
  int32_t connect_steps;  /* steps of the connect loop validated per round (<= 0: 8, at most 32)      */
  int32_t n_slots;        /* concurrent searches (<= 0 or > n_queries: n_queries).  A slot whose query ends takes the
                             next pending query, so a batch larger than n_slots flows through without idle slots */
  int32_t max_rounds_per_query; /* the PER-QUERY budget (the role of OMPL's `time` for one query): rounds a query
                             may occupy a slot before it ends as MPB_RRT_ACTIVE ("Timeout"); <= 0: none */
} mpb_rrt_params;

typedef struct mpb_rrt_stats {
  int64_t rounds;          /* lock-step rounds executed                       */
  int64_t grow_calls;      /* growTree calls summed over queries              */
  int64_t states_checked;  /* states sent through the validity kernels        */
  int64_t n_solved;
  int64_t kernel_launches;
  float gpu_ms;            /* device time of the whole call (CUDA events)     */
  float host_ms_total;
} mpb_rrt_stats;

/* RRTConnect branch of OMPLPlannerTpl::plan (src/ompl_planner.cpp:97-243) for n_queries independent queries in
 * lock-step on the device; start and goal states must already be valid (the host layer runs the start fix-up and
 * the goal-image enumeration of src/ompl_planner.cpp:110-158 first).  Host pointers only.
 *   starts   [n_queries][dim]             goals  [n_queries][max_goals][dim]     n_goals [n_queries]
 *   status   [n_queries] MPB_RRT_*        path_len [n_queries]
 *   paths    [n_queries][max_path][dim]   start ... goal waypoints of solved queries */
int mpb_rrtconnect_batch(mpb_scene *scene, const mpb_space_desc *space, const float *starts, const float *goals,
                         const int32_t *n_goals, int32_t max_goals, int64_t n_queries, const mpb_rrt_params *params,
                         int32_t *status, int32_t *path_len, float *paths, mpb_rrt_stats *stats);

/* Query geometry constants of the compiled scene (for tests / host layer). */
int mpb_scene_info(const mpb_scene *scene, int32_t *n_pairs, int32_t *state_dim, int32_t *n_links,
                   int32_t *n_objects, int32_t *device);

/* FP32 FMA micro-benchmark on `device`: achieved TFLOP/s of a register-resident FFMA chain on every SM
 * (the in-run confirmation of SMs x 128 x 2 x f that BASELINE.md 2 asks for; no reference counterpart). */
int mpb_fp32_peak(int device, float *tflops);

/* Diagnostics (no reference counterpart): the support cell table the scene compiler builds for a convex hull of <= 255
 * vertices -- uint16 first[6 G G + 1] (G = 8 cells per cube-map face edge), then uint8 candidate vertex indices per cell.
 * The hull support on the device scans only the candidates of the direction's cell; tests check on the host that the
 * arg-max vertex of any direction is among them.  Needs no GPU.  out may be NULL (size query). */
int mpb_debug_cell_table(const double *verts, int32_t n_verts, uint8_t *out, int64_t cap, int64_t *n_bytes);

/* PinocchioModelTpl::computeIKCLIK (src/pinocchio_model.cpp:377-432) for n independent problems, double precision:
 * damped least squares on the 6 x nv joint Jacobian of `link`'s parent joint until |log6 error| < eps or max_iter.
 * The unknowns are the scene's batch state (state_dim joints); joints outside it keep their q_base value.
 *   link     user link index (as mpb_fk_batch orders its output)
 *   poses    [n][7] goal pose of the link: x y z, qw qx qy qz
 *   q_init   [n][state_dim]        mask [state_dim] or NULL: 1 = joint frozen (its Jacobian column is zeroed)
 *   q_out    [n][state_dim]        success [n] 0/1        err [n][6] last log6 error (linear, angular)
 * Reference defaults: eps 1e-5, max_iter 1000, dt 0.1, damp 1e-12 (python/pybind_pinocchio.hpp:54-57).  Host pointers. */
int mpb_ik_clik_batch(mpb_scene *scene, int32_t link, const double *poses, const double *q_init, const uint8_t *mask, int64_t n,
                      double eps, int32_t max_iter, double dt, double damp, double *q_out, uint8_t *success, double *err);

/* PinocchioModelTpl::computeFullJacobian + getLinkJacobian(index, local) (src/pinocchio_model.cpp:335-360) for n states,
 * double precision: the 6 x state_dim Jacobian of `link`'s parent joint, rows (linear, angular), in Pinocchio's WORLD
 * frame (local == 0) or in the link frame (local != 0; equals computeSingleLinkLocalJacobian, :362-375).  Columns follow
 * the scene's batch state; joints off the chain give zero columns.
 *   q [n][state_dim]   jac [n][6][state_dim]   link_pose [n][12] nullable: the link's world placement, 3x4 row-major.
 * Host pointers. */
int mpb_link_jacobian_batch(mpb_scene *scene, int32_t link, const double *q, int64_t n, int32_t local, double *jac, double *link_pose);

#ifdef __cplusplus
}
#endif
#endif /* MPB_H_ */
