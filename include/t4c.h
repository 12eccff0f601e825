/*
 * t4c.h -- C ABI of the B200-native collision-query engine ("tesseract_b200").
 *
 * This is the drop-in boundary for ONE hot path of ros-industrial-consortium/tesseract
 * (v0.5.0): batched discrete and continuous (cast-hull) contact tests of a kinematic
 * robot against itself and its environment.  Everything here is plain C: pointers,
 * sizes, PODs.  No torch / Eigen / Bullet types.  The C++ classes in
 * tesseract_b200/csrc/b200_contact_managers.h (DiscreteContactManager /
 * ContinuousContactManager implementations) and the Python test tooling both sit on
 * top of exactly these entry points.
 *
 * Reference interfaces each entry replaces (paths relative to the reference checkout):
 *   t4c_scene_create / _clone / _destroy
 *       DiscreteContactManager::addCollisionObject / clone
 *       tesseract_collision/include/tesseract_collision/core/discrete_contact_manager.h:65-110
 *       (shape conversion rules: tesseract_collision/src/bullet/bullet_utils.cpp:59-365)
 *   t4c_scene_set_active / _set_enabled / _set_acm / _set_margins / _set_link_pose
 *       setActiveCollisionObjects / enable|disableCollisionObject / setIsContactAllowedFn /
 *       setCollisionMarginData / setCollisionObjectsTransform (static links)
 *       discrete_contact_manager.h:112-200, continuous_contact_manager.h:112-250
 *   t4c_fk_model_create, t4c_fk
 *       StateSolver::getState (tesseract_environment/src/kdl/kdl_state_solver.cpp:144-161,232-258)
 *       ForwardKinematics::calcFwdKinAll (tesseract_kinematics/src/kdl/kdl_fwd_kin_chain.cpp:71-106)
 *   t4c_check_discrete / t4c_check_link_poses_discrete
 *       checkTrajectoryState + DiscreteContactManager::contactTest
 *       (tesseract_environment/include/tesseract_environment/core/utils.h:129-160;
 *        tesseract_collision/src/bullet/bullet_discrete_bvh_manager.cpp:267-283)
 *   t4c_check_cast / t4c_check_link_poses_cast / t4c_check_trajectories_cast
 *       checkTrajectorySegment / checkTrajectory(LVS_CONTINUOUS) + ContinuousContactManager::contactTest
 *       (core/utils.h:87-119,172-297; tesseract_collision/src/bullet/bullet_cast_bvh_manager.cpp:260-332,446-462)
 *   t4c_check_trajectories_discrete
 *       checkTrajectory(LVS_DISCRETE) (core/utils.h:309-428)
 *
 * Threading: one t4c_scene per host thread (the reference's "clone per thread" rule,
 * discrete_contact_manager.h:59-65).  All calls are synchronous with respect to the caller.
 * There is NO CPU fallback: every compute entry fails with T4C_ERR_CUDA when no device
 * is usable.
 */
#ifndef T4C_H_
#define T4C_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define T4C_ABI_VERSION 2

typedef int32_t t4c_status;
enum {
  T4C_OK = 0,
  T4C_ERR_INVALID = -1,     /* bad argument / inconsistent description */
  T4C_ERR_CUDA = -2,        /* CUDA runtime error or no device; see t4c_last_error() */
  T4C_ERR_CAPACITY = -3,    /* contact buffer too small: n_contacts holds the required count */
  T4C_ERR_UNSUPPORTED = -4, /* e.g. cast check of a non-convex leaf (reference throws runtime_error) */
  T4C_ERR_NOMEM = -5
};

/* Shape kinds after the reference's conversion rules (bullet_utils.cpp:59-243):
 * Mesh -> one T4C_TRIANGLE per face, Octree -> one BOX / SPHERE per occupied leaf. */
enum {
  T4C_SPHERE = 0,   /* params[0] = r (core is a point, margin r) */
  T4C_BOX = 1,      /* params = half extents */
  T4C_CYLINDER = 2, /* Z axis; params[0] = r, params[1] = half length */
  T4C_CONE = 3,     /* Z axis, origin at mid height; params[0] = r, params[1] = height */
  T4C_CAPSULE = 4,  /* Z axis; params[0] = r, params[1] = half length of the core segment */
  T4C_HULL = 5,     /* convex vertex cloud: vertex_offset / vertex_count */
  T4C_TRIANGLE = 6  /* 3 vertices at vertex_offset */
};

typedef struct t4c_shape {
  int32_t type;
  int32_t link;          /* owning link */
  int32_t shape_id;      /* ContactResult::shape_id  (index in the link's geometry list) */
  int32_t subshape_id;   /* ContactResult::subshape_id (triangle / voxel index) or -1 */
  double params[3];
  int32_t vertex_offset; /* into t4c_scene_desc::vertices (HULL, TRIANGLE) */
  int32_t vertex_count;
  double pose[12];       /* shape pose in the link frame, 3x4 row-major [R|p] */
} t4c_shape;

typedef struct t4c_link {
  const char* name; /* used only for diagnostics; pair keys are built by the caller */
  int32_t type_id;  /* user mask id (ContactResult::type_id) */
  int32_t enabled;
  int32_t active;   /* 1: kinematic (pose given per state); 0: static */
  double pose[12];  /* world pose (static links; start pose for active links) */
} t4c_link;

typedef struct t4c_scene_desc {
  int32_t n_links;
  const t4c_link* links;
  int32_t n_shapes;
  const t4c_shape* shapes;
  int32_t n_vertices;
  const double* vertices;         /* xyz triples */
  int32_t n_acm_pairs;
  const int32_t* acm_pairs;       /* 2*n link indices: contact allowed => pair never tested */
  double default_margin;          /* CollisionMarginData default */
  int32_t n_margin_pairs;
  const int32_t* margin_pairs;    /* 2*n link indices */
  const double* margin_values;    /* n per-pair margins */
} t4c_scene_desc;

enum { T4C_JOINT_FIXED = 0, T4C_JOINT_REVOLUTE = 1, T4C_JOINT_PRISMATIC = 2 };

typedef struct t4c_joint {
  int32_t type;
  int32_t parent_link;
  int32_t child_link;
  int32_t q_index;   /* column of the joint vector, -1 for fixed joints */
  double origin[12]; /* parent_to_joint_origin_transform, 3x4 row-major */
  double axis[3];
} t4c_joint;

/* Joints must be ordered parent-before-child.  Links that no joint reaches keep their
 * scene pose.  root_link's pose is taken from the scene (t4c_link::pose). */
typedef struct t4c_fk_model_desc {
  int32_t n_links;
  int32_t n_joints;
  const t4c_joint* joints;
  int32_t n_q;
  int32_t root_link;
} t4c_fk_model_desc;

enum { T4C_TEST_FIRST = 0, T4C_TEST_CLOSEST = 1, T4C_TEST_ALL = 2 };

typedef struct t4c_request {
  int32_t type;                  /* ContactTestType */
  int32_t calculate_penetration; /* ContactRequest::calculate_penetration */
  int32_t calculate_distance;    /* ContactRequest::calculate_distance */
} t4c_request;

enum { T4C_CC_NONE = 0, T4C_CC_TIME0 = 1, T4C_CC_TIME1 = 2, T4C_CC_BETWEEN = 3 };

/* One contact, 80 bytes, 16-byte aligned (five 128-bit stores on the device). */
typedef struct t4c_contact {
  int32_t state;          /* state index (discrete) or sub-segment index (cast) */
  int32_t link[2];
  int32_t subshape_id[2];
  int16_t shape_id[2];
  float distance;
  float point[2][3];      /* nearest points on link[0], link[1]; world frame */
  float normal[3];        /* from link[0] towards link[1] */
  float cc_time[2];       /* cast only, else -1 */
  uint8_t cc_type[2];     /* T4C_CC_* */
  uint8_t reserved[2];
  uint32_t flags;         /* bit 0: produced by the penetration (EPA) pass */
} t4c_contact;

/* Work counters of one call (roofline accounting).  An "iteration" is one support-PAIR evaluation
 * (support of A and of B along one direction).  The FP32 cull, the exact FP64 GJK kernel and the EPA kernel are
 * counted separately so that each kernel's algorithmic work can be recomputed from the line alone. */
typedef struct t4c_stats {
  uint64_t states;            /* states / sub-segments processed */
  uint64_t bvh_nodes;         /* BVH node visits (two child boxes tested per visit) */
  uint64_t candidates;        /* (shape, shape) pairs passing the AABB + filter tests */
  uint64_t culled;            /* candidates rejected by the bounding-sphere / link-box pretests */
  uint64_t gjk_pairs;         /* pairs entering the FP32 cull GJK */
  uint64_t gjk_iterations;    /* total support-pair evaluations: cull + exact GJK (phases 1, 2, 4) + EPA kernel's GJK */
  uint64_t epa_pairs;         /* pairs entering the penetration solver */
  uint64_t epa_iterations;    /* polytope expansion iterations (one support pair each) */
  uint64_t contacts;          /* records emitted */
  uint64_t colliding_states;
  /* ---- ABI 2 */
  uint64_t cull_iterations;   /* FP32 cull kernel: support pairs */
  uint64_t exact_pairs;       /* pairs entering the FP64 exact GJK kernel (survivors of the cull) */
  uint64_t phase1_iterations; /* exact GJK kernel: support pairs of the boolean GJK */
  uint64_t phase2_iterations; /* exact GJK kernel: support pairs of the Voronoi GJK + the 3 of phase 4 */
  uint64_t epa_gjk_iterations;/* EPA kernel: support pairs of its seed GJK + the 3 of phase 4 */
  uint64_t support_vertices;  /* hull vertices whose dot product was evaluated, all kernels (table scans touch few) */
  uint64_t support_vertices_fp64; /* ... of which in the FP64 kernels */
  uint64_t bvh_stack_overflows;   /* traversal stack overflows (must be 0: the call fails otherwise) */
  uint64_t phase1_skipped;    /* exact pairs whose boolean GJK outcome was already certified by the cull */
  uint64_t support_vertices_epa;  /* part of support_vertices_fp64 evaluated in the EPA kernel */
  uint64_t phase1_skip_violations; /* T4C_VERIFY_SKIP test hook: certified pairs whose boolean GJK disagreed (must be 0) */
  uint64_t closest_pruned;    /* CLOSEST: cull survivors whose bounds show they cannot be the minimum of their (state,
                                 link pair) key -- compound links (octomaps, meshes) -- and that skip the exact pass */
} t4c_stats;

typedef struct t4c_scene t4c_scene;
typedef struct t4c_fk_model t4c_fk_model;

/* Host-buffer result view (caller owns all memory). */
typedef struct t4c_result {
  uint8_t* hit;          /* optional, n bytes: 1 if the state has >=1 accepted contact */
  t4c_contact* contacts; /* optional, capacity records */
  int64_t capacity;
  int64_t n_contacts;    /* out */
  t4c_stats stats;       /* out */
  float* cc_local;       /* optional, cast checks only: 6 floats per contact = ContactResult::nearest_points_local
                            of the cast side(s) (bullet_utils.h:452-525; the "Between" case moves it to the
                            mid-point of the two supports); rows of non-cast sides are zero */
} t4c_result;

/* Device-buffer variant: every pointer is a device pointer on the scene's device. */
enum { T4C_Q_F32_SOA = 0, T4C_Q_F64_ROWMAJOR = 1, T4C_Q_F32_ROWMAJOR = 2 };
typedef struct t4c_device_io {
  const void* q;          /* joint states, layout per q_layout; SoA: q[j*n + i] */
  int32_t q_layout;
  uint8_t* hit;           /* optional */
  t4c_contact* contacts;  /* optional */
  int64_t capacity;
  float* cc_local;        /* optional (cast): see t4c_result::cc_local */
  int64_t n_contacts;     /* out (host value) */
  t4c_stats stats;        /* out (host value) */
} t4c_device_io;

const char* t4c_last_error(void);
int32_t t4c_abi_version(void);
/* "cuda sm_100a" for the product library; "hostsim" for the host simulation of the kernel bodies that the CPU test
 * tier builds (tests/hostsim) -- the Python binding refuses to load the latter as the product library */
const char* t4c_build_info(void);
int32_t t4c_device_count(void);

t4c_status t4c_scene_create(const t4c_scene_desc* desc, int32_t device, t4c_scene** out);
t4c_status t4c_scene_clone(const t4c_scene* scene, t4c_scene** out);
void t4c_scene_destroy(t4c_scene* scene);
t4c_status t4c_scene_set_active(t4c_scene* scene, const int32_t* active_flags /* n_links */);
t4c_status t4c_scene_set_enabled(t4c_scene* scene, int32_t link, int32_t enabled);
t4c_status t4c_scene_set_acm(t4c_scene* scene, int32_t n_pairs, const int32_t* pairs);
t4c_status t4c_scene_set_margins(t4c_scene* scene, double default_margin, int32_t n_pairs,
                                 const int32_t* pairs, const double* values);
t4c_status t4c_scene_set_link_pose(t4c_scene* scene, int32_t link, const double* pose12);
/* several links at once (one BVH refit for all moved static objects); unchanged poses cost nothing */
t4c_status t4c_scene_set_link_poses(t4c_scene* scene, int32_t n, const int32_t* links, const double* poses12);
int32_t t4c_scene_num_active(const t4c_scene* scene); /* kinematic links, ascending link index */
/* addCollisionObject / removeCollisionObject on a live scene (discrete_contact_manager.h:75-98).  Shape records: link
 * field ignored, vertex offsets relative to `vertices`.  An existing link name is replaced.  A removed object keeps its
 * link index (shapeless, disabled) so that the indices of the others do not move.  Moving a STATIC object with
 * t4c_scene_set_link_pose refits the environment BVH (no rebuild); adding / removing rebuilds it at the next check. */
t4c_status t4c_scene_add_object(t4c_scene* scene, const t4c_link* link, int32_t n_shapes, const t4c_shape* shapes,
                                int32_t n_vertices, const double* vertices, int32_t* link_index);
t4c_status t4c_scene_remove_object(t4c_scene* scene, int32_t link);

t4c_status t4c_fk_model_create(const t4c_fk_model_desc* desc, const t4c_scene* scene, t4c_fk_model** out);
void t4c_fk_model_destroy(t4c_fk_model* model);

/* Batched forward kinematics: q is n x n_q row-major doubles (host); out is n x n_links x 12 doubles
 * (every scene link; links no joint reaches keep their scene pose). */
t4c_status t4c_fk(const t4c_fk_model* model, const t4c_scene* scene, const double* q, int64_t n, double* link_poses);
/* t4c_fk for callers that may not allocate device memory (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments
 * and results as t4c_fk (bit for bit: the same kernel); q and the poses are staged in a fixed 8 MiB array of device
 * memory that belongs to the library image (T4C_FK_NOALLOC_SCRATCH_BYTES; every GPU that loads the library holds it
 * until exit), so n * (n_q * 8 + n_links * 96) + 256 bytes must fit: T4C_ERR_CAPACITY otherwise.  The scene and the
 * model must already be resident on the device (one earlier t4c_fk / check call after the last edit): this entry never
 * uploads or allocates, it returns T4C_ERR_INVALID instead.  All calls on a device share the array (contents of the
 * previous call are overwritten, nothing needs clearing); overlapping use on one GPU would race, so calls of one
 * process are serialised per device inside the entry. */
#define T4C_FK_NOALLOC_SCRATCH_BYTES (8u << 20)
t4c_status t4c_fk_noalloc(const t4c_fk_model* model, const t4c_scene* scene, const double* q, int64_t n,
                          double* link_poses);

/* Geometric Jacobian of `link` for n joint vectors (host buffers): J is n x 6 x n_q row-major, rows 0-2 linear and 3-5
 * angular velocity of the link per unit joint velocity, world frame, taken at the link origin or at point_local (a
 * point fixed in the link frame; may be NULL).  ForwardKinematics::calcJacobian
 * (tesseract_kinematics/src/kdl/kdl_fwd_kin_chain.cpp:127-178) + jacobianChangeRefPoint. */
t4c_status t4c_jacobian(const t4c_fk_model* model, const t4c_scene* scene, const double* q, int64_t n, int32_t link,
                        const double* point_local, double* J);
/* d(distance)/dq of contact records: grad is n_contacts x n_q, grad[i] = n_i . (J_lin,link1(p1) - J_lin,link0(p0)) with
 * the record's world points and normal (TrajOpt's collision-term gradient).  contacts[i].state indexes q.  The device
 * variant takes device pointers (q in any t4c_device_io layout) so that records never leave the GPU. */
t4c_status t4c_contact_gradients(t4c_scene* scene, const t4c_fk_model* model, const double* q, int64_t n,
                                 const t4c_contact* contacts, int64_t n_contacts, double* grad);
t4c_status t4c_contact_gradients_device(t4c_scene* scene, const t4c_fk_model* model, const void* q, int32_t q_layout,
                                        int64_t n, const t4c_contact* contacts, int64_t n_contacts, double* grad);

/* Discrete batched check: FK + self + environment test of n joint vectors (host buffers). */
t4c_status t4c_check_discrete(t4c_scene* scene, const t4c_fk_model* model, const double* q, int64_t n,
                              const t4c_request* request, t4c_result* result);
/* Same, FK bypassed: poses = n x n_active x 12 doubles (active links in ascending link index). */
t4c_status t4c_check_link_poses_discrete(t4c_scene* scene, const double* poses, int64_t n,
                                         const t4c_request* request, t4c_result* result);
/* Cast check of n independent segments q0[i] -> q1[i] (each n x n_q row-major). */
t4c_status t4c_check_cast(t4c_scene* scene, const t4c_fk_model* model, const double* q0, const double* q1,
                          int64_t n, const t4c_request* request, t4c_result* result);
t4c_status t4c_check_link_poses_cast(t4c_scene* scene, const double* poses0, const double* poses1, int64_t n,
                                     const t4c_request* request, t4c_result* result);
/* LVS trajectory checks (core/utils.h:172-428): traj is n_traj x n_wp x n_q row-major doubles.
 * Work items are LVS sub-states (discrete) / sub-segments (cast); result contact.state and the hit array are
 * indexed by the flat work-item index (t4c_lvs_items maps it back to trajectory / step / sub-step).  n_items
 * receives the item count.  lvs = +infinity disables the subdivision (CollisionEvaluatorType::DISCRETE /
 * CONTINUOUS). */
t4c_status t4c_check_trajectories_discrete(t4c_scene* scene, const t4c_fk_model* model, const double* traj,
                                           int64_t n_traj, int32_t n_wp, double lvs, const t4c_request* request,
                                           t4c_result* result, int64_t* n_items);
t4c_status t4c_check_trajectories_cast(t4c_scene* scene, const t4c_fk_model* model, const double* traj,
                                       int64_t n_traj, int32_t n_wp, double lvs, const t4c_request* request,
                                       t4c_result* result, int64_t* n_items);

/* number of LVS work items checkTrajectory would visit (core/utils.h:189-205,328-347); no device work */
t4c_status t4c_lvs_count(const double* traj, int64_t n_traj, int32_t n_wp, int32_t n_q, double lvs, int32_t cast,
                         int64_t* n_items);
/* Map flat work-item indices (t4c_contact::state / the hit array of t4c_check_trajectories_*) back to
 * (trajectory, waypoint step iStep, sub-step iSubStep) in the reference's visiting order.  Each output array is
 * optional and holds `capacity` entries; items beyond the capacity are only counted.  Host only. */
t4c_status t4c_lvs_items(const double* traj, int64_t n_traj, int32_t n_wp, int32_t n_q, double lvs, int32_t cast,
                         int64_t capacity, int32_t* item_traj, int32_t* item_step, int32_t* item_sub, int64_t* n_items);
/* Joint vectors of selected work items (strictly ascending flat indices): q0, and q1 for cast sub-segments, are
 * n_sel x n_q row-major -- the subtraj.row(iSubStep) / row(iSubStep + 1) of core/utils.h:194-205.  Host only. */
t4c_status t4c_lvs_item_states(const double* traj, int64_t n_traj, int32_t n_wp, int32_t n_q, double lvs, int32_t cast,
                               int64_t n_sel, const int64_t* items, double* q0, double* q1);

/* Device-resident entries (inputs already in HBM; used by bench.py "value" and by multi-GPU
 * callers that gather with NCCL themselves). */
t4c_status t4c_check_discrete_device(t4c_scene* scene, const t4c_fk_model* model, int64_t n,
                                     const t4c_request* request, t4c_device_io* io);
t4c_status t4c_check_cast_device(t4c_scene* scene, const t4c_fk_model* model, int64_t n /* segments */,
                                 const void* q1, const t4c_request* request, t4c_device_io* io);
/* ---- multi-GPU: contiguous shard of the state batch per GPU, scene replicated, ONE gather of the
 * per-GPU results (hit flags + contact records + counters) to the root GPU over NVLink:
 *     ncclGroupStart(); ncclSend (ranks != root) / ncclRecv (root); ncclGroupEnd();
 * Replaces the reference's clone-per-thread parallelism (discrete_contact_manager.h:59-65;
 * test_suite/collision_multi_threaded_unit.hpp:79-127).  NCCL is loaded with dlopen("libnccl.so.2"). */
typedef struct t4c_comm t4c_comm;
/* rank 0 makes the 128-byte id (ncclUniqueId) and hands it to the other ranks by any transport */
t4c_status t4c_comm_unique_id(uint8_t* id128);
t4c_status t4c_comm_create(const uint8_t* id128, int32_t rank, int32_t world, int32_t device, t4c_comm** out);
/* one process driving n devices: out receives n communicators, out[i] for devices[i] (ncclCommInitAll) */
t4c_status t4c_comm_create_local(const int32_t* devices, int32_t n, t4c_comm** out);
void t4c_comm_destroy(t4c_comm* comm);
/* gathered results on the root (device buffers on the root's device, caller-owned; hit / contacts optional) */
typedef struct t4c_gathered {
  uint8_t* hit;          /* one byte per state of the WHOLE batch, indexed by state_offset + local index */
  t4c_contact* contacts; /* shards concatenated in rank order; t4c_contact::state is the global state index */
  int64_t capacity;
  int64_t n_contacts;    /* out */
  int64_t n_states;      /* out: states of all shards */
  t4c_stats stats;       /* out: counters summed over the ranks */
} t4c_gathered;
/* Every rank calls with ITS shard (io: device buffers of this rank; io->q holds n_local joint vectors, records are
 * also left in io->contacts); state_offset = global index of the shard's first state.  `gathered` on the root only. */
t4c_status t4c_check_discrete_sharded(t4c_scene* scene, const t4c_fk_model* model, t4c_comm* comm, int32_t root,
                                      int64_t n_local, int64_t state_offset, const t4c_request* request,
                                      t4c_device_io* io, t4c_gathered* gathered);
t4c_status t4c_check_cast_sharded(t4c_scene* scene, const t4c_fk_model* model, t4c_comm* comm, int32_t root,
                                  int64_t n_local, int64_t state_offset, const void* q1, const t4c_request* request,
                                  t4c_device_io* io, t4c_gathered* gathered);
/* One process, one host thread per device: scenes[i] / comms[i] live on device i of t4c_comm_create_local; q is the
 * whole batch (n x n_q doubles, host); the result (host buffers) is what t4c_check_discrete would return for it. */
t4c_status t4c_check_discrete_multi(t4c_scene* const* scenes, const t4c_fk_model* model, t4c_comm* const* comms,
                                    int32_t n_devices, const double* q, int64_t n, const t4c_request* request,
                                    t4c_result* result);

/* number of this library's kernel launches since load (bench.py "gpu_launches") */
uint64_t t4c_kernel_launch_count(void);
/* Per-kernel device timing (CUDA events recorded on the launching stream around every launch of this library).
 * Kernel order: fk, convert_poses, broad_static, broad_active, cull_fp32, gjk_fp64, epa_fp64, closest_insert, closest_emit,
 * count_hit.  Process-wide, not thread safe: meant for bench.py / profiling runs only. */
void t4c_profile_enable(int32_t on);
void t4c_profile_reset(void);
int32_t t4c_profile_get(double* ms, uint64_t* launches, int32_t capacity); /* returns the number of kernels */
/* Test hook, host only: hull support of `shape` for n directions (xyz triples) through the support table and by the
 * plain scan; out[4*i..4*i+3] = vertex index {table FP64, scan FP64, table FP32, scan FP32}; mean_candidates =
 * average table cell length (-1: the hull has no table). */
t4c_status t4c_debug_hull_support(t4c_scene* scene, int32_t shape, const double* dirs, int64_t n, int32_t* out,
                                  double* mean_candidates);
/* Test hook: sine and cosine of n doubles as the FK kernel computes them (t4_sincos.h), evaluated ON THE DEVICE. */
t4c_status t4c_debug_sincos(int32_t device, const double* x, int64_t n, double* sin_out, double* cos_out);
/* measured-peak helpers: an FP32 FFMA loop / an FP64 DFMA loop on every SM, TFLOP/s (roofline denominators) */
t4c_status t4c_measure_fp32_peak(int32_t device, double* tflops);
t4c_status t4c_measure_fp64_peak(int32_t device, double* tflops);

#ifdef __cplusplus
}
#endif
#endif /* T4C_H_ */
