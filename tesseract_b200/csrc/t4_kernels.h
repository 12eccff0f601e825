/*
 * t4_kernels.h -- per-work-item bodies of every kernel of the batched collision pipeline.
 *
 *   fkBody           joint vector -> world poses of the kinematic links          (StateSolver::getState,
 *                    tesseract_environment/src/kdl/kdl_state_solver.cpp:144-161,232-258; rule ofkt_nodes.cpp:200-289)
 *   broadStaticBody  (state, active shape) x static-environment BVH -> candidate pairs
 *                    (btDbvtBroadphase pair set + TesseractOverlapFilterCallback, bullet_utils.h:864-881,170-177)
 *   broadActiveBody  (state, active-active shape pair) AABB test -> candidate pairs
 *   cullBody         FP32 GJK lower bound vs. the contact cut-off -> survivors (+ early accept for boolean queries)
 *   exactBody        FP64 narrowphase + result bridging / processResult
 *                    (bullet_utils.h:388-441,527-611,757-802 ; core/common.h:114-167)
 *   closestInsert / closestEmit   ContactTestType::CLOSEST reduction per (state, link pair) key
 *
 * The bodies are plain functions of (item index, KParams); t4c.cu wraps them in __global__ kernels for sm_100a.
 */
#ifndef T4_KERNELS_H_
#define T4_KERNELS_H_

#include "../../include/t4c.h"
#include "t4_pair.h"
#include "t4_sincos.h"

namespace t4
{
enum
{
  ST_STATES = 0,
  ST_BVH_NODES,
  ST_CANDIDATES,
  ST_CULLED,
  ST_GJK_PAIRS,
  ST_GJK_ITERS,
  ST_EPA_PAIRS,
  ST_EPA_ITERS,
  ST_CONTACTS,
  ST_COLLIDING,
  ST_CULL_ITERS,     // support pairs of the FP32 cull
  ST_EXACT_PAIRS,    // pairs entering the exact GJK kernel
  ST_P1_ITERS,       // exact kernel: boolean GJK
  ST_P2_ITERS,       // exact kernel: Voronoi GJK + phase 4
  ST_EPA_GJK_ITERS,  // EPA kernel: seed GJK + phase 4
  ST_SUPV,           // hull vertices evaluated by the support mapping, all kernels
  ST_SUPV64,         // ... in the FP64 kernels
  ST_BVH_OVERFLOW,   // traversal stack overflows (the call fails when non-zero)
  ST_P1_SKIPPED,     // exact pairs whose phase-1 outcome was certified by the cull
  ST_SUPV_EPA,       // part of ST_SUPV64 evaluated in the EPA kernel
  ST_P1_VIOLATIONS,  // verify mode: certified pairs whose boolean GJK said "intersect" (must be 0)
  ST_COUNT
};

// work-queue counters (u32, device)
enum
{
  Q_CAND = 0,   // candidate pairs appended by the broadphase
  Q_SURV,       // survivors of the FP32 cull (FIRST with records: one certain contact per state)
  Q_SURV2,      // FIRST with records: one certainly PENETRATING pair per state (looked at only if still no contact)
  Q_SURV3,      // FIRST with records: one certainly penetrating pair per state against a CURVED primitive (slow EPA)
  Q_SURV4,      // FIRST with records: every other survivor (looked at only if still no contact after the others)
  Q_TASK,       // pairs handed from the GJK kernel to the EPA kernel
  Q_FETCH,      // (unused)
  Q_FETCH_EPA,  // (unused)
  Q_RAW,        // raw contacts (CLOSEST staging) or final contacts (ALL / FIRST)
  Q_OUT,        // final contacts emitted by the CLOSEST reduction
  Q_LONG,       // survivors whose GJK ran out of its iteration budget: run again, without one, beside the later passes
  Q_TASK_LONG,  // EPA tasks of the long-pair pass (it runs on its own stream: own task queue)
  Q_LIVE,       // FIRST in rounds: states still without a contact at the start of the round (liveKeyBody)
  Q_COUNT
};

#define T4_MAX_JOINTS 64
// depth-first traversal stack of the environment BVH: a walk keeps at most one pending sibling per level, so a tree of
// depth D needs D entries; the scene compiler refuses trees deeper than this (t4_host_scene.h)
#define T4_BVH_STACK 64

struct alignas(16) DJoint
{
  double origin[12];
  double axis[3];  // normalised
  int32_t type, parent_joint, q_index, pad;
  double pre[12];  // world pose of the parent link when parent_joint < 0
  int32_t alink;   // active-link ordinal whose pose this joint's frame is, or -1
  int32_t keep;    // 1: a joint other than the next one has this joint as its parent (tree branch): keep the frame
  int32_t pad2[2];
};

// one pair handed from the GJK kernel (phases 1-2) to the EPA kernel (phases 3-4)
struct alignas(8) EpaTask
{
  uint32_t cand;
  int32_t is_valid;
  double distance;
  double normalInB[3], pointOnA[3], pointOnB[3], orgNormalInB[3];
};

struct KParams
{
  // ---- scene (read only)
  const DShape* shapes;
  const DStaticW* statics;  // indexed by shape index (valid for static shapes)
  const double* shape_d;    // [shape][16]: params[3], pad, lpose[12]  -- FP64 copies for the exact pass
  const double* statics_d;  // [shape][12]: world pose of static shapes, FP64
  const f4* verts;
  const uint32_t* sup_cells;  // hull support tables (t4_shapes.h), CSR offsets into sup_cand
  const uint16_t* sup_cand;
  const DBvhNode* bvh;
  int32_t n_bvh_nodes;
  const int32_t* active_shapes;
  int32_t n_active_shapes;
  const int32_t* aa_pairs;  // [3 * n_aa_pairs]: shape, active ordinal of the other link, swap flag (t4_host_scene.h)
  int32_t n_aa_pairs;
  const DBvhNode* link_bvh;  // child trees of the kinematic links (boxes in the link frame)
  const int32_t* link_root;  // [n_alinks]
  const uint32_t* acm_bits;  // [n_alinks][acm_words]
  int32_t acm_words;
  const double* pair_margin;  // [n_alinks][n_links] or null
  int32_t n_links, n_alinks;
  float default_margin, max_margin;  // rounded UP (conservative culling)
  double default_margin_d, max_margin_d;  // exact values for the accept / reject decisions
  // ---- FK model
  const DJoint* joints;
  int32_t n_joints, n_q;
  const int32_t* alink_joint;   // [n_alinks] joint index driving the link or -1
  const double* alink_static;   // [n_alinks][12]
  const int32_t* link_joint;    // [n_links] (t4c_fk full output) or null
  const double* link_static;    // [n_links][12]
  // ---- batch
  int64_t n;  // states / segments in this chunk
  int64_t state_base;
  const void* q0;
  const void* q1;
  int32_t q_layout;
  int64_t q_stride_n;  // total n of the SoA array (SoA layout only)
  float* poses0;       // [n_alinks][n][12]: link-major, so that a warp working on 32 states of ONE link
                       // (broadphase item order, FK stores) touches consecutive 48 B records
  float* poses1;
  double* poses0d;     // same poses in FP64 (exact pass: identical inputs to the reference's FP64 path)
  double* poses1d;
  const double* poses_in0;  // FK bypass input (device, double)
  const double* poses_in1;
  double* fk_out;  // t4c_fk
  int32_t cast;
  // ---- request
  int32_t test_type, calc_pen, calc_dist, hit_only;
  int32_t skip_hit;  // FIRST in rounds: states that already have a contact take no further part (ContactTestData::done)
  // FIRST in rounds: the states of this round's broadphase, in Morton order of the round's first shape, states that
  // already have a contact last (marked T4_DEAD_STATE); null: all states in index order
  const uint32_t* live;
  uint32_t n_live;         // entries of `live` the broadphase is launched over
  const uint32_t* live_prev;  // liveKeyBody: the previous round's list (null: identity)
  uint32_t* live_key;      // liveKeyBody out: sort key per entry
  uint32_t* live_val;      // liveKeyBody out: state per entry
  float live_org[3], live_scale[3];  // scene box -> Morton grid
  // ---- queues / outputs
  uint32_t* cand_state;
  uint32_t* cand_a;
  uint32_t* cand_b;
  uint32_t* long_w;       // Q_LONG queue (candidate indices) or null: no budget
  uint32_t long_cap;
  int task_q;             // Q_TASK, or Q_TASK_LONG in the long-pair pass
  int gjk_budget;         // support pairs a pair may spend in phases 1-2 before it is deferred (0: unlimited)
  int verify_skip;        // test hook: run phase 1 for certified pairs too and count contradictions
  uint8_t* cand_class;    // per candidate: iterations the FP32 cull needed (sort key of the exact queues) or null
  // CLOSEST pruning for compound links (octomaps, meshes): only the minimum per (state, link pair) is reported, so the
  // FP32 bounds of the cull decide which pairs of a key can still be that minimum (t4_kernels2.h pruneBody)
  const int32_t* link_group;  // [n_links] dense id of links with more than one shape, else -1; null: pruning off
  int32_t n_groups;
  uint32_t* key_min_ub;       // [n][n_alinks][n_groups] ordered bits of the smallest distance UPPER bound of the key
  float* cand_lb;             // [cand_cap] distance LOWER bound of every surviving candidate
  uint32_t cand_cap;
  uint32_t cull_base;  // first candidate of the range the cull kernel works on (FIRST in rounds: candidates append)
  uint32_t* surv;      // queue read by the GJK kernel in this launch
  uint32_t surv_cap;
  uint32_t* surv_w;    // queues written by the cull kernel
  uint32_t* surv2_w;
  uint32_t* surv3_w;
  uint32_t* surv4_w;
  uint32_t* cull_claim;  // [n] FIRST with records: bit 0 = has its certain separated pair, bit 1 / 2 = its certain penetrating pair (flat / curved)
  EpaTask* tasks;
  const uint32_t* task_order;  // EPA tasks sorted by shape-type pair (null: queue order)
  uint32_t task_cap;
  uint32_t* qcount;       // [Q_COUNT]
  unsigned long long* stats;  // [ST_COUNT]
  uint8_t* hit;           // [n] (chunk-relative) or null
  uint32_t* first_claim;  // [n] FIRST: 0 free / 1 claimed
  t4c_contact* raw;       // staging (CLOSEST) or final buffer
  float* raw_local;       // cast: nearest_points_local [cap][6] or null
  double* raw_d64;        // CLOSEST: FP64 distance of every raw record (the reference compares doubles) or null
  uint32_t raw_cap;
  t4c_contact* out;       // CLOSEST final
  float* out_local;
  uint32_t out_cap;
  int32_t* ht_rep;        // CLOSEST hash table: representative raw index per slot (-1 empty)
  unsigned long long* ht_best;
  uint32_t ht_size;       // power of two
};

// ---------------------------------------------------------------------------------------------------------------
// device / host-simulator primitives
// ---------------------------------------------------------------------------------------------------------------
#ifdef T4_DEVICE_BUILD
// warp-aggregated append: lanes that target the SAME counter are grouped (the tiered survivor queues make lanes of
// one converged warp append to different queues), one atomicAdd per group
__device__ __forceinline__ uint32_t qAppend(uint32_t* ctr)
{
  const unsigned active = __activemask();
  const unsigned m = __match_any_sync(active, (unsigned long long)ctr);
  const int lane = threadIdx.x & 31;
  const int leader = __ffs(m) - 1;
  uint32_t base = 0;
  if (lane == leader)
    base = atomicAdd(ctr, (uint32_t)__popc(m));
  base = __shfl_sync(m, base, leader);
  return base + (uint32_t)__popc(m & ((1u << lane) - 1u));
}
__device__ __forceinline__ void statAdd(unsigned long long* stats, int k, uint32_t v)
{
  const unsigned m = __activemask();
  const uint32_t s = __reduce_add_sync(m, v);
  if ((threadIdx.x & 31) == __ffs(m) - 1 && s)
    atomicAdd(&stats[k], (unsigned long long)s);
}
__device__ __forceinline__ uint32_t casU32(uint32_t* p, uint32_t cmp, uint32_t val) { return atomicCAS(p, cmp, val); }
__device__ __forceinline__ uint32_t orU32(uint32_t* p, uint32_t val) { return atomicOr(p, val); }
__device__ __forceinline__ int32_t casI32(int32_t* p, int32_t cmp, int32_t val) { return atomicCAS(p, cmp, val); }
__device__ __forceinline__ void minU64(unsigned long long* p, unsigned long long v) { atomicMin(p, v); }
__device__ __forceinline__ unsigned long long casU64(unsigned long long* p, unsigned long long cmp, unsigned long long val)
{
  return atomicCAS(p, cmp, val);
}
#else
inline uint32_t qAppend(uint32_t* ctr) { return (*ctr)++; }
inline void statAdd(unsigned long long* stats, int k, uint32_t v) { stats[k] += v; }
inline uint32_t orU32(uint32_t* p, uint32_t val)
{
  uint32_t o = *p;
  *p = o | val;
  return o;
}
inline uint32_t casU32(uint32_t* p, uint32_t cmp, uint32_t val)
{
  uint32_t o = *p;
  if (o == cmp)
    *p = val;
  return o;
}
inline int32_t casI32(int32_t* p, int32_t cmp, int32_t val)
{
  int32_t o = *p;
  if (o == cmp)
    *p = val;
  return o;
}
inline void minU64(unsigned long long* p, unsigned long long v)
{
  if (v < *p)
    *p = v;
}
inline unsigned long long casU64(unsigned long long* p, unsigned long long cmp, unsigned long long val)
{
  unsigned long long o = *p;
  if (o == cmp)
    *p = val;
  return o;
}
#endif

// element offset of the pose of active link `alink` in state s of the current chunk
T4_KD size_t poseOffset(const KParams& P, int64_t s, int alink) { return ((size_t)alink * (size_t)P.n + (size_t)s) * 12; }

template <class T>
T4_KD Xf<T> loadPose(const float* p)
{
  Xf<T> x;
  const f4* q = reinterpret_cast<const f4*>(p);
  const f4 a = q[0], b = q[1], c = q[2];
  x.r[0] = (T)a.x, x.r[1] = (T)a.y, x.r[2] = (T)a.z, x.p.x = (T)a.w;
  x.r[3] = (T)b.x, x.r[4] = (T)b.y, x.r[5] = (T)b.z, x.p.y = (T)b.w;
  x.r[6] = (T)c.x, x.r[7] = (T)c.y, x.r[8] = (T)c.z, x.p.z = (T)c.w;
  return x;
}

// world pose(s) of shape `sh` in state s
template <class T>
T4_KD void shapePose(const KParams& P, const DShape& S, int sh, int64_t s, Xf<T>& x0, Xf<T>& x1)
{
  if (S.alink < 0)
  {
    x0 = loadPose<T>(P.statics[sh].pose);
    x1 = x0;
    return;
  }
  const size_t off = poseOffset(P, s, S.alink);
  x0 = loadPose<T>(P.poses0 + off);
  if (P.cast)
    x1 = loadPose<T>(P.poses1 + off);
  if (!S.identity_lpose)
  {
    const Xf<T> l = loadPose<T>(S.lpose);
    x0 = x0 * l;
    if (P.cast)
      x1 = x1 * l;
  }
  if (!P.cast)
    x1 = x0;
}

// FP64 variant for the exact pass
T4_KD void shapePoseD(const KParams& P, const DShape& S, int sh, int64_t s, Xf<double>& x0, Xf<double>& x1)
{
  if (S.alink < 0)
  {
    x0 = Xf<double>::from12(P.statics_d + (size_t)sh * 12);
    x1 = x0;
    return;
  }
  const size_t off = poseOffset(P, s, S.alink);
  x0 = Xf<double>::from12(P.poses0d + off);
  if (P.cast)
    x1 = Xf<double>::from12(P.poses1d + off);
  if (!S.identity_lpose)
  {
    const Xf<double> l = Xf<double>::from12(P.shape_d + (size_t)sh * 16 + 4);
    x0 = x0 * l;
    if (P.cast)
      x1 = x1 * l;
  }
  if (!P.cast)
    x1 = x0;
}

T4_KD bool acmAllowed(const KParams& P, int alink, int linkB)
{
  return (P.acm_bits[(size_t)alink * P.acm_words + (linkB >> 5)] >> (linkB & 31)) & 1u;
}

T4_KD double pairMargin(const KParams& P, int alink, int linkB)
{
  return P.pair_margin ? P.pair_margin[(size_t)alink * P.n_links + linkB] : P.default_margin_d;
}

// ---------------------------------------------------------------------------------------------------------------
// FK
// ---------------------------------------------------------------------------------------------------------------
T4_KD double loadQ(const KParams& P, const void* q, int64_t s, int j)
{
  if (P.q_layout == T4C_Q_F32_SOA)
    return (double)((const float*)q)[(size_t)j * (size_t)P.q_stride_n + (size_t)s];
  if (P.q_layout == T4C_Q_F64_ROWMAJOR)
    return ((const double*)q)[(size_t)s * P.n_q + j];
  return (double)((const float*)q)[(size_t)s * P.n_q + j];
}

// frame of joint j relative to its parent: T_origin * motion(q_j)
T4_KD Xf<double> fkLocal(const KParams& P, const DJoint& J, const void* q, int64_t s)
{
  Xf<double> local = Xf<double>::from12(J.origin);
  if (J.type == T4C_JOINT_REVOLUTE)
  {
    const double a = loadQ(P, q, s, J.q_index);
    const double sn = t4sin(a), cs = t4cos(a);
    const double t = 1.0 - cs, x = J.axis[0], y = J.axis[1], z = J.axis[2];
    Xf<double> rot;
    rot.r[0] = t * x * x + cs, rot.r[1] = t * x * y - sn * z, rot.r[2] = t * x * z + sn * y;
    rot.r[3] = t * x * y + sn * z, rot.r[4] = t * y * y + cs, rot.r[5] = t * y * z - sn * x;
    rot.r[6] = t * x * z - sn * y, rot.r[7] = t * y * z + sn * x, rot.r[8] = t * z * z + cs;
    local = local * rot;
  }
  else if (J.type == T4C_JOINT_PRISMATIC)
  {
    const double a = loadQ(P, q, s, J.q_index);
    Xf<double> tr;
    tr.p = V3<double>(J.axis[0] * a, J.axis[1] * a, J.axis[2] * a);
    local = local * tr;
  }
  return local;
}

T4_KD void fkOne(const KParams& P, const void* q, int64_t s, Xf<double>* Tj)
{
  for (int j = 0; j < P.n_joints; ++j)
  {
    const DJoint& J = P.joints[j];
    Xf<double> local = Xf<double>::from12(J.origin);
    if (J.type == T4C_JOINT_REVOLUTE)
    {
      // Eigen::AngleAxisd(q, axis.normalized()) (ofkt_nodes.cpp:200,217-219)
      const double a = loadQ(P, q, s, J.q_index);
      // the host libm's bits (t4_sincos.h): link poses identical to the reference's FK on the same host
      const double sn = t4sin(a), cs = t4cos(a);
      const double t = 1.0 - cs, x = J.axis[0], y = J.axis[1], z = J.axis[2];
      Xf<double> rot;
      rot.r[0] = t * x * x + cs, rot.r[1] = t * x * y - sn * z, rot.r[2] = t * x * z + sn * y;
      rot.r[3] = t * x * y + sn * z, rot.r[4] = t * y * y + cs, rot.r[5] = t * y * z - sn * x;
      rot.r[6] = t * x * z - sn * y, rot.r[7] = t * y * z + sn * x, rot.r[8] = t * z * z + cs;
      local = local * rot;
    }
    else if (J.type == T4C_JOINT_PRISMATIC)
    {
      // T_origin * Translation(q * axis.normalized()) (ofkt_nodes.cpp:272,287-289)
      const double a = loadQ(P, q, s, J.q_index);
      Xf<double> tr;
      tr.p = V3<double>(J.axis[0] * a, J.axis[1] * a, J.axis[2] * a);
      local = local * tr;
    }
    const Xf<double> parent = J.parent_joint >= 0 ? Tj[J.parent_joint] : Xf<double>::from12(J.pre);
    Tj[j] = parent * local;
  }
}

T4_KD void storePoseF(float* dst, const Xf<double>& x)
{
  f4* d = reinterpret_cast<f4*>(dst);
  f4 a, b, c;
  a.x = (float)x.r[0], a.y = (float)x.r[1], a.z = (float)x.r[2], a.w = (float)x.p.x;
  b.x = (float)x.r[3], b.y = (float)x.r[4], b.z = (float)x.r[5], b.w = (float)x.p.y;
  c.x = (float)x.r[6], c.y = (float)x.r[7], c.z = (float)x.r[8], c.w = (float)x.p.z;
  d[0] = a, d[1] = b, d[2] = c;
}

// world pose of active link `a` from the joint frames of one state
T4_KD Xf<double> fkAlinkPose(const KParams& P, const Xf<double>* Tj, int a)
{
  const int j = P.alink_joint[a];
  return j >= 0 ? Tj[j] : Xf<double>::from12(P.alink_static + (size_t)a * 12);
}

// t4c_fk: poses of ALL links of state s, FP64
T4_KD void fkStoreAllLinks(const KParams& P, const Xf<double>* Tj, int64_t s)
{
  for (int l = 0; l < P.n_links; ++l)
  {
    const int j = P.link_joint[l];
    const Xf<double> x = j >= 0 ? Tj[j] : Xf<double>::from12(P.link_static + (size_t)l * 12);
    x.to12(P.fk_out + ((size_t)s * P.n_links + l) * 12);
  }
}

// item = state.  Writes poses0 (and poses1 for cast) of the kinematic links, or fk_out (all links, FP64).
// (Serial form: host simulator.  The device kernel stages the poses of a warp through shared memory and writes whole
// 128-byte lines, t4c_exact.cu fkKernelT.)
template <int MAXJ>
T4_KD void fkBodyT(int64_t s, const KParams& P)
{
  Xf<double> Tj[MAXJ];  // joint frames of this state (thread-local; 96 B each)
  for (int pass = 0; pass < (P.cast ? 2 : 1); ++pass)
  {
    fkOne(P, pass ? P.q1 : P.q0, s, Tj);
    if (P.fk_out)
    {
      fkStoreAllLinks(P, Tj, s);
      return;
    }
    float* dst = pass ? P.poses1 : P.poses0;
    double* dstd = pass ? P.poses1d : P.poses0d;
    for (int a = 0; a < P.n_alinks; ++a)
    {
      const Xf<double> x = fkAlinkPose(P, Tj, a);
      storePoseF(dst + poseOffset(P, s, a), x);
      x.to12(dstd + poseOffset(P, s, a));
    }
  }
}

T4_KD void fkBody(int64_t s, const KParams& P) { fkBodyT<T4_MAX_JOINTS>(s, P); }

// item = (state, active link): FK bypass, FP64 host poses -> FP32 device poses
T4_KD void convertPosesBody(int64_t i, const KParams& P)
{
  const int64_t s = i / P.n_alinks;
  const int a = (int)(i - s * P.n_alinks);
  const size_t o = poseOffset(P, s, a);
  const double* src0 = P.poses_in0 + (size_t)i * 12;
  storePoseF(P.poses0 + o, Xf<double>::from12(src0));
  for (int k = 0; k < 12; ++k)
    P.poses0d[o + k] = src0[k];
  if (P.cast)
  {
    const double* src1 = P.poses_in1 + (size_t)i * 12;
    storePoseF(P.poses1 + o, Xf<double>::from12(src1));
    for (int k = 0; k < 12; ++k)
      P.poses1d[o + k] = src1[k];
  }
}

// ---------------------------------------------------------------------------------------------------------------
// LVS subdivision of trajectories (tesseract_environment/core/utils.h:189-205 continuous, :328-347 discrete)
// ---------------------------------------------------------------------------------------------------------------
struct LvsParams
{
  const double* traj;  // [n_traj][n_wp][n_q]
  int64_t n_traj;
  int32_t n_wp, n_q, cast;
  double lvs;
  uint64_t* counts;         // [n_traj * steps] work items of every waypoint step
  const uint64_t* offsets;  // exclusive prefix sum of counts
  double* q0;               // [n_items][n_q]
  double* q1;               // cast only
  T4_HD int steps() const { return cast ? n_wp - 1 : n_wp; }
};

// Eigen::VectorXd::LinSpaced(cnt, low, high)(i) (core/utils.h:194-200)
T4_HD double lvsLinspaced(int64_t cnt, double low, double high, int64_t i)
{
  const int64_t size1 = cnt == 1 ? 1 : cnt - 1;
  const double step = cnt == 1 ? 0.0 : (high - low) / double(cnt - 1);
  const bool flip = t4abs(high) < t4abs(low);
  if (flip)
    return (i == 0) ? low : (high - double(size1 - i) * step);
  return (i == size1) ? high : (low + double(i) * step);
}

// number of LinSpaced points of waypoint step (a -> b), or 0 when the step is not subdivided
T4_HD int64_t lvsPoints(const LvsParams& L, const double* a, const double* b, bool has_next)
{
  if (!has_next)
    return 0;
  double acc = 0;
  for (int k = 0; k < L.n_q; ++k)
    acc += (b[k] - a[k]) * (b[k] - a[k]);
  const double dist = t4sqrt(acc);
  if (dist > 0 && dist > L.lvs)
    return (int64_t)ceil(dist / L.lvs) + 1;
  return 0;
}

// item = t * steps + s
T4_KD void lvsCountBody(int64_t i, const LvsParams& L)
{
  const int steps = L.steps();
  const int64_t t = i / steps;
  const int s = (int)(i - t * steps);
  const double* a = L.traj + ((size_t)t * L.n_wp + s) * L.n_q;
  const int64_t cnt = lvsPoints(L, a, a + L.n_q, s < L.n_wp - 1);
  L.counts[i] = cnt > 0 ? (uint64_t)(cnt - 1) : 1ull;
}

T4_KD void lvsExpandBody(int64_t i, const LvsParams& L)
{
  const int steps = L.steps();
  const int64_t t = i / steps;
  const int s = (int)(i - t * steps);
  const double* a = L.traj + ((size_t)t * L.n_wp + s) * L.n_q;
  const double* b = a + L.n_q;
  const int64_t cnt = lvsPoints(L, a, b, s < L.n_wp - 1);
  const size_t base = (size_t)L.offsets[i];
  if (cnt == 0)
  {
    for (int k = 0; k < L.n_q; ++k)
    {
      L.q0[base * L.n_q + k] = a[k];
      if (L.cast)
        L.q1[base * L.n_q + k] = b[k];
    }
    return;
  }
  for (int64_t j = 0; j < cnt - 1; ++j)
    for (int k = 0; k < L.n_q; ++k)
    {
      L.q0[(base + j) * L.n_q + k] = lvsLinspaced(cnt, a[k], b[k], j);
      if (L.cast)
        L.q1[(base + j) * L.n_q + k] = lvsLinspaced(cnt, a[k], b[k], j + 1);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// broadphase
// ---------------------------------------------------------------------------------------------------------------
T4_KD bool boxesOverlap(const V3<float>& mn, const V3<float>& mx, const float* bmn, const float* bmx)
{
  return !(mn.x > bmx[0] || mx.x < bmn[0] || mn.y > bmx[1] || mx.y < bmn[1] || mn.z > bmx[2] || mx.z < bmn[2]);
}

T4_KD void shapeWorldBox(const KParams& P, const DShape& S, const Xf<float>& x0, const Xf<float>& x1, V3<float>& mn,
                         V3<float>& mx)
{
  const V3<float> c(S.bc[0], S.bc[1], S.bc[2]), h(S.bh[0], S.bh[1], S.bh[2]);
  boxAabb(x0, c, h, mn, mx);
  if (P.cast && S.alink >= 0)
  {
    V3<float> mn1, mx1;
    boxAabb(x1, c, h, mn1, mx1);
    mn = V3<float>(t4min(mn.x, mn1.x), t4min(mn.y, mn1.y), t4min(mn.z, mn1.z));
    mx = V3<float>(t4max(mx.x, mx1.x), t4max(mx.y, mx1.y), t4max(mx.z, mx1.z));
  }
}

T4_KD void pushCandidate(const KParams& P, int64_t s, int a, int b)
{
  const uint32_t k = qAppend(&P.qcount[Q_CAND]);
  if (k < P.cand_cap)
  {
    P.cand_state[k] = (uint32_t)s;
    P.cand_a[k] = (uint32_t)a;
    P.cand_b[k] = (uint32_t)b;
  }
}

// conservative slack (metres) absorbing FP32 rounding of poses / boxes in every cull
#define T4_CULL_SLACK 2e-5f

// item = a_ord * n + s  (shape-major, so that a warp works on 32 states of the SAME robot shape)
// resident 128-thread CTAs per SM of the EPA kernel (register budget 65536 / (128 * this)); sweeps: profiles/r2a, r3z
#ifndef T4_EPA_MIN_CTAS
#define T4_EPA_MIN_CTAS 4
#endif
#define T4_DEAD_STATE 0xffffffffu
#define T4_LIVE_BITS 7 /* Morton grid: 2^7 cells per axis; key = 21 bits + the dead flag */

T4_KD uint32_t spread3(uint32_t v)
{
  // bit i of the low 10 bits -> bit 3 i
  v = (v | (v << 16)) & 0x030000ffu;
  v = (v | (v << 8)) & 0x0300f00fu;
  v = (v | (v << 4)) & 0x030c30c3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}

// item = entry of the previous round's state list.  Key of a state for this round's broadphase: dead flag (the state
// has its contact) above the Morton code of the world position of the round's first shape.  After the radix sort the
// 32 states of a warp have that shape in the same 3 cm cell of the scene: they walk the same nodes of the
// environment tree, and the states that are done sit at the end of the list where whole warps leave at once.
// (Before: a warp took 32 consecutive states of a random batch -- 8 of 32 lanes active in the traversal, and in the
// later rounds most lanes belonged to states that were done.)
T4_KD void liveKeyBody(int64_t i, const KParams& P)
{
  const uint32_t s = P.live_prev ? P.live_prev[i] : (uint32_t)i;
  uint32_t key = 1u << (3 * T4_LIVE_BITS);
  if (s != T4_DEAD_STATE && !(P.skip_hit && P.hit[s]))
  {
    const int sh = P.active_shapes[0];
    const DShape& S = P.shapes[sh];
    Xf<float> x0, x1;
    shapePose<float>(P, S, sh, (int64_t)s, x0, x1);
    const V3<float> c = x0(V3<float>(S.bc[0], S.bc[1], S.bc[2]));
    const float hi = (float)((1 << T4_LIVE_BITS) - 1);
    const float gx = t4min(t4max((c.x - P.live_org[0]) * P.live_scale[0], 0.f), hi);
    const float gy = t4min(t4max((c.y - P.live_org[1]) * P.live_scale[1], 0.f), hi);
    const float gz = t4min(t4max((c.z - P.live_org[2]) * P.live_scale[2], 0.f), hi);
    key = spread3((uint32_t)gx) | (spread3((uint32_t)gy) << 1) | (spread3((uint32_t)gz) << 2);
    qAppend(&P.qcount[Q_LIVE]);
  }
  P.live_key[i] = key;
  P.live_val[i] = key >> (3 * T4_LIVE_BITS) ? T4_DEAD_STATE : s;
}

T4_KD void broadStaticBody(int64_t item, const KParams& P)
{
  const int64_t per_shape = P.live ? (int64_t)P.n_live : P.n;
  const int a_ord = (int)(item / per_shape);
  int64_t s = item - (int64_t)a_ord * per_shape;
  if (P.live)
  {
    const uint32_t ls = P.live[s];
    if (ls == T4_DEAD_STATE)
      return;
    s = (int64_t)ls;
  }
  if (P.skip_hit && P.hit[s])
    return;
  const int sh = P.active_shapes[a_ord];
  const DShape& S = P.shapes[sh];
  Xf<float> x0, x1;
  shapePose<float>(P, S, sh, s, x0, x1);
  V3<float> mn, mx;
  shapeWorldBox(P, S, x0, x1, mn, mx);
  const float grow = P.max_margin + T4_CULL_SLACK;
  mn = mn - V3<float>(grow, grow, grow);
  mx = mx + V3<float>(grow, grow, grow);
  const V3<float> bc(S.bc[0], S.bc[1], S.bc[2]), bh(S.bh[0], S.bh[1], S.bh[2]);
  const V3<float> cw = x0(bc);
  const float ra = S.bc[3];
  uint32_t nodes = 0, culled = 0, cands = 0, overflow = 0;
  int stack[T4_BVH_STACK];
  int sp = 0;
  stack[sp++] = 0;
  while (sp > 0)
  {
    const int ni = stack[--sp];
    const DBvhNode& N = P.bvh[ni];
    ++nodes;
    for (int side = 0; side < 2; ++side)
    {
      const float* bmn = side ? N.rmn : N.lmn;
      const float* bmx = side ? N.rmx : N.lmx;
      if (!boxesOverlap(mn, mx, bmn, bmx))
        continue;
      const int child = side ? N.right : N.left;
      if (child >= 0)
      {
        if (sp < T4_BVH_STACK)
          stack[sp++] = child;
        else
          ++overflow;  // cannot happen for a tree the scene compiler accepted (depth check); reported, never silent
        continue;
      }
      const int b = ~child;
      const DShape& Sb = P.shapes[b];
      if (Sb.link == S.link || acmAllowed(P, S.alink, Sb.link))
        continue;
      if (!P.cast)
      {
        // bounding-sphere and link-frame box culls (exact tests follow in the narrowphase)
        const DStaticW& W = P.statics[b];
        const V3<float> cb(W.bs[0], W.bs[1], W.bs[2]);
        const float reach = W.bs[3] + grow;
        const V3<float> d = cb - cw;
        const float rr = ra + reach;
        if (d.len2() > rr * rr)
        {
          ++culled;
          continue;
        }
        const V3<float> pl = x0.invApply(cb) - bc;
        const V3<float> e(t4max(t4abs(pl.x) - bh.x, 0.f), t4max(t4abs(pl.y) - bh.y, 0.f),
                          t4max(t4abs(pl.z) - bh.z, 0.f));
        if (e.len2() > reach * reach)
        {
          ++culled;
          continue;
        }
      }
      ++cands;
      pushCandidate(P, s, sh, b);
    }
  }
  statAdd(P.stats, ST_BVH_NODES, nodes);
  statAdd(P.stats, ST_CULLED, culled);
  statAdd(P.stats, ST_CANDIDATES, cands + culled);
  if (overflow)
    statAdd(P.stats, ST_BVH_OVERFLOW, overflow);
}

// Kinematic x kinematic broadphase.  item = e * n + s ; entry e = (shape A of one kinematic link, the OTHER link of an
// admissible link pair): the margin-grown world box of A is taken into the other link's frame and walks that link's
// child tree -- what the reference's compound algorithms do with the compound's btDbvt
// (tesseract_compound_collision_algorithm.cpp:323-338, ..._compound_compound_...:256-325).  Single-shape links have a
// one-leaf tree, so robot self-collision pairs of convex hull links cost one box test as before.
T4_KD void broadActiveBody(int64_t item, const KParams& P)
{
  const int e = (int)(item / P.n);
  const int64_t s = item - (int64_t)e * P.n;
  if (P.skip_hit && P.hit[s])
    return;
  const int a = P.aa_pairs[3 * e], alb = P.aa_pairs[3 * e + 1], swap = P.aa_pairs[3 * e + 2];
  const DShape& Sa = P.shapes[a];
  Xf<float> a0, a1;
  shapePose<float>(P, Sa, a, s, a0, a1);
  V3<float> mn, mx;
  shapeWorldBox(P, Sa, a0, a1, mn, mx);
  // the narrowphase keeps a pair only when its distance is within the margin, and a box distance never exceeds the true
  // distance: growing ONE side by the max margin loses nothing (the reference grows both, bullet_utils.h:170-177, which
  // only adds candidates its narrowphase then rejects)
  const float grow = P.max_margin + T4_CULL_SLACK;
  const V3<float> cw = (mn + mx) * 0.5f;
  const V3<float> hw = (mx - mn) * 0.5f + V3<float>(grow, grow, grow);
  // the box in the other link's frame (conservative: box of the rotated box); cast: union over both poses of that link
  V3<float> qmn(1e30f, 1e30f, 1e30f), qmx(-1e30f, -1e30f, -1e30f);
  for (int pass = 0; pass < (P.cast ? 2 : 1); ++pass)
  {
    const Xf<float> xb = loadPose<float>((pass ? P.poses1 : P.poses0) + poseOffset(P, s, alb));
    const V3<float> c = xb.invApply(cw);
    const V3<float> h(t4abs(xb.r[0]) * hw.x + t4abs(xb.r[3]) * hw.y + t4abs(xb.r[6]) * hw.z,
                      t4abs(xb.r[1]) * hw.x + t4abs(xb.r[4]) * hw.y + t4abs(xb.r[7]) * hw.z,
                      t4abs(xb.r[2]) * hw.x + t4abs(xb.r[5]) * hw.y + t4abs(xb.r[8]) * hw.z);
    const float pad = 1e-5f;  // FP32 rounding of the frame change
    qmn = V3<float>(t4min(qmn.x, c.x - h.x - pad), t4min(qmn.y, c.y - h.y - pad), t4min(qmn.z, c.z - h.z - pad));
    qmx = V3<float>(t4max(qmx.x, c.x + h.x + pad), t4max(qmx.y, c.y + h.y + pad), t4max(qmx.z, c.z + h.z + pad));
  }
  uint32_t cands = 0, overflow = 0;
  int stack[T4_BVH_STACK];
  int sp = 0;
  stack[sp++] = P.link_root[alb];
  while (sp > 0)
  {
    const DBvhNode& N = P.link_bvh[stack[--sp]];
    for (int side = 0; side < 2; ++side)
    {
      if (!boxesOverlap(qmn, qmx, side ? N.rmn : N.lmn, side ? N.rmx : N.lmx))
        continue;
      const int child = side ? N.right : N.left;
      if (child >= 0)
      {
        if (sp < T4_BVH_STACK)
          stack[sp++] = child;
        else
          ++overflow;
        continue;
      }
      ++cands;
      if (swap)
        pushCandidate(P, s, ~child, a);
      else
        pushCandidate(P, s, a, ~child);
    }
  }
  statAdd(P.stats, ST_CANDIDATES, cands);
  if (overflow)
    statAdd(P.stats, ST_BVH_OVERFLOW, overflow);
}

}  // namespace t4
#endif
