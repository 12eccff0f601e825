/*
 * t4o.h -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * FP64 restatement of the reference's collision hot path (tesseract v0.5.0 Bullet
 * managers + the Bullet / KDL arithmetic they call).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this library.  The product
 * (tesseract_b200/csrc) never links or calls it.
 *
 * PARITY STATUS: pinned against every known-answer test the reference holds for this
 * path (tests/test_oracle_golden.py) at the reference's own
 * tolerances (1e-3 / 1e-4).  Parity tighter than that against a real Bullet build is
 * UNPINNED in this container (Bullet, Eigen, KDL are not installed and cannot be built
 * offline): the Bullet-internal pieces (btVoronoiSimplexSolver, btGjkEpa2, sphere-sphere,
 * sphere-triangle, shape support / AABB functions) are restated from the published
 * Bullet 2.8x/3.x algorithms in double precision (BT_USE_DOUBLE_PRECISION constants).
 */
#ifndef T4O_H_
#define T4O_H_

#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "../include/t4c.h"

namespace t4o
{
// Bullet double-precision constants (btScalar.h): SIMD_EPSILON = DBL_EPSILON, BT_LARGE_FLOAT = 1e30
constexpr double SIMD_EPS = std::numeric_limits<double>::epsilon();
constexpr double LARGE = 1e30;

struct V3
{
  double x = 0, y = 0, z = 0;
  V3() = default;
  V3(double a, double b, double c) : x(a), y(b), z(c) {}
  double& operator[](int i) { return (&x)[i]; }
  double operator[](int i) const { return (&x)[i]; }
  V3 operator+(const V3& o) const { return { x + o.x, y + o.y, z + o.z }; }
  V3 operator-(const V3& o) const { return { x - o.x, y - o.y, z - o.z }; }
  V3 operator-() const { return { -x, -y, -z }; }
  V3 operator*(double s) const { return { x * s, y * s, z * s }; }
  V3 operator/(double s) const { return { x / s, y / s, z / s }; }
  V3& operator+=(const V3& o) { x += o.x; y += o.y; z += o.z; return *this; }
  V3& operator-=(const V3& o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
  V3& operator*=(double s) { x *= s; y *= s; z *= s; return *this; }
  V3& operator/=(double s) { x /= s; y /= s; z /= s; return *this; }
  bool operator==(const V3& o) const { return x == o.x && y == o.y && z == o.z; }
  double dot(const V3& o) const { return x * o.x + y * o.y + z * o.z; }
  V3 cross(const V3& o) const { return { y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x }; }
  double length2() const { return dot(*this); }
  double length() const { return std::sqrt(length2()); }
  bool fuzzyZero() const { return length2() < SIMD_EPS * SIMD_EPS; }
};
inline V3 operator*(double s, const V3& v) { return v * s; }
inline bool fuzzyZero(double x) { return std::fabs(x) < SIMD_EPS; }

struct M3
{
  double m[3][3] = { { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 } };
  V3 row(int i) const { return { m[i][0], m[i][1], m[i][2] }; }
  V3 col(int j) const { return { m[0][j], m[1][j], m[2][j] }; }
  V3 operator*(const V3& v) const { return { row(0).dot(v), row(1).dot(v), row(2).dot(v) }; }
  // Bullet's "v * basis" (row vector times matrix) == basis^T v
  V3 tmul(const V3& v) const { return { col(0).dot(v), col(1).dot(v), col(2).dot(v) }; }
  M3 operator*(const M3& o) const
  {
    M3 r;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        r.m[i][j] = m[i][0] * o.m[0][j] + m[i][1] * o.m[1][j] + m[i][2] * o.m[2][j];
    return r;
  }
  M3 transpose() const
  {
    M3 r;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        r.m[i][j] = m[j][i];
    return r;
  }
  M3 absolute() const
  {
    M3 r;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        r.m[i][j] = std::fabs(m[i][j]);
    return r;
  }
};

struct Xf
{
  M3 R;
  V3 p;
  V3 operator()(const V3& v) const { return R * v + p; }
  V3 operator*(const V3& v) const { return R * v + p; }
  Xf operator*(const Xf& o) const { return { R * o.R, R * o.p + p }; }
  Xf inverse() const
  {
    M3 rt = R.transpose();
    return { rt, -(rt * p) };
  }
  // this^-1 * o
  Xf inverseTimes(const Xf& o) const
  {
    M3 rt = R.transpose();
    return { rt * o.R, rt * (o.p - p) };
  }
  static Xf from12(const double* a)
  {
    Xf t;
    for (int i = 0; i < 3; ++i)
    {
      for (int j = 0; j < 3; ++j)
        t.R.m[i][j] = a[4 * i + j];
      t.p[i] = a[4 * i + 3];
    }
    return t;
  }
  void to12(double* a) const
  {
    for (int i = 0; i < 3; ++i)
    {
      for (int j = 0; j < 3; ++j)
        a[4 * i + j] = R.m[i][j];
      a[4 * i + 3] = p[i];
    }
  }
  bool isIdentity() const
  {
    for (int i = 0; i < 3; ++i)
    {
      for (int j = 0; j < 3; ++j)
        if (R.m[i][j] != (i == j ? 1.0 : 0.0))
          return false;
      if (p[i] != 0.0)
        return false;
    }
    return true;
  }
};

// ---------------------------------------------------------------------------------------------
// Convex shapes (bullet_utils.cpp:59-158 conversion rules; supports per Bullet btConvexShape.cpp)
// ---------------------------------------------------------------------------------------------
struct Convex
{
  int type = T4C_SPHERE;
  double r = 0, h = 0;  // radius; half length (cylinder, capsule) or full height (cone)
  V3 he;                // box half extents
  const V3* verts = nullptr;
  int nverts = 0;
  // cast hull (bullet_utils.h:217-304): support = argmax over the inner shape at pose 0 and pose t01
  const Convex* inner = nullptr;
  Xf t01;
  bool is_cast = false;

  double margin() const;  // getMargin(): sphere r, capsule r, everything else BULLET_MARGIN = 0
  V3 supportNoMargin(const V3& d) const;     // localGetSupportVertexWithoutMarginNonVirtual
  V3 supportWithMargin(const V3& d) const;   // localGetSupportVertexNonVirtual / localGetSupportingVertex
  void aabb(const Xf& t, V3& mn, V3& mx) const;  // btCollisionShape::getAabb (margin 0 shapes)
  bool polyhedral() const { return type == T4C_BOX || type == T4C_HULL || type == T4C_TRIANGLE; }
};

struct Counters
{
  uint64_t candidates = 0, gjk_pairs = 0, gjk_iterations = 0, epa_pairs = 0, epa_iterations = 0, contacts = 0;
};

struct PairRequest
{
  bool calculate_penetration = true;
  bool calculate_distance = true;
};

struct PairResult
{
  bool has = false;
  V3 normalOnB;  // world, points from B towards A
  V3 pointOnB;   // world
  double distance = 0;
};

// TesseractConvexConvexAlgorithm::processCollision + TesseractGjkPairDetector (+ Bullet's
// sphere-sphere / sphere-triangle algorithms chosen by the dispatch matrix).
// threshold = m_closestPointDistanceThreshold (max collision margin).
void collideConvexConvex(const Convex& a, const Xf& ta, const Convex& b, const Xf& tb, double threshold,
                         const PairRequest& req, PairResult& out, Counters& cnt);

// Exposed pieces for unit tests
struct GjkEpaResults
{
  int status = 0;  // 0 separated, 1 penetrating, 2 gjk failed, 3 epa failed
  V3 witnesses[2];
  V3 normal;
  double distance = 0;
  int epa_iterations = 0;
};
bool gjkEpaPenetration(const Convex& a, const Xf& ta, const Convex& b, const Xf& tb, const V3& guess,
                       GjkEpaResults& res, bool usemargins = true);
bool gjkEpaDistance(const Convex& a, const Xf& ta, const Convex& b, const Xf& tb, const V3& guess,
                    GjkEpaResults& res);

}  // namespace t4o

// -------------------------------------------------------------------------------------------------
// C API used by the Python tests (ctypes)
// -------------------------------------------------------------------------------------------------
extern "C" {
typedef struct t4o_contact
{
  int64_t state;
  int32_t link[2];
  int32_t shape_id[2];
  int32_t subshape_id[2];
  int32_t cc_type[2];
  double distance;
  double point[2][3];
  double point_local[2][3];
  double normal[3];
  double cc_time[2];
  double transform[2][12];
  double cc_transform[2][12];
} t4o_contact;

typedef struct t4o_manager t4o_manager;

// cast != 0 builds the BulletCastBVHManager restatement, else BulletDiscreteBVHManager
t4o_manager* t4o_manager_create(const t4c_scene_desc* desc, int cast);
t4o_manager* t4o_manager_clone(const t4o_manager* m);
void t4o_manager_destroy(t4o_manager* m);
void t4o_manager_set_active(t4o_manager* m, const int32_t* active_flags);
void t4o_manager_set_enabled(t4o_manager* m, int32_t link, int32_t enabled);
void t4o_manager_set_acm(t4o_manager* m, int32_t n_pairs, const int32_t* pairs);
void t4o_manager_set_margins(t4o_manager* m, double default_margin, int32_t n_pairs, const int32_t* pairs,
                             const double* values);
void t4o_manager_set_pose(t4o_manager* m, int32_t link, const double* pose12);
void t4o_manager_set_pose_cast(t4o_manager* m, int32_t link, const double* pose0, const double* pose1);
// contactTest: appends to out (up to capacity); returns total number of contacts produced
int64_t t4o_manager_contact_test(t4o_manager* m, const t4c_request* req, int64_t state_tag, t4o_contact* out,
                                 int64_t capacity);
void t4o_manager_get_counters(const t4o_manager* m, uint64_t* out6);

// FK (ofkt_nodes.cpp:200-289 / kdl_parser.h:86-118): out = n x n_links x 12
void t4o_fk(const t4c_scene_desc* scene, const t4c_fk_model_desc* model, const double* q, int64_t n, double* out);

// Batched drivers (checkTrajectoryState / checkTrajectorySegment loops), n_threads clones
int64_t t4o_check_discrete(const t4c_scene_desc* scene, const t4c_fk_model_desc* model, const double* q, int64_t n,
                           const t4c_request* req, uint8_t* hit, t4o_contact* out, int64_t capacity, int n_threads,
                           uint64_t* counters6);
int64_t t4o_check_cast(const t4c_scene_desc* scene, const t4c_fk_model_desc* model, const double* q0,
                       const double* q1, int64_t n, const t4c_request* req, uint8_t* hit, t4o_contact* out,
                       int64_t capacity, int n_threads, uint64_t* counters6);
// LVS expansion (core/utils.h:189-205, 328-347): returns number of work items; fills item arrays if non-null.
// cast=1: sub-segments (q0,q1); cast=0: sub-states (q0 only)
int64_t t4o_lvs_expand(const double* traj, int64_t n_traj, int32_t n_wp, int32_t n_q, double lvs, int cast,
                       double* q0, double* q1, int64_t* item_traj, int32_t* item_step, int64_t capacity);

// narrowphase probe for unit tests: one convex pair, shapes described by t4c_shape + vertex pool
int t4o_pair_distance(const t4c_shape* sa, const double* pose_a, const t4c_shape* sb, const double* pose_b,
                      const double* vertices, double threshold, int calc_pen, int calc_dist, double* out_dist,
                      double* out_normal_on_b, double* out_point_on_b);
}

#endif  // T4O_H_
