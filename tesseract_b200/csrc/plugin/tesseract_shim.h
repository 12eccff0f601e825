/*
 * tesseract_shim.h -- the part of the Tesseract / Eigen API surface the B200 contact managers touch.
 *
 * With -DT4_USE_TESSERACT_HEADERS the REAL headers are used (tesseract_collision, tesseract_geometry, Eigen): that is
 * the build a Tesseract workspace does.  This container has neither Eigen nor Boost nor Bullet, so
 * the default build declares signature-identical stand-ins in the same namespaces, restricted to what the managers
 * and their tests use.  Declarations follow:
 *   tesseract_collision/include/tesseract_collision/core/types.h:52-319
 *   tesseract_collision/include/tesseract_collision/core/discrete_contact_manager.h:44-208
 *   tesseract_collision/include/tesseract_collision/core/continuous_contact_manager.h:44-257
 *   tesseract_common/include/tesseract_common/collision_margin_data.h:40-270, types.h:60-75
 *   tesseract_geometry/include/tesseract_geometry/impl/{box,sphere,cylinder,cone,capsule,convex_mesh,mesh,octree}.h
 * Only the adapter functions at the bottom (pose <-> double[12], vertex access) differ between the two builds.
 */
#ifndef T4_TESSERACT_SHIM_H_
#define T4_TESSERACT_SHIM_H_

#ifdef T4_USE_TESSERACT_HEADERS
#include <tesseract_collision/core/continuous_contact_manager.h>
#include <tesseract_collision/core/discrete_contact_manager.h>
#include <tesseract_collision/core/types.h>
#include <tesseract_geometry/geometries.h>
#else

#include <array>
#include <cmath>
#include <functional>
#include <limits>
#include <map>
#include <memory>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#define EIGEN_MAKE_ALIGNED_OPERATOR_NEW

namespace Eigen
{
struct Vector3d
{
  double v[3]{ 0, 0, 0 };
  Vector3d() = default;
  Vector3d(double x, double y, double z) : v{ x, y, z } {}
  double& operator[](int i) { return v[i]; }
  const double& operator[](int i) const { return v[i]; }
  double& operator()(int i) { return v[i]; }
  const double& operator()(int i) const { return v[i]; }
  double x() const { return v[0]; }
  double y() const { return v[1]; }
  double z() const { return v[2]; }
  void setZero() { v[0] = v[1] = v[2] = 0; }
  Vector3d operator-(const Vector3d& o) const { return { v[0] - o.v[0], v[1] - o.v[1], v[2] - o.v[2] }; }
  Vector3d operator+(const Vector3d& o) const { return { v[0] + o.v[0], v[1] + o.v[1], v[2] + o.v[2] }; }
  double dot(const Vector3d& o) const { return v[0] * o.v[0] + v[1] * o.v[1] + v[2] * o.v[2]; }
  double norm() const { return std::sqrt(dot(*this)); }
};

// rigid transform, rows of [R | p]
struct Isometry3d
{
  double m[12]{ 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0 };
  static Isometry3d Identity() { return Isometry3d(); }
  double& operator()(int r, int c) { return m[4 * r + c]; }
  const double& operator()(int r, int c) const { return m[4 * r + c]; }
  Vector3d translation() const { return { m[3], m[7], m[11] }; }
  void setTranslation(double x, double y, double z) { m[3] = x, m[7] = y, m[11] = z; }
  Isometry3d operator*(const Isometry3d& o) const
  {
    Isometry3d r;
    for (int i = 0; i < 3; ++i)
    {
      for (int j = 0; j < 3; ++j)
        r.m[4 * i + j] = m[4 * i] * o.m[j] + m[4 * i + 1] * o.m[4 + j] + m[4 * i + 2] * o.m[8 + j];
      r.m[4 * i + 3] = m[4 * i] * o.m[3] + m[4 * i + 1] * o.m[7] + m[4 * i + 2] * o.m[11] + m[4 * i + 3];
    }
    return r;
  }
  Vector3d operator*(const Vector3d& p) const
  {
    return { m[0] * p[0] + m[1] * p[1] + m[2] * p[2] + m[3], m[4] * p[0] + m[5] * p[1] + m[6] * p[2] + m[7],
             m[8] * p[0] + m[9] * p[1] + m[10] * p[2] + m[11] };
  }
  Isometry3d inverse() const
  {
    Isometry3d r;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        r.m[4 * i + j] = m[4 * j + i];
    for (int i = 0; i < 3; ++i)
      r.m[4 * i + 3] = -(r.m[4 * i] * m[3] + r.m[4 * i + 1] * m[7] + r.m[4 * i + 2] * m[11]);
    return r;
  }
};
using VectorXi = std::vector<int>;
}  // namespace Eigen

namespace tesseract_common
{
template <typename T>
using AlignedVector = std::vector<T>;
template <typename K, typename V>
using AlignedMap = std::map<K, V>;
using VectorIsometry3d = AlignedVector<Eigen::Isometry3d>;
using VectorVector3d = std::vector<Eigen::Vector3d>;
using TransformMap = AlignedMap<std::string, Eigen::Isometry3d>;
// Eigen::Matrix<double, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor> (tesseract_common/types.h:52-53):
// one joint vector per row, contiguous row-major storage
class TrajArray
{
public:
  TrajArray() = default;
  TrajArray(long rows, long cols) : rows_(rows), cols_(cols), v_((size_t)(rows * cols), 0.0) {}
  long rows() const { return rows_; }
  long cols() const { return cols_; }
  double& operator()(long r, long c) { return v_[(size_t)(r * cols_ + c)]; }
  double operator()(long r, long c) const { return v_[(size_t)(r * cols_ + c)]; }
  const double* data() const { return v_.data(); }
  double* data() { return v_.data(); }

private:
  long rows_ = 0, cols_ = 0;
  std::vector<double> v_;
};
using LinkNamesPair = std::pair<std::string, std::string>;

struct PairHash
{
  std::size_t operator()(const LinkNamesPair& p) const { return std::hash<std::string>()(p.first + p.second); }
};

inline LinkNamesPair makeOrderedLinkPair(const std::string& a, const std::string& b)
{
  return a <= b ? std::make_pair(a, b) : std::make_pair(b, a);
}

enum class CollisionMarginOverrideType
{
  NONE,
  REPLACE,
  OVERRIDE_DEFAULT_MARGIN,
  OVERRIDE_PAIR_MARGIN,
  MODIFY_PAIR_MARGIN
};
using PairsCollisionMarginData = std::unordered_map<LinkNamesPair, double, PairHash>;

// collision_margin_data.h:66-270
class CollisionMarginData
{
public:
  CollisionMarginData(double default_collision_margin = 0)
    : default_collision_margin_(default_collision_margin), max_collision_margin_(default_collision_margin)
  {
  }
  CollisionMarginData(double default_collision_margin, PairsCollisionMarginData pairs)
    : default_collision_margin_(default_collision_margin), lookup_table_(std::move(pairs))
  {
    updateMax();
  }
  void setDefaultCollisionMargin(double m)
  {
    default_collision_margin_ = m;
    updateMax();
  }
  double getDefaultCollisionMargin() const { return default_collision_margin_; }
  void setPairCollisionMargin(const std::string& a, const std::string& b, double m)
  {
    lookup_table_[makeOrderedLinkPair(a, b)] = m;
    updateMax();
  }
  double getPairCollisionMargin(const std::string& a, const std::string& b) const
  {
    const auto it = lookup_table_.find(makeOrderedLinkPair(a, b));
    return it != lookup_table_.end() ? it->second : default_collision_margin_;
  }
  const PairsCollisionMarginData& getPairCollisionMargins() const { return lookup_table_; }
  double getMaxCollisionMargin() const { return max_collision_margin_; }
  // collision_margin_data.h:163-181: every margin moves, so the maximum moves with them
  void incrementMargins(const double& increment)
  {
    default_collision_margin_ += increment;
    max_collision_margin_ += increment;
    for (auto& p : lookup_table_)
      p.second += increment;
  }
  void scaleMargins(const double& scale)
  {
    default_collision_margin_ *= scale;
    max_collision_margin_ *= scale;
    for (auto& p : lookup_table_)
      p.second *= scale;
  }
  void apply(const CollisionMarginData& d, CollisionMarginOverrideType t)
  {
    switch (t)
    {
      case CollisionMarginOverrideType::REPLACE:
        *this = d;
        break;
      case CollisionMarginOverrideType::OVERRIDE_DEFAULT_MARGIN:
        default_collision_margin_ = d.default_collision_margin_;
        updateMax();
        break;
      case CollisionMarginOverrideType::OVERRIDE_PAIR_MARGIN:
        lookup_table_ = d.lookup_table_;
        updateMax();
        break;
      case CollisionMarginOverrideType::MODIFY_PAIR_MARGIN:
        for (const auto& p : d.lookup_table_)
          lookup_table_[p.first] = p.second;
        updateMax();
        break;
      case CollisionMarginOverrideType::NONE:
        break;
    }
  }

private:
  double default_collision_margin_{ 0 };
  double max_collision_margin_{ 0 };
  PairsCollisionMarginData lookup_table_;
  void updateMax()
  {
    max_collision_margin_ = default_collision_margin_;
    for (const auto& p : lookup_table_)
      if (p.second > max_collision_margin_)
        max_collision_margin_ = p.second;
  }
};
}  // namespace tesseract_common

namespace tesseract_geometry
{
enum GeometryType
{
  SPHERE,
  CYLINDER,
  CAPSULE,
  CONE,
  BOX,
  PLANE,
  MESH,
  CONVEX_MESH,
  SDF_MESH,
  OCTREE
};

class Geometry
{
public:
  using Ptr = std::shared_ptr<Geometry>;
  using ConstPtr = std::shared_ptr<const Geometry>;
  explicit Geometry(GeometryType type) : type_(type) {}
  virtual ~Geometry() = default;
  Geometry(const Geometry&) = delete;
  Geometry& operator=(const Geometry&) = delete;
  GeometryType getType() const { return type_; }

private:
  GeometryType type_;
};

class Box : public Geometry
{
public:
  Box(double x, double y, double z) : Geometry(BOX), x_(x), y_(y), z_(z) {}
  double getX() const { return x_; }
  double getY() const { return y_; }
  double getZ() const { return z_; }

private:
  double x_, y_, z_;
};
class Sphere : public Geometry
{
public:
  explicit Sphere(double r) : Geometry(SPHERE), r_(r) {}
  double getRadius() const { return r_; }

private:
  double r_;
};
#define T4_SHIM_RL(NAME, TYPE)                                                                                         \
  class NAME : public Geometry                                                                                         \
  {                                                                                                                    \
  public:                                                                                                              \
    NAME(double r, double l) : Geometry(TYPE), r_(r), l_(l) {}                                                         \
    double getRadius() const { return r_; }                                                                            \
    double getLength() const { return l_; }                                                                            \
                                                                                                                       \
  private:                                                                                                             \
    double r_, l_;                                                                                                     \
  }
T4_SHIM_RL(Cylinder, CYLINDER);
T4_SHIM_RL(Cone, CONE);
T4_SHIM_RL(Capsule, CAPSULE);
#undef T4_SHIM_RL

// mesh.h:148-166 / convex_mesh.h:146-164: triangles / faces are "n, i0, .., in-1" runs in one index vector
class Mesh : public Geometry
{
public:
  Mesh(std::shared_ptr<const tesseract_common::VectorVector3d> vertices, std::shared_ptr<const Eigen::VectorXi> triangles,
       GeometryType type = MESH)
    : Geometry(type), vertices_(std::move(vertices)), triangles_(std::move(triangles))
  {
    vertice_count_ = (int)vertices_->size();
    triangle_count_ = 0;
    for (size_t i = 0; i < triangles_->size(); i += (size_t)(*triangles_)[i] + 1)
      ++triangle_count_;
  }
  const std::shared_ptr<const tesseract_common::VectorVector3d>& getVertices() const { return vertices_; }
  const std::shared_ptr<const Eigen::VectorXi>& getTriangles() const { return triangles_; }
  int getVerticeCount() const { return vertice_count_; }
  int getTriangleCount() const { return triangle_count_; }

private:
  std::shared_ptr<const tesseract_common::VectorVector3d> vertices_;
  std::shared_ptr<const Eigen::VectorXi> triangles_;
  int vertice_count_, triangle_count_;
};
class ConvexMesh : public Geometry
{
public:
  ConvexMesh(std::shared_ptr<const tesseract_common::VectorVector3d> vertices,
             std::shared_ptr<const Eigen::VectorXi> faces = nullptr)
    : Geometry(CONVEX_MESH), vertices_(std::move(vertices)), faces_(std::move(faces))
  {
  }
  const std::shared_ptr<const tesseract_common::VectorVector3d>& getVertices() const { return vertices_; }
  const std::shared_ptr<const Eigen::VectorXi>& getFaces() const { return faces_; }
  int getVerticeCount() const { return (int)vertices_->size(); }

private:
  std::shared_ptr<const tesseract_common::VectorVector3d> vertices_;
  std::shared_ptr<const Eigen::VectorXi> faces_;
};

// octree.h:47-120 wraps an octomap::OcTree; the stand-in carries the occupied max-depth leaves the reference
// iterates (bullet_utils.cpp:164-172): centre xyz + edge length, in leaf-iterator order.
class Octree : public Geometry
{
public:
  enum SubType
  {
    BOX,
    SPHERE_INSIDE,
    SPHERE_OUTSIDE
  };
  struct Leaf
  {
    double x, y, z, size;
  };
  Octree(std::vector<Leaf> leaves, SubType sub_type) : Geometry(OCTREE), leaves_(std::move(leaves)), sub_type_(sub_type)
  {
  }
  SubType getSubType() const { return sub_type_; }
  const std::vector<Leaf>& getLeaves() const { return leaves_; }

private:
  std::vector<Leaf> leaves_;
  SubType sub_type_;
};
}  // namespace tesseract_geometry

namespace tesseract_collision
{
using CollisionShapesConst = std::vector<tesseract_geometry::Geometry::ConstPtr>;
using CollisionShapeConstPtr = tesseract_geometry::Geometry::ConstPtr;
using CollisionMarginData = tesseract_common::CollisionMarginData;
using CollisionMarginOverrideType = tesseract_common::CollisionMarginOverrideType;
using PairsCollisionMarginData = tesseract_common::PairsCollisionMarginData;
using IsContactAllowedFn = std::function<bool(const std::string&, const std::string&)>;

enum class ContinuousCollisionType
{
  CCType_None,
  CCType_Time0,
  CCType_Time1,
  CCType_Between
};
enum class ContactTestType
{
  FIRST = 0,
  CLOSEST = 1,
  ALL = 2,
  LIMITED = 3
};

// types.h:89-161
struct ContactResult
{
  double distance;
  std::array<int, 2> type_id;
  std::array<std::string, 2> link_names;
  std::array<int, 2> shape_id;
  std::array<int, 2> subshape_id;
  std::array<Eigen::Vector3d, 2> nearest_points;
  std::array<Eigen::Vector3d, 2> nearest_points_local;
  std::array<Eigen::Isometry3d, 2> transform;
  Eigen::Vector3d normal;
  std::array<double, 2> cc_time;
  std::array<ContinuousCollisionType, 2> cc_type;
  std::array<Eigen::Isometry3d, 2> cc_transform;
  bool single_contact_point = false;
  ContactResult() { clear(); }
  void clear()
  {
    distance = std::numeric_limits<double>::max();
    for (int i = 0; i < 2; ++i)
    {
      nearest_points[i].setZero();
      nearest_points_local[i].setZero();
      transform[i] = Eigen::Isometry3d::Identity();
      cc_transform[i] = Eigen::Isometry3d::Identity();
      link_names[i] = "";
      shape_id[i] = -1;
      subshape_id[i] = -1;
      type_id[i] = 0;
      cc_time[i] = -1;
      cc_type[i] = ContinuousCollisionType::CCType_None;
    }
    normal.setZero();
    single_contact_point = false;
  }
};
using ContactResultVector = tesseract_common::AlignedVector<ContactResult>;
using ContactResultMap = tesseract_common::AlignedMap<std::pair<std::string, std::string>, ContactResultVector>;
using IsContactResultValidFn = std::function<bool(const ContactResult&)>;

struct ContactRequest
{
  ContactTestType type = ContactTestType::ALL;
  bool calculate_penetration = true;
  bool calculate_distance = true;
  long contact_limit = 0;
  IsContactResultValidFn is_valid = nullptr;
  ContactRequest(ContactTestType type = ContactTestType::ALL) : type(type) {}
};

// types.h:278-319
enum class CollisionEvaluatorType
{
  DISCRETE,
  LVS_DISCRETE,
  CONTINUOUS,
  LVS_CONTINUOUS
};
struct CollisionCheckConfig
{
  CollisionCheckConfig(double default_margin = 0, ContactRequest request = ContactRequest(),
                       CollisionEvaluatorType type = CollisionEvaluatorType::DISCRETE,
                       double longest_valid_segment_length = 0.005)
    : collision_margin_data(default_margin)
    , contact_request(std::move(request))
    , type(type)
    , longest_valid_segment_length(longest_valid_segment_length)
  {
  }
  CollisionMarginOverrideType collision_margin_override_type{ CollisionMarginOverrideType::NONE };
  CollisionMarginData collision_margin_data;
  ContactRequest contact_request;
  CollisionEvaluatorType type;
  double longest_valid_segment_length{ 0.005 };
};

// discrete_contact_manager.h:65-207
class DiscreteContactManager
{
public:
  using Ptr = std::shared_ptr<DiscreteContactManager>;
  using ConstPtr = std::shared_ptr<const DiscreteContactManager>;
  DiscreteContactManager() = default;
  virtual ~DiscreteContactManager() = default;
  DiscreteContactManager(const DiscreteContactManager&) = delete;
  DiscreteContactManager& operator=(const DiscreteContactManager&) = delete;
  virtual std::shared_ptr<DiscreteContactManager> clone() const = 0;
  virtual bool addCollisionObject(const std::string& name, const int& mask_id, const CollisionShapesConst& shapes,
                                  const tesseract_common::VectorIsometry3d& shape_poses, bool enabled = true) = 0;
  virtual const CollisionShapesConst& getCollisionObjectGeometries(const std::string& name) const = 0;
  virtual const tesseract_common::VectorIsometry3d&
  getCollisionObjectGeometriesTransforms(const std::string& name) const = 0;
  virtual bool hasCollisionObject(const std::string& name) const = 0;
  virtual bool removeCollisionObject(const std::string& name) = 0;
  virtual bool enableCollisionObject(const std::string& name) = 0;
  virtual bool disableCollisionObject(const std::string& name) = 0;
  virtual void setCollisionObjectsTransform(const std::string& name, const Eigen::Isometry3d& pose) = 0;
  virtual void setCollisionObjectsTransform(const std::vector<std::string>& names,
                                            const tesseract_common::VectorIsometry3d& poses) = 0;
  virtual void setCollisionObjectsTransform(const tesseract_common::TransformMap& transforms) = 0;
  virtual const std::vector<std::string>& getCollisionObjects() const = 0;
  virtual void setActiveCollisionObjects(const std::vector<std::string>& names) = 0;
  virtual const std::vector<std::string>& getActiveCollisionObjects() const = 0;
  virtual void
  setCollisionMarginData(CollisionMarginData collision_margin_data,
                         CollisionMarginOverrideType override_type = CollisionMarginOverrideType::REPLACE) = 0;
  virtual void setDefaultCollisionMarginData(double default_collision_margin) = 0;
  virtual void setPairCollisionMarginData(const std::string& name1, const std::string& name2,
                                          double collision_margin) = 0;
  virtual const CollisionMarginData& getCollisionMarginData() const = 0;
  virtual void setIsContactAllowedFn(IsContactAllowedFn fn) = 0;
  virtual IsContactAllowedFn getIsContactAllowedFn() const = 0;
  virtual void contactTest(ContactResultMap& collisions, const ContactRequest& request) = 0;
};
}  // namespace tesseract_collision

namespace tesseract_collision
{
// continuous_contact_manager.h:65-256
class ContinuousContactManager
{
public:
  using Ptr = std::shared_ptr<ContinuousContactManager>;
  using ConstPtr = std::shared_ptr<const ContinuousContactManager>;
  ContinuousContactManager() = default;
  virtual ~ContinuousContactManager() = default;
  ContinuousContactManager(const ContinuousContactManager&) = delete;
  ContinuousContactManager& operator=(const ContinuousContactManager&) = delete;
  virtual std::shared_ptr<ContinuousContactManager> clone() const = 0;
  virtual bool addCollisionObject(const std::string& name, const int& mask_id, const CollisionShapesConst& shapes,
                                  const tesseract_common::VectorIsometry3d& shape_poses, bool enabled = true) = 0;
  virtual const CollisionShapesConst& getCollisionObjectGeometries(const std::string& name) const = 0;
  virtual const tesseract_common::VectorIsometry3d&
  getCollisionObjectGeometriesTransforms(const std::string& name) const = 0;
  virtual bool hasCollisionObject(const std::string& name) const = 0;
  virtual bool removeCollisionObject(const std::string& name) = 0;
  virtual bool enableCollisionObject(const std::string& name) = 0;
  virtual bool disableCollisionObject(const std::string& name) = 0;
  virtual void setCollisionObjectsTransform(const std::string& name, const Eigen::Isometry3d& pose) = 0;
  virtual void setCollisionObjectsTransform(const std::vector<std::string>& names,
                                            const tesseract_common::VectorIsometry3d& poses) = 0;
  virtual void setCollisionObjectsTransform(const tesseract_common::TransformMap& transforms) = 0;
  virtual void setCollisionObjectsTransform(const std::string& name, const Eigen::Isometry3d& pose1,
                                            const Eigen::Isometry3d& pose2) = 0;
  virtual void setCollisionObjectsTransform(const std::vector<std::string>& names,
                                            const tesseract_common::VectorIsometry3d& pose1,
                                            const tesseract_common::VectorIsometry3d& pose2) = 0;
  virtual void setCollisionObjectsTransform(const tesseract_common::TransformMap& pose1,
                                            const tesseract_common::TransformMap& pose2) = 0;
  virtual const std::vector<std::string>& getCollisionObjects() const = 0;
  virtual void setActiveCollisionObjects(const std::vector<std::string>& names) = 0;
  virtual const std::vector<std::string>& getActiveCollisionObjects() const = 0;
  virtual void
  setCollisionMarginData(CollisionMarginData collision_margin_data,
                         CollisionMarginOverrideType override_type = CollisionMarginOverrideType::REPLACE) = 0;
  virtual void setDefaultCollisionMarginData(double default_collision_margin) = 0;
  virtual void setPairCollisionMarginData(const std::string& name1, const std::string& name2,
                                          double collision_margin) = 0;
  virtual const CollisionMarginData& getCollisionMarginData() const = 0;
  virtual void setIsContactAllowedFn(IsContactAllowedFn fn) = 0;
  virtual IsContactAllowedFn getIsContactAllowedFn() const = 0;
  virtual void contactTest(ContactResultMap& collisions, const ContactRequest& request) = 0;
};
}  // namespace tesseract_collision
#endif  // !T4_USE_TESSERACT_HEADERS

// ---------------------------------------------------------------------------------------------------------------
// adapters: the only code that depends on which of the two header sets is in use
// ---------------------------------------------------------------------------------------------------------------
namespace tesseract_b200
{
inline void toPose12(const Eigen::Isometry3d& T, double* p)
{
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c)
      p[4 * r + c] = T(r, c);
}
inline Eigen::Isometry3d fromPose12(const double* p)
{
  Eigen::Isometry3d T = Eigen::Isometry3d::Identity();
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c)
      T(r, c) = p[4 * r + c];
  return T;
}

struct OctreeLeaf
{
  double x, y, z, size;
};
// occupied leaves at maximum depth, leaf-iterator order (bullet_utils.cpp:164-172)
inline void octreeLeaves(const tesseract_geometry::Octree& g, std::vector<OctreeLeaf>& out)
{
#ifdef T4_USE_TESSERACT_HEADERS
  const octomap::OcTree& octree = *(g.getOctree());
  for (auto it = octree.begin(static_cast<unsigned char>(octree.getTreeDepth())), end = octree.end(); it != end; ++it)
    if (octree.isNodeOccupied(*it))
      out.push_back({ it.getX(), it.getY(), it.getZ(), it.getSize() });
#else
  for (const auto& l : g.getLeaves())
    out.push_back({ l.x, l.y, l.z, l.size });
#endif
}
}  // namespace tesseract_b200
#endif  // T4_TESSERACT_SHIM_H_
