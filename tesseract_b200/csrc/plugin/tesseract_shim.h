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
