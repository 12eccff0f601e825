/*
 * b200_contact_managers.cpp -- see b200_contact_managers.h.  Host logic only: every query goes through the C ABI
 * (libt4c.so, CUDA); a failing C-ABI call throws std::runtime_error with t4c_last_error().
 */
#include "b200_contact_managers.h"

#include <algorithm>
#include <cstring>
#include <stdexcept>

namespace tesseract_b200
{
using namespace tesseract_collision;

static const CollisionShapesConst EMPTY_SHAPES;
static const tesseract_common::VectorIsometry3d EMPTY_POSES;

static void check(t4c_status st, const char* what)
{
  if (st != T4C_OK)
    throw std::runtime_error(std::string("tesseract_b200: ") + what + ": " + t4c_last_error());
}

ManagerCore::ManagerCore(bool cast, int device) : cast_(cast), device_(device), margins_(0) {}

ManagerCore::~ManagerCore()
{
  if (scene_)
    t4c_scene_destroy(scene_);
}

// clone(): identical objects, transforms, active set, margins, ACM function; geometries stay shared
// (bullet_discrete_bvh_manager.cpp:81-105).  Every consumer of an Environment takes a clone
// (tesseract_environment/src/core/environment.cpp:482-488), so a clone must be cheap: when this manager's device scene
// is built, the clone takes a reference to the same compiled image (t4c_scene_clone: no compile, no upload); otherwise
// its scene is built on its first query.
std::unique_ptr<ManagerCore> ManagerCore::clone(int device) const
{
  const bool same = device < 0 || device == device_;
  std::unique_ptr<ManagerCore> c(new ManagerCore(cast_, same ? device_ : device));
  c->objects_ = objects_;
  c->names_ = names_;
  c->index_ = index_;
  c->active_ = active_;
  c->margins_ = margins_;
  c->fn_ = fn_;
  if (same && scene_ && !dirty_structure_ && t4c_scene_clone(scene_, &c->scene_) == T4C_OK)
  {
    c->dirty_structure_ = false;
    c->dirty_soft_ = true;  // settings are compared by the library: unchanged ones cost nothing
    c->acm_uploaded_ = acm_uploaded_;
    c->acm_valid_ = acm_valid_ && !dirty_soft_;
  }
  return c;
}

int ManagerCore::find(const std::string& name) const
{
  const auto it = index_.find(name);
  return it == index_.end() ? -1 : it->second;
}

bool ManagerCore::isActive(int i) const
{
  // isLinkActive: an empty list means every link is active (core/common.h:66-69)
  return active_.empty() || std::find(active_.begin(), active_.end(), objects_[i].name) != active_.end();
}

bool ManagerCore::addCollisionObject(const std::string& name, int mask_id, const CollisionShapesConst& shapes,
                                     const tesseract_common::VectorIsometry3d& shape_poses, bool enabled)
{
  if (find(name) >= 0)
    removeCollisionObject(name);  // re-adding replaces (bullet_discrete_bvh_manager.cpp:113-114)
  if (shapes.empty() || shapes.size() != shape_poses.size())
    return false;  // bullet_utils.h:926-930
  for (const auto& g : shapes)
    if (!g)
      return false;
  Object o;
  o.name = name;
  o.mask_id = mask_id;
  o.shapes = shapes;
  o.shape_poses = shape_poses;
  o.enabled = enabled;
  index_[name] = (int)objects_.size();
  objects_.push_back(std::move(o));
  names_.push_back(name);
  dirty_structure_ = true;
  return true;
}

const CollisionShapesConst& ManagerCore::getGeometries(const std::string& name) const
{
  const int i = find(name);
  return i >= 0 ? objects_[i].shapes : EMPTY_SHAPES;  // static empty vector (bullet_discrete_bvh_manager.cpp:50-51)
}

const tesseract_common::VectorIsometry3d& ManagerCore::getGeometriesTransforms(const std::string& name) const
{
  const int i = find(name);
  return i >= 0 ? objects_[i].shape_poses : EMPTY_POSES;
}

bool ManagerCore::hasCollisionObject(const std::string& name) const { return find(name) >= 0; }

bool ManagerCore::removeCollisionObject(const std::string& name)
{
  const int i = find(name);
  if (i < 0)
    return false;
  objects_.erase(objects_.begin() + i);
  names_.erase(names_.begin() + i);
  index_.clear();
  for (size_t k = 0; k < objects_.size(); ++k)
    index_[objects_[k].name] = (int)k;
  dirty_structure_ = true;
  return true;
}

bool ManagerCore::setEnabled(const std::string& name, bool enabled)
{
  const int i = find(name);
  if (i < 0)
    return false;
  if (objects_[i].enabled != enabled)
  {
    objects_[i].enabled = enabled;
    dirty_soft_ = true;
  }
  return true;
}

void ManagerCore::setTransform(const std::string& name, const Eigen::Isometry3d& pose)
{
  const int i = find(name);
  if (i < 0)
    return;  // unknown name: silent no-op (bullet_discrete_bvh_manager.cpp:198-199)
  objects_[i].pose0 = pose;
  objects_[i].pose1 = pose;
  if (!isActive(i))
    dirty_soft_ = true;  // static link: its world pose is part of the device scene
}

void ManagerCore::setTransform(const std::string& name, const Eigen::Isometry3d& pose1, const Eigen::Isometry3d& pose2)
{
  const int i = find(name);
  if (i < 0)
    return;
  objects_[i].pose0 = pose1;
  objects_[i].pose1 = pose2;
  if (!isActive(i))
    dirty_soft_ = true;
}

void ManagerCore::setActive(const std::vector<std::string>& names)
{
  active_ = names;
  dirty_soft_ = true;
}

void ManagerCore::setMarginData(CollisionMarginData data, CollisionMarginOverrideType type)
{
  margins_.apply(data, type);
  dirty_soft_ = true;
}
void ManagerCore::setDefaultMargin(double m)
{
  margins_.setDefaultCollisionMargin(m);
  dirty_soft_ = true;
}
void ManagerCore::setPairMargin(const std::string& a, const std::string& b, double m)
{
  margins_.setPairCollisionMargin(a, b, m);
  dirty_soft_ = true;
}

// Geometry -> convex shape records: the reference's createShapePrimitive family
// (tesseract_collision/src/bullet/bullet_utils.cpp:59-365); see include/t4c.h for the parameter conventions.
static void appendShapes(const tesseract_geometry::Geometry& g, int link, int shape_id, int plain_sub,
                         const Eigen::Isometry3d& lpose, std::vector<t4c_shape>& shapes, std::vector<double>& verts)
{
  using namespace tesseract_geometry;
  t4c_shape s;
  std::memset(&s, 0, sizeof(s));
  s.link = link;
  s.shape_id = shape_id;
  s.subshape_id = plain_sub;
  toPose12(lpose, s.pose);
  switch (g.getType())
  {
    case BOX:
    {
      const auto& b = static_cast<const Box&>(g);
      s.type = T4C_BOX;
      s.params[0] = b.getX() / 2, s.params[1] = b.getY() / 2, s.params[2] = b.getZ() / 2;
      shapes.push_back(s);
      return;
    }
    case SPHERE:
      s.type = T4C_SPHERE;
      s.params[0] = static_cast<const Sphere&>(g).getRadius();
      shapes.push_back(s);
      return;
    case CYLINDER:
      s.type = T4C_CYLINDER;
      s.params[0] = static_cast<const Cylinder&>(g).getRadius();
      s.params[1] = static_cast<const Cylinder&>(g).getLength() / 2;
      shapes.push_back(s);
      return;
    case CONE:
      s.type = T4C_CONE;
      s.params[0] = static_cast<const Cone&>(g).getRadius();
      s.params[1] = static_cast<const Cone&>(g).getLength();
      shapes.push_back(s);
      return;
    case CAPSULE:
      s.type = T4C_CAPSULE;
      s.params[0] = static_cast<const Capsule&>(g).getRadius();
      s.params[1] = static_cast<const Capsule&>(g).getLength() / 2;
      shapes.push_back(s);
      return;
    case CONVEX_MESH:
    {
      const auto& m = static_cast<const ConvexMesh&>(g);
      s.type = T4C_HULL;
      s.vertex_offset = (int32_t)(verts.size() / 3);
      s.vertex_count = m.getVerticeCount();
      for (const auto& v : *m.getVertices())
        verts.insert(verts.end(), { v[0], v[1], v[2] });
      shapes.push_back(s);
      return;
    }
    case MESH:
    {
      // one triangle shape per face, subshape_id = face index (bullet_utils.cpp:111-132)
      const auto& m = static_cast<const Mesh&>(g);
      const auto& V = *m.getVertices();
      const auto& T = *m.getTriangles();
      int face = 0;
      for (long i = 0; i < (long)T.size(); i += T[i] + 1, ++face)
      {
        s.type = T4C_TRIANGLE;
        s.subshape_id = face;
        s.vertex_offset = (int32_t)(verts.size() / 3);
        s.vertex_count = 3;
        for (int k = 1; k <= 3; ++k)
        {
          const auto& v = V[(size_t)T[i + k]];
          verts.insert(verts.end(), { v[0], v[1], v[2] });
        }
        shapes.push_back(s);
      }
      return;
    }
    case OCTREE:
    {
      // one box / inscribed sphere / circumscribed sphere per occupied leaf (bullet_utils.cpp:160-243)
      const auto& o = static_cast<const Octree&>(g);
      std::vector<OctreeLeaf> leaves;
      octreeLeaves(o, leaves);
      int v = 0;
      for (const auto& l : leaves)
      {
        Eigen::Isometry3d t = Eigen::Isometry3d::Identity();
        t(0, 3) = l.x, t(1, 3) = l.y, t(2, 3) = l.z;
        toPose12(lpose * t, s.pose);
        s.subshape_id = v++;
        const double h = l.size / 2;
        if (o.getSubType() == Octree::BOX)
        {
          s.type = T4C_BOX;
          s.params[0] = s.params[1] = s.params[2] = h;
        }
        else
        {
          s.type = T4C_SPHERE;
          s.params[0] = o.getSubType() == Octree::SPHERE_INSIDE ? h : std::sqrt(2 * h * h);
          s.params[1] = s.params[2] = 0;
        }
        shapes.push_back(s);
      }
      return;
    }
    default:
      throw std::runtime_error("tesseract_b200: unsupported geometry type (PLANE / SDF_MESH are not supported by the "
                               "reference's Bullet managers either)");
  }
}

static bool isIdentity(const Eigen::Isometry3d& T)
{
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c)
      if (T(r, c) != (r == c ? 1.0 : 0.0))
        return false;
  return true;
}

void ManagerCore::sync()
{
  const int nl = (int)objects_.size();
  if (dirty_structure_ || !scene_)
  {
    if (scene_)
      t4c_scene_destroy(scene_);
    scene_ = nullptr;
    std::vector<t4c_link> links((size_t)nl);
    std::vector<t4c_shape> shapes;
    std::vector<double> verts;
    for (int i = 0; i < nl; ++i)
    {
      const Object& o = objects_[i];
      links[i].name = o.name.c_str();
      links[i].type_id = o.mask_id;
      links[i].enabled = o.enabled ? 1 : 0;
      links[i].active = isActive(i) ? 1 : 0;
      toPose12(o.pose0, links[i].pose);
      // a single shape with identity pose is used directly, anything else is a compound whose child index becomes
      // subshape_id (bullet_utils.cpp:334-360, bullet_utils.h:422-423)
      const bool single = o.shapes.size() == 1 && isIdentity(o.shape_poses[0]);
      for (size_t j = 0; j < o.shapes.size(); ++j)
        appendShapes(*o.shapes[j], i, (int)j, single ? -1 : (int)j, o.shape_poses[j], shapes, verts);
    }
    t4c_scene_desc d;
    std::memset(&d, 0, sizeof(d));
    d.n_links = nl;
    d.links = links.data();
    d.n_shapes = (int32_t)shapes.size();
    d.shapes = shapes.data();
    d.n_vertices = (int32_t)(verts.size() / 3);
    d.vertices = verts.data();
    d.default_margin = margins_.getDefaultCollisionMargin();
    check(t4c_scene_create(&d, device_, &scene_), "t4c_scene_create");
    dirty_structure_ = false;
    dirty_soft_ = true;
    acm_valid_ = false;
  }
  if (dirty_soft_)
  {
    std::vector<int32_t> act((size_t)std::max(nl, 1), 0);
    std::vector<int32_t> moved;
    std::vector<double> moved_pose;
    for (int i = 0; i < nl; ++i)
    {
      act[i] = isActive(i) ? 1 : 0;
      check(t4c_scene_set_enabled(scene_, i, objects_[i].enabled ? 1 : 0), "t4c_scene_set_enabled");
      if (!act[i])
      {
        moved.push_back(i);
        moved_pose.resize(moved_pose.size() + 12);
        toPose12(objects_[i].pose0, &moved_pose[moved_pose.size() - 12]);
      }
    }
    if (nl > 0)
      check(t4c_scene_set_active(scene_, act.data()), "t4c_scene_set_active");
    // static objects that actually moved are refitted into the environment BVH in ONE pass (unchanged poses are
    // recognised by the library and cost nothing)
    if (!moved.empty())
      check(t4c_scene_set_link_poses(scene_, (int32_t)moved.size(), moved.data(), moved_pose.data()),
            "t4c_scene_set_link_poses");
    std::vector<int32_t> mp;
    std::vector<double> mv;
    for (const auto& kv : margins_.getPairCollisionMargins())
    {
      const int a = find(kv.first.first), b = find(kv.first.second);
      if (a < 0 || b < 0)
        continue;
      mp.push_back(a), mp.push_back(b);
      mv.push_back(kv.second);
    }
    check(t4c_scene_set_margins(scene_, margins_.getDefaultCollisionMargin(), (int32_t)mv.size(), mp.data(), mv.data()),
          "t4c_scene_set_margins");
    dirty_soft_ = false;
    acm_valid_ = false;
  }
  // IsContactAllowedFn is an opaque std::function that may change behaviour without notice (bullet_utils.h:871-872):
  // it is evaluated into the pair list at every query, for (active link) x (every other link) only -- static x static
  // pairs are never tested (bullet_utils.h:900-917).  The device bit-matrix is re-uploaded only when the list changed.
  std::vector<int32_t> acm;
  if (fn_)
    for (int i = 0; i < nl; ++i)
    {
      const bool ai = isActive(i);
      for (int j = i + 1; j < nl; ++j)
        if ((ai || isActive(j)) && fn_(objects_[i].name, objects_[j].name))
          acm.push_back(i), acm.push_back(j);
    }
  if (!acm_valid_ || acm != acm_uploaded_)
  {
    check(t4c_scene_set_acm(scene_, (int32_t)(acm.size() / 2), acm.data()), "t4c_scene_set_acm");
    acm_uploaded_.swap(acm);
    acm_valid_ = true;
  }
}

t4c_scene* ManagerCore::scene()
{
  sync();
  return scene_;
}

void ManagerCore::run(const double* p0, const double* p1, int64_t n, const ContactRequest& request, int device_type,
                      std::vector<uint8_t>* hit, std::vector<t4c_contact>& contacts, std::vector<float>& local)
{
  sync();
  t4c_request rq;
  rq.type = device_type;
  rq.calculate_penetration = request.calculate_penetration ? 1 : 0;
  rq.calculate_distance = request.calculate_distance ? 1 : 0;
  int64_t cap = std::max<int64_t>(256, 4 * n);
  for (;;)
  {
    contacts.resize((size_t)cap);
    if (cast_)
      local.assign((size_t)cap * 6, 0.f);
    if (hit)
      hit->assign((size_t)std::max<int64_t>(n, 1), 0);
    t4c_result r;
    std::memset(&r, 0, sizeof(r));
    r.hit = hit ? hit->data() : nullptr;
    r.contacts = contacts.data();
    r.capacity = cap;
    r.cc_local = cast_ ? local.data() : nullptr;
    const t4c_status st = cast_ ? t4c_check_link_poses_cast(scene_, p0, p1, n, &rq, &r) :
                                  t4c_check_link_poses_discrete(scene_, p0, n, &rq, &r);
    if (st == T4C_ERR_CAPACITY)
    {
      cap = r.n_contacts;
      continue;
    }
    // a cast check of a non-convex leaf is a runtime_error in the reference too (bullet_cast_bvh_manager.cpp:321-326)
    check(st, cast_ ? "t4c_check_link_poses_cast" : "t4c_check_link_poses_discrete");
    contacts.resize((size_t)r.n_contacts);
    stats_ = r.stats;
    return;
  }
}

void ManagerCore::checkLinkPoses(const double* poses0, const double* poses1, int64_t n, const ContactRequest& request,
                                 std::vector<uint8_t>& hit, std::vector<t4c_contact>& contacts)
{
  std::vector<float> local;
  run(poses0, cast_ ? poses1 : nullptr, n, request, (int)request.type, &hit, contacts, local);
}

// contactTest: one state.  Device records -> ContactResult (addDiscreteSingleResult / addCastSingleResult,
// bullet_utils.h:388-441,527-611) -> processResult (core/common.h:114-167).  The caller's map is NOT cleared
// (bullet_discrete_bvh_manager.cpp:269).
void ManagerCore::contactTest(ContactResultMap& collisions, const ContactRequest& request)
{
  if (request.type == ContactTestType::LIMITED)
    throw std::runtime_error("tesseract_b200: ContactTestType::LIMITED is not implemented (nor in the reference, "
                             "core/common.h:159-164)");
  const int nl = (int)objects_.size();
  std::vector<double> p0, p1;
  for (int i = 0; i < nl; ++i)
    if (isActive(i))
    {
      double a[12], b[12];
      toPose12(objects_[i].pose0, a);
      toPose12(objects_[i].pose1, b);
      p0.insert(p0.end(), a, a + 12);
      p1.insert(p1.end(), b, b + 12);
    }
  // a user validity callback has to see every candidate contact before the FIRST / CLOSEST reduction
  const int device_type = request.is_valid ? (int)T4C_TEST_ALL : (int)request.type;
  std::vector<t4c_contact> raw;
  std::vector<float> local;
  run(p0.data(), p1.data(), 1, request, device_type, nullptr, raw, local);
  std::vector<const t4c_contact*> recs(raw.size());
  std::vector<const float*> locs(raw.size(), nullptr);
  for (size_t k = 0; k < raw.size(); ++k)
  {
    recs[k] = &raw[k];
    if (cast_)
      locs[k] = &local[k * 6];
  }
  processRecords(collisions, request, recs.data(), locs.data(), raw.size(), nullptr, nullptr);
}

bool ManagerCore::processRecords(ContactResultMap& collisions, const ContactRequest& request,
                                 const t4c_contact* const* recs, const float* const* locals, size_t n,
                                 const double* poses0, const double* poses1) const
{
  for (size_t k = 0; k < n; ++k)
  {
    const t4c_contact& c = *recs[k];
    ContactResult cr;
    cr.distance = c.distance;
    cr.normal = Eigen::Vector3d(c.normal[0], c.normal[1], c.normal[2]);
    for (int j = 0; j < 2; ++j)
    {
      const Object& o = objects_[(size_t)c.link[j]];
      // only active links are moved by a state (checkTrajectoryState, core/utils.h:136-137)
      const bool moved = poses0 && isActive(c.link[j]);
      const Eigen::Isometry3d T0 = moved ? fromPose12(poses0 + (size_t)c.link[j] * 12) : o.pose0;
      cr.link_names[j] = o.name;
      cr.type_id[j] = o.mask_id;
      cr.shape_id[j] = c.shape_id[j];
      cr.subshape_id[j] = c.subshape_id[j];
      cr.nearest_points[j] = Eigen::Vector3d(c.point[j][0], c.point[j][1], c.point[j][2]);
      cr.transform[j] = T0;
      cr.nearest_points_local[j] = T0.inverse() * cr.nearest_points[j];
      if (cast_ && c.cc_type[j] != T4C_CC_NONE)
      {
        cr.cc_time[j] = c.cc_time[j];
        cr.cc_type[j] = static_cast<ContinuousCollisionType>(c.cc_type[j]);
        cr.cc_transform[j] = (moved && poses1) ? fromPose12(poses1 + (size_t)c.link[j] * 12) : o.pose1;
        if (locals && locals[k])
        {
          const float* L = locals[k] + 3 * (size_t)j;
          cr.nearest_points_local[j] = Eigen::Vector3d(L[0], L[1], L[2]);
        }
      }
    }
    // processResult (core/common.h:114-167)
    if (request.is_valid && !request.is_valid(cr))
      continue;
    const auto key = cr.link_names[0] < cr.link_names[1] ? std::make_pair(cr.link_names[0], cr.link_names[1]) :
                                                            std::make_pair(cr.link_names[1], cr.link_names[0]);
    if ((request.calculate_distance || request.calculate_penetration) &&
        cr.distance > margins_.getPairCollisionMargin(key.first, key.second))
      continue;
    auto it = collisions.find(key);
    if (it == collisions.end() || it->second.empty())
    {
      collisions[key].push_back(cr);
      if (request.type == ContactTestType::FIRST)
        return true;
    }
    else if (request.type == ContactTestType::ALL)
      it->second.push_back(cr);
    else if (request.type == ContactTestType::CLOSEST && cr.distance < it->second[0].distance)
      it->second[0] = cr;
  }
  return false;
}

// checkTrajectory for many trajectories at once (tesseract_environment/core/utils.h:172-428): the LVS subdivision,
// FK and all contact tests run as one device batch (t4c_check_trajectories_*); the host then rebuilds, only for the
// colliding work items, what checkTrajectoryState / checkTrajectorySegment (:87-160) would have pushed.
void ManagerCore::checkTrajectories(const t4c_fk_model* model, const double* traj, int64_t n_traj, int32_t n_wp,
                                    int32_t n_q, double lvs, const ContactRequest& request,
                                    std::vector<std::vector<ContactResultMap>>& contacts, std::vector<uint8_t>& found)
{
  if (request.type == ContactTestType::LIMITED)
    throw std::runtime_error("tesseract_b200: ContactTestType::LIMITED is not implemented (nor in the reference, "
                             "core/common.h:159-164)");
  if (!model || !traj || n_traj < 0 || n_q <= 0)
    throw std::runtime_error("tesseract_b200: checkTrajectories was given a null model / trajectory");
  if (cast_ && n_wp == 1)
    throw std::runtime_error("checkTrajectory was given continuous contact manager with a trajectory that only has "
                             "one state.");  // core/utils.h:184-186
  contacts.assign((size_t)n_traj, {});
  found.assign((size_t)n_traj, 0);
  if (n_traj == 0)
    return;
  sync();
  int64_t n_items = 0;
  check(t4c_lvs_count(traj, n_traj, n_wp, n_q, lvs, cast_ ? 1 : 0, &n_items), "t4c_lvs_count");
  std::vector<int32_t> item_traj((size_t)std::max<int64_t>(n_items, 1));
  check(t4c_lvs_items(traj, n_traj, n_wp, n_q, lvs, cast_ ? 1 : 0, n_items, item_traj.data(), nullptr, nullptr,
                      &n_items),
        "t4c_lvs_items");
  t4c_request rq;
  rq.type = request.is_valid ? (int)T4C_TEST_ALL : (int)request.type;
  rq.calculate_penetration = request.calculate_penetration ? 1 : 0;
  rq.calculate_distance = request.calculate_distance ? 1 : 0;
  std::vector<uint8_t> hit((size_t)std::max<int64_t>(n_items, 1), 0);
  std::vector<t4c_contact> raw;
  std::vector<float> local;
  int64_t cap = std::max<int64_t>(256, 2 * n_items);
  for (;;)
  {
    raw.resize((size_t)cap);
    if (cast_)
      local.assign((size_t)cap * 6, 0.f);
    t4c_result r;
    std::memset(&r, 0, sizeof(r));
    r.hit = hit.data();
    r.contacts = raw.data();
    r.capacity = cap;
    r.cc_local = cast_ ? local.data() : nullptr;
    int64_t ni = 0;
    const t4c_status st = cast_ ? t4c_check_trajectories_cast(scene_, model, traj, n_traj, n_wp, lvs, &rq, &r, &ni) :
                                  t4c_check_trajectories_discrete(scene_, model, traj, n_traj, n_wp, lvs, &rq, &r, &ni);
    if (st == T4C_ERR_CAPACITY)
    {
      cap = r.n_contacts;
      continue;
    }
    check(st, cast_ ? "t4c_check_trajectories_cast" : "t4c_check_trajectories_discrete");
    raw.resize((size_t)r.n_contacts);
    stats_ = r.stats;
    break;
  }
  if (raw.empty())
    return;
  // records grouped by work item, items ascending (= the reference's visiting order)
  std::vector<uint32_t> order(raw.size());
  for (size_t k = 0; k < raw.size(); ++k)
    order[k] = (uint32_t)k;
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return raw[a].state < raw[b].state; });
  struct Group
  {
    int64_t item;
    size_t begin, end;
  };
  std::vector<Group> groups;
  for (size_t k = 0; k < order.size();)
  {
    size_t e = k;
    while (e < order.size() && raw[order[e]].state == raw[order[k]].state)
      ++e;
    groups.push_back({ raw[order[k]].state, k, e });
    k = e;
  }
  const bool first = request.type == ContactTestType::FIRST;
  if (first && !request.is_valid)
  {
    // only the first colliding item of a trajectory is ever reported (core/utils.h:226-231)
    std::vector<Group> keep;
    for (const Group& g : groups)
      if (keep.empty() || item_traj[(size_t)keep.back().item] != item_traj[(size_t)g.item])
        keep.push_back(g);
    groups.swap(keep);
  }
  // link poses of the selected items, in blocks bounded by the size of the FK output (every scene link)
  const size_t nl = objects_.size();
  const size_t block = std::max<size_t>(1, (size_t)(64u << 20) / (nl * 12 * sizeof(double) * (cast_ ? 2 : 1)));
  std::vector<int64_t> sel;
  std::vector<double> q0, q1, T0, T1;
  std::vector<const t4c_contact*> recs;
  std::vector<const float*> locs;
  for (size_t g0 = 0; g0 < groups.size(); g0 += block)
  {
    const size_t g1 = std::min(groups.size(), g0 + block), nb = g1 - g0;
    sel.resize(nb);
    for (size_t g = g0; g < g1; ++g)
      sel[g - g0] = groups[g].item;
    q0.resize(nb * (size_t)n_q);
    T0.resize(nb * nl * 12);
    if (cast_)
    {
      q1.resize(nb * (size_t)n_q);
      T1.resize(nb * nl * 12);
    }
    check(t4c_lvs_item_states(traj, n_traj, n_wp, n_q, lvs, cast_ ? 1 : 0, (int64_t)nb, sel.data(), q0.data(),
                              cast_ ? q1.data() : nullptr),
          "t4c_lvs_item_states");
    check(t4c_fk(model, scene_, q0.data(), (int64_t)nb, T0.data()), "t4c_fk");
    if (cast_)
      check(t4c_fk(model, scene_, q1.data(), (int64_t)nb, T1.data()), "t4c_fk");
    for (size_t g = g0; g < g1; ++g)
    {
      const Group& G = groups[g];
      const size_t t = (size_t)item_traj[(size_t)G.item];
      if (first && found[t])
        continue;
      recs.clear();
      locs.clear();
      for (size_t k = G.begin; k < G.end; ++k)
      {
        recs.push_back(&raw[order[k]]);
        locs.push_back(cast_ ? &local[(size_t)order[k] * 6] : nullptr);
      }
      ContactResultMap collisions;
      processRecords(collisions, request, recs.data(), locs.data(), recs.size(), &T0[(g - g0) * nl * 12],
                     cast_ ? &T1[(g - g0) * nl * 12] : nullptr);
      if (!collisions.empty())
      {
        contacts[t].push_back(std::move(collisions));
        found[t] = 1;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// the two interface classes: thin forwards to ManagerCore
// ---------------------------------------------------------------------------------------------------------------
#define T4_FORWARD_COMMON(CLASS)                                                                                       \
  bool CLASS::addCollisionObject(const std::string& name, const int& mask_id, const CollisionShapesConst& shapes,      \
                                 const tesseract_common::VectorIsometry3d& shape_poses, bool enabled)                  \
  {                                                                                                                    \
    return core_->addCollisionObject(name, mask_id, shapes, shape_poses, enabled);                                     \
  }                                                                                                                    \
  const CollisionShapesConst& CLASS::getCollisionObjectGeometries(const std::string& name) const                       \
  {                                                                                                                    \
    return core_->getGeometries(name);                                                                                 \
  }                                                                                                                    \
  const tesseract_common::VectorIsometry3d& CLASS::getCollisionObjectGeometriesTransforms(const std::string& name)     \
      const                                                                                                            \
  {                                                                                                                    \
    return core_->getGeometriesTransforms(name);                                                                       \
  }                                                                                                                    \
  bool CLASS::hasCollisionObject(const std::string& name) const { return core_->hasCollisionObject(name); }            \
  bool CLASS::removeCollisionObject(const std::string& name) { return core_->removeCollisionObject(name); }            \
  bool CLASS::enableCollisionObject(const std::string& name) { return core_->setEnabled(name, true); }                 \
  bool CLASS::disableCollisionObject(const std::string& name) { return core_->setEnabled(name, false); }               \
  void CLASS::setCollisionObjectsTransform(const std::string& name, const Eigen::Isometry3d& pose)                     \
  {                                                                                                                    \
    core_->setTransform(name, pose);                                                                                   \
  }                                                                                                                    \
  void CLASS::setCollisionObjectsTransform(const std::vector<std::string>& names,                                      \
                                           const tesseract_common::VectorIsometry3d& poses)                            \
  {                                                                                                                    \
    for (size_t i = 0; i < names.size() && i < poses.size(); ++i)                                                      \
      core_->setTransform(names[i], poses[i]);                                                                         \
  }                                                                                                                    \
  void CLASS::setCollisionObjectsTransform(const tesseract_common::TransformMap& transforms)                           \
  {                                                                                                                    \
    for (const auto& t : transforms)                                                                                   \
      core_->setTransform(t.first, t.second);                                                                          \
  }                                                                                                                    \
  const std::vector<std::string>& CLASS::getCollisionObjects() const { return core_->getCollisionObjects(); }          \
  void CLASS::setActiveCollisionObjects(const std::vector<std::string>& names) { core_->setActive(names); }            \
  const std::vector<std::string>& CLASS::getActiveCollisionObjects() const { return core_->getActive(); }              \
  void CLASS::setCollisionMarginData(CollisionMarginData d, CollisionMarginOverrideType t)                             \
  {                                                                                                                    \
    core_->setMarginData(std::move(d), t);                                                                             \
  }                                                                                                                    \
  void CLASS::setDefaultCollisionMarginData(double m) { core_->setDefaultMargin(m); }                                  \
  void CLASS::setPairCollisionMarginData(const std::string& a, const std::string& b, double m)                         \
  {                                                                                                                    \
    core_->setPairMargin(a, b, m);                                                                                     \
  }                                                                                                                    \
  const CollisionMarginData& CLASS::getCollisionMarginData() const { return core_->getMarginData(); }                  \
  void CLASS::setIsContactAllowedFn(IsContactAllowedFn fn) { core_->setIsContactAllowedFn(std::move(fn)); }            \
  IsContactAllowedFn CLASS::getIsContactAllowedFn() const { return core_->getIsContactAllowedFn(); }                   \
  void CLASS::contactTest(ContactResultMap& collisions, const ContactRequest& request)                                 \
  {                                                                                                                    \
    core_->contactTest(collisions, request);                                                                           \
  }

B200DiscreteManager::B200DiscreteManager(int device) : core_(new ManagerCore(false, device)) {}
B200DiscreteManager::B200DiscreteManager(std::unique_ptr<ManagerCore> core) : core_(std::move(core)) {}
std::shared_ptr<DiscreteContactManager> B200DiscreteManager::clone() const
{
  return std::shared_ptr<DiscreteContactManager>(new B200DiscreteManager(core_->clone()));
}
T4_FORWARD_COMMON(B200DiscreteManager)

B200CastManager::B200CastManager(int device) : core_(new ManagerCore(true, device)) {}
B200CastManager::B200CastManager(std::unique_ptr<ManagerCore> core) : core_(std::move(core)) {}
std::shared_ptr<ContinuousContactManager> B200CastManager::clone() const
{
  return std::shared_ptr<ContinuousContactManager>(new B200CastManager(core_->clone()));
}
T4_FORWARD_COMMON(B200CastManager)
void B200CastManager::setCollisionObjectsTransform(const std::string& name, const Eigen::Isometry3d& pose1,
                                                   const Eigen::Isometry3d& pose2)
{
  core_->setTransform(name, pose1, pose2);
}
void B200CastManager::setCollisionObjectsTransform(const std::vector<std::string>& names,
                                                   const tesseract_common::VectorIsometry3d& pose1,
                                                   const tesseract_common::VectorIsometry3d& pose2)
{
  for (size_t i = 0; i < names.size() && i < pose1.size() && i < pose2.size(); ++i)
    core_->setTransform(names[i], pose1[i], pose2[i]);
}
void B200CastManager::setCollisionObjectsTransform(const tesseract_common::TransformMap& pose1,
                                                   const tesseract_common::TransformMap& pose2)
{
  // both maps hold the same links (bullet_cast_bvh_manager.cpp:345-359)
  auto it1 = pose1.begin();
  auto it2 = pose2.begin();
  for (; it1 != pose1.end() && it2 != pose2.end(); ++it1, ++it2)
    core_->setTransform(it1->first, it1->second, it2->second);
}
}  // namespace tesseract_b200

extern "C" {
tesseract_collision::DiscreteContactManager* tesseract_b200_create_discrete_manager(void)
{
  return new tesseract_b200::B200DiscreteManager();
}
tesseract_collision::ContinuousContactManager* tesseract_b200_create_cast_manager(void)
{
  return new tesseract_b200::B200CastManager();
}
}

#ifdef T4_USE_TESSERACT_HEADERS
#include <tesseract_common/class_loader.h>
TESSERACT_ADD_PLUGIN(tesseract_b200::B200DiscreteManager, B200DiscreteManager)
TESSERACT_ADD_PLUGIN(tesseract_b200::B200CastManager, B200CastManager)
#endif
