/*
 * t4o_manager.cpp -- CPU ORACLE manager semantics (test infrastructure, not product code).
 *
 * Restates, in FP64:
 *   - BulletDiscreteBVHManager / BulletCastBVHManager object, filter, margin and contactTest semantics
 *       tesseract_collision/src/bullet/bullet_discrete_bvh_manager.cpp:53-305
 *       tesseract_collision/src/bullet/bullet_cast_bvh_manager.cpp:260-332,361-415,446-510
 *   - pair admissibility                  bullet_utils.h:381-386,900-917 ; core/common.h:66-104
 *   - broadphase pair set                 bullet_utils.h:170-177 (AABB grown by the max margin on every object);
 *                                         the tree shape is irrelevant to results, a static median-split
 *                                         AABB tree + brute-force active x active is used
 *   - compound child descent              tesseract_compound_collision_algorithm.cpp:142-240,323-338
 *   - result bridging + processResult     bullet_utils.h:388-441,757-802 ; core/common.h:114-167
 *   - cast continuous data                bullet_utils.h:306-349,452-611
 *   - FK                                  tesseract_environment/src/ofkt/ofkt_nodes.cpp:200-289,
 *                                         tesseract_scene_graph/.../parser/kdl_parser.h:86-118
 *   - LVS subdivision                     tesseract_environment/include/tesseract_environment/core/utils.h:189-205,328-347
 */
#include "t4o.h"

#include <algorithm>
#include <map>
#include <thread>
#include <unordered_map>
#include <unordered_set>

using namespace t4o;

namespace
{
constexpr double BULLET_SUPPORT_FUNC_TOLERANCE = 0.01;
constexpr double BULLET_LENGTH_TOLERANCE = 0.001;
constexpr double BULLET_EPSILON = 1e-3;

struct Shape
{
  Convex cv;
  Convex cast_cv;  // CastHullShape twin (cast managers, active links)
  int link = 0, shape_id = 0, subshape_id = -1;
  int vert_offset = 0;
  Xf local;
};

struct Link
{
  std::string name;
  int type_id = 0;
  bool enabled = true;
  bool active = false;
  Xf pose, pose1;
  std::vector<int> shapes;
  V3 mn, mx;  // world AABB, grown by the max margin
};

struct BvhNode
{
  V3 mn, mx;
  int left = -1, right = -1;  // children, or left = -(link+1) for leaves
};

inline uint64_t pairKey(int a, int b)
{
  if (a > b)
    std::swap(a, b);
  return (uint64_t)(uint32_t)a << 32 | (uint32_t)b;
}

inline bool overlap(const V3& mn0, const V3& mx0, const V3& mn1, const V3& mx1)
{
  // TestAabbAgainstAabb2: inclusive
  return !(mn0.x > mx1.x || mx0.x < mn1.x || mn0.y > mx1.y || mx0.y < mn1.y || mn0.z > mx1.z || mx0.z < mn1.z);
}
}  // namespace

struct t4o_manager
{
  bool cast = false;
  std::vector<V3> verts;
  std::vector<Shape> shapes;
  std::vector<Link> links;
  std::unordered_set<uint64_t> acm;
  double default_margin = 0;
  std::unordered_map<uint64_t, double> pair_margins;
  double max_margin = 0;
  std::vector<BvhNode> bvh;
  std::vector<int> static_links, active_links;
  bool bvh_dirty = true;
  Counters cnt;

  // per-contactTest state
  t4c_request req;
  bool done = false;
  std::map<uint64_t, std::vector<t4o_contact>> res;

  void fixPointers()
  {
    for (auto& s : shapes)
    {
      if (s.cv.type == T4C_HULL || s.cv.type == T4C_TRIANGLE)
        s.cv.verts = verts.data() + s.vert_offset;
      s.cast_cv.inner = &s.cv;
    }
  }

  double pairMargin(int a, int b) const
  {
    auto it = pair_margins.find(pairKey(a, b));
    return it == pair_margins.end() ? default_margin : it->second;
  }
  void updateMaxMargin()
  {
    max_margin = default_margin;
    for (auto& kv : pair_margins)
      max_margin = std::max(max_margin, kv.second);
    bvh_dirty = true;
  }
  void refreshLists()
  {
    static_links.clear();
    active_links.clear();
    for (int i = 0; i < (int)links.size(); ++i)
    {
      if (links[i].shapes.empty())
        continue;
      (links[i].active ? active_links : static_links).push_back(i);
    }
    bvh_dirty = true;
  }

  const Convex& shapeConvex(const Shape& s) const { return (cast && links[s.link].active) ? s.cast_cv : s.cv; }

  void updateCast(int li)
  {
    Link& l = links[li];
    for (int si : l.shapes)
    {
      Shape& s = shapes[si];
      // bullet_cast_bvh_manager.cpp:281-305: delta = (tf1 * local)^-1 * (tf2 * local)
      s.cast_cv.t01 = (l.pose * s.local).inverseTimes(l.pose1 * s.local);
    }
  }

  void shapeAabb(const Shape& s, V3& mn, V3& mx) const
  {
    const Link& l = links[s.link];
    shapeConvex(s).aabb(l.pose * s.local, mn, mx);
  }

  void linkAabb(int li)
  {
    Link& l = links[li];
    l.mn = V3(LARGE, LARGE, LARGE);
    l.mx = V3(-LARGE, -LARGE, -LARGE);
    for (int si : l.shapes)
    {
      V3 mn, mx;
      shapeAabb(shapes[si], mn, mx);
      for (int k = 0; k < 3; ++k)
      {
        l.mn[k] = std::min(l.mn[k], mn[k]);
        l.mx[k] = std::max(l.mx[k], mx[k]);
      }
    }
    // COW::getAABB (bullet_utils.h:170-177): grow by the contact processing threshold = max margin
    l.mn -= V3(max_margin, max_margin, max_margin);
    l.mx += V3(max_margin, max_margin, max_margin);
  }

  int buildNode(std::vector<int>& ids, int lo, int hi)
  {
    BvhNode node;
    node.mn = V3(LARGE, LARGE, LARGE);
    node.mx = V3(-LARGE, -LARGE, -LARGE);
    for (int i = lo; i < hi; ++i)
      for (int k = 0; k < 3; ++k)
      {
        node.mn[k] = std::min(node.mn[k], links[ids[i]].mn[k]);
        node.mx[k] = std::max(node.mx[k], links[ids[i]].mx[k]);
      }
    int idx = (int)bvh.size();
    bvh.push_back(node);
    if (hi - lo == 1)
    {
      bvh[idx].left = -(ids[lo] + 1);
      return idx;
    }
    V3 ext = node.mx - node.mn;
    int ax = ext.x > ext.y ? (ext.x > ext.z ? 0 : 2) : (ext.y > ext.z ? 1 : 2);
    int mid = (lo + hi) / 2;
    std::nth_element(ids.begin() + lo, ids.begin() + mid, ids.begin() + hi, [&](int a, int b) {
      return links[a].mn[ax] + links[a].mx[ax] < links[b].mn[ax] + links[b].mx[ax];
    });
    int l = buildNode(ids, lo, mid);
    int r = buildNode(ids, mid, hi);
    bvh[idx].left = l;
    bvh[idx].right = r;
    return idx;
  }

  void rebuildStatic()
  {
    bvh.clear();
    for (int li : static_links)
      linkAabb(li);
    if (!static_links.empty())
    {
      std::vector<int> ids = static_links;
      bvh.reserve(ids.size() * 2);
      buildNode(ids, 0, (int)ids.size());
    }
    bvh_dirty = false;
  }

  bool needsCollisionCheck(int a, int b) const
  {
    const Link &la = links[a], &lb = links[b];
    // bullet_utils.h:381-386 + filter groups :900-917 (static x static never tested)
    if (!la.enabled || !lb.enabled)
      return false;
    if (!la.active && !lb.active)
      return false;
    if (a == b)
      return false;
    if (acm.count(pairKey(a, b)))
      return false;
    return true;
  }

  // GetAverageSupport (bullet_utils.h:306-349)
  static void averageSupport(const Convex& s, const V3& n, double& outsupport, V3& outpt)
  {
    if (s.polyhedral())
    {
      V3 ptSum(0, 0, 0);
      double ptCount = 0;
      double maxSupport = -1000;
      int npts = s.type == T4C_BOX ? 8 : s.nverts;
      for (int i = 0; i < npts; ++i)
      {
        V3 pt;
        if (s.type == T4C_BOX)
          pt = V3(s.he.x * (1 - (i & 1)) - s.he.x * (i & 1), s.he.y * (1 - ((i & 2) >> 1)) - s.he.y * ((i & 2) >> 1),
                  s.he.z * (1 - ((i & 4) >> 2)) - s.he.z * ((i & 4) >> 2));
        else
          pt = s.verts[i];
        double sup = pt.dot(n);
        if (sup > maxSupport + BULLET_EPSILON)
        {
          ptCount = 1;
          ptSum = pt;
          maxSupport = sup;
        }
        else if (sup < maxSupport - BULLET_EPSILON)
        {
        }
        else
        {
          ptCount += 1;
          ptSum += pt;
        }
      }
      outsupport = maxSupport;
      outpt = ptSum / ptCount;
    }
    else
    {
      // localGetSupportingVertex (with margin, virtual path)
      V3 sv = s.supportNoMargin(n);
      double m = s.margin();
      if (m != 0.0)
      {
        V3 nn = n;
        if (nn.length2() < SIMD_EPS * SIMD_EPS)
          nn = V3(-1, -1, -1);
        sv += nn / nn.length() * m;
      }
      outpt = sv;
      outsupport = n.dot(outpt);
    }
  }

  // calculateContinuousData (bullet_utils.h:452-525)
  void continuousData(t4o_contact& col, const Shape& s, const V3& pt_world, const V3& normal_world,
                      const Xf& link_tf_inv, int link_index)
  {
    const Link& l = links[s.link];
    Xf tfWorld0 = l.pose * s.local;
    Xf tfWorld1 = tfWorld0 * s.cast_cv.t01;
    Xf link_tf = Xf::from12(col.transform[link_index]);
    Xf sxf = link_tf.inverse() * tfWorld0;
    (tfWorld1 * sxf.inverse()).to12(col.cc_transform[link_index]);
    V3 nl0 = tfWorld0.R.tmul(normal_world);
    V3 nl1 = tfWorld1.R.tmul(normal_world);
    V3 ptLocal0, ptLocal1;
    double localsup0, localsup1;
    averageSupport(s.cv, nl0, localsup0, ptLocal0);
    V3 ptWorld0 = tfWorld0 * ptLocal0;
    averageSupport(s.cv, nl1, localsup1, ptLocal1);
    V3 ptWorld1 = tfWorld1 * ptLocal1;
    double sup0 = normal_world.dot(ptWorld0);
    double sup1 = normal_world.dot(ptWorld1);
    if (sup0 - sup1 > BULLET_SUPPORT_FUNC_TOLERANCE)
    {
      col.cc_time[link_index] = 0;
      col.cc_type[link_index] = T4C_CC_TIME0;
    }
    else if (sup1 - sup0 > BULLET_SUPPORT_FUNC_TOLERANCE)
    {
      col.cc_time[link_index] = 1;
      col.cc_type[link_index] = T4C_CC_TIME1;
    }
    else
    {
      double l0c = (pt_world - ptWorld0).length();
      double l1c = (pt_world - ptWorld1).length();
      V3 pl = link_tf_inv * (tfWorld0 * ((ptLocal0 + ptLocal1) / 2.0));
      for (int k = 0; k < 3; ++k)
        col.point_local[link_index][k] = pl[k];
      col.cc_type[link_index] = T4C_CC_BETWEEN;
      if (l0c + l1c < BULLET_LENGTH_TOLERANCE)
        col.cc_time[link_index] = .5;
      else
        col.cc_time[link_index] = l0c / (l0c + l1c);
    }
  }

  // TesseractBroadphaseBridgedManifoldResult::addContactPoint -> add{Discrete,Cast}SingleResult -> processResult
  void addContact(const Shape& s0, const Shape& s1, const PairResult& pr, int64_t state_tag)
  {
    if (done || pr.distance > max_margin)
      return;
    V3 pointA = pr.pointOnB + pr.normalOnB * pr.distance;
    const Link &l0 = links[s0.link], &l1 = links[s1.link];
    t4o_contact c;
    std::memset(&c, 0, sizeof(c));
    c.state = state_tag;
    c.link[0] = s0.link;
    c.link[1] = s1.link;
    c.shape_id[0] = s0.shape_id;
    c.shape_id[1] = s1.shape_id;
    c.subshape_id[0] = s0.subshape_id;
    c.subshape_id[1] = s1.subshape_id;
    c.distance = pr.distance;
    Xf tf0_inv = l0.pose.inverse(), tf1_inv = l1.pose.inverse();
    V3 la = tf0_inv * pointA, lb = tf1_inv * pr.pointOnB;
    for (int k = 0; k < 3; ++k)
    {
      c.point[0][k] = pointA[k];
      c.point[1][k] = pr.pointOnB[k];
      c.point_local[0][k] = la[k];
      c.point_local[1][k] = lb[k];
      c.normal[k] = -pr.normalOnB[k];
    }
    l0.pose.to12(c.transform[0]);
    l1.pose.to12(c.transform[1]);
    Xf ident;
    ident.to12(c.cc_transform[0]);
    ident.to12(c.cc_transform[1]);
    c.cc_time[0] = c.cc_time[1] = -1;
    c.cc_type[0] = c.cc_type[1] = T4C_CC_NONE;

    // processResult (core/common.h:114-167); is_valid callback is a host-side concern (not modelled)
    if ((req.calculate_distance || req.calculate_penetration) && c.distance > pairMargin(s0.link, s1.link))
      return;
    uint64_t key = pairKey(s0.link, s1.link);
    auto it = res.find(key);
    t4o_contact* col = nullptr;
    if (it == res.end())
    {
      if (req.type == T4C_TEST_FIRST)
        done = true;
      auto& v = res[key];
      v.push_back(c);
      col = &v.back();
    }
    else if (req.type == T4C_TEST_ALL)
    {
      it->second.push_back(c);
      col = &it->second.back();
    }
    else if (req.type == T4C_TEST_CLOSEST)
    {
      if (c.distance < it->second[0].distance)
      {
        it->second[0] = c;
        col = &it->second[0];
      }
    }
    if (!col || !cast)
      return;

    // addCastSingleResult (bullet_utils.h:581-608)
    V3 normalOnB = pr.normalOnB;
    if (l0.active && l1.active)
    {
      continuousData(*col, s0, pointA, -normalOnB, tf0_inv, 0);
      continuousData(*col, s1, pr.pointOnB, normalOnB, tf1_inv, 1);
    }
    else
    {
      bool castShapeIsFirst = l0.active;
      V3 normalWorldFromCast = normalOnB * (-(castShapeIsFirst ? 1.0 : -1.0));
      const Shape& first = castShapeIsFirst ? s0 : s1;
      const Xf& first_tf_inv = castShapeIsFirst ? tf0_inv : tf1_inv;
      V3 ptOnCast = castShapeIsFirst ? pointA : pr.pointOnB;
      if (castShapeIsFirst)
      {
        for (int k = 0; k < 3; ++k)
        {
          std::swap(col->point[0][k], col->point[1][k]);
          std::swap(col->point_local[0][k], col->point_local[1][k]);
          col->normal[k] *= -1;
        }
        for (int k = 0; k < 12; ++k)
          std::swap(col->transform[0][k], col->transform[1][k]);
        std::swap(col->link[0], col->link[1]);
        std::swap(col->shape_id[0], col->shape_id[1]);
        std::swap(col->subshape_id[0], col->subshape_id[1]);
      }
      continuousData(*col, first, ptOnCast, normalWorldFromCast, first_tf_inv, 1);
    }
  }

  void collideLinks(int a, int b, int64_t state_tag)
  {
    if (done || !needsCollisionCheck(a, b))
      return;
    const Link &la = links[a], &lb = links[b];
    PairRequest preq;
    preq.calculate_penetration = req.calculate_penetration != 0;
    preq.calculate_distance = req.calculate_distance != 0;
    const bool single = la.shapes.size() == 1 && lb.shapes.size() == 1;
    const V3 ext(max_margin, max_margin, max_margin);
    for (int sa : la.shapes)
    {
      const Shape& A = shapes[sa];
      V3 mnA, mxA;
      if (!single)
      {
        shapeAabb(A, mnA, mxA);
        mnA -= ext;
        mxA += ext;
      }
      for (int sb : lb.shapes)
      {
        if (done)
          return;
        const Shape& B = shapes[sb];
        if (!single)
        {
          // compound child AABB re-test (tesseract_compound_collision_algorithm.cpp:158-171)
          V3 mnB, mxB;
          shapeAabb(B, mnB, mxB);
          if (!overlap(mnA, mxA, mnB, mxB))
            continue;
        }
        PairResult pr;
        collideConvexConvex(shapeConvex(A), la.pose * A.local, shapeConvex(B), lb.pose * B.local, max_margin, preq, pr,
                            cnt);
        if (pr.has)
          addContact(A, B, pr, state_tag);
      }
    }
  }

  void queryStatic(int node, int a, int64_t state_tag)
  {
    const BvhNode& n = bvh[node];
    const Link& la = links[a];
    if (!overlap(la.mn, la.mx, n.mn, n.mx))
      return;
    if (n.left < 0)
    {
      collideLinks(a, -n.left - 1, state_tag);
      return;
    }
    queryStatic(n.left, a, state_tag);
    if (!done)
      queryStatic(n.right, a, state_tag);
  }

  int64_t contactTest(const t4c_request& r, int64_t state_tag, t4o_contact* out, int64_t capacity)
  {
    req = r;
    done = false;
    res.clear();
    if (bvh_dirty)
      rebuildStatic();
    for (int li : active_links)
    {
      if (cast)
        updateCast(li);
      linkAabb(li);
    }
    for (size_t i = 0; i < active_links.size() && !done; ++i)
    {
      int a = active_links[i];
      if (!links[a].enabled)
        continue;
      for (size_t j = i + 1; j < active_links.size() && !done; ++j)
      {
        int b = active_links[j];
        if (overlap(links[a].mn, links[a].mx, links[b].mn, links[b].mx))
          collideLinks(a, b, state_tag);
      }
      if (!bvh.empty() && !done)
        queryStatic(0, a, state_tag);
    }
    int64_t n = 0;
    for (auto& kv : res)
      for (auto& c : kv.second)
      {
        if (out && n < capacity)
          out[n] = c;
        ++n;
      }
    cnt.contacts += (uint64_t)n;
    return n;
  }
};

// -------------------------------------------------------------------------------------------------
// FK
// -------------------------------------------------------------------------------------------------
namespace
{
M3 angleAxis(const V3& axis_in, double q)
{
  // Eigen::AngleAxisd(q, axis.normalized()).toRotationMatrix()  (ofkt_nodes.cpp:200,217-219)
  V3 a = axis_in / axis_in.length();
  double c = std::cos(q), s = std::sin(q), t = 1 - c;
  M3 r;
  r.m[0][0] = t * a.x * a.x + c;
  r.m[0][1] = t * a.x * a.y - s * a.z;
  r.m[0][2] = t * a.x * a.z + s * a.y;
  r.m[1][0] = t * a.x * a.y + s * a.z;
  r.m[1][1] = t * a.y * a.y + c;
  r.m[1][2] = t * a.y * a.z - s * a.x;
  r.m[2][0] = t * a.x * a.z - s * a.y;
  r.m[2][1] = t * a.y * a.z + s * a.x;
  r.m[2][2] = t * a.z * a.z + c;
  return r;
}

void fkOne(const t4c_scene_desc* scene, const t4c_fk_model_desc* model, const double* q, std::vector<Xf>& T)
{
  T.resize((size_t)model->n_links);
  for (int i = 0; i < model->n_links; ++i)
    T[i] = i < scene->n_links ? Xf::from12(scene->links[i].pose) : Xf();
  for (int j = 0; j < model->n_joints; ++j)
  {
    const t4c_joint& jt = model->joints[j];
    Xf origin = Xf::from12(jt.origin);
    Xf local = origin;
    if (jt.type == T4C_JOINT_REVOLUTE)
    {
      Xf rot;
      rot.R = angleAxis(V3(jt.axis[0], jt.axis[1], jt.axis[2]), q[jt.q_index]);
      local = origin * rot;
    }
    else if (jt.type == T4C_JOINT_PRISMATIC)
    {
      // ofkt_nodes.cpp:272,287-289: T_origin * Translation(q * axis.normalized())
      Xf tr;
      V3 ax(jt.axis[0], jt.axis[1], jt.axis[2]);
      tr.p = ax / ax.length() * q[jt.q_index];
      local = origin * tr;
    }
    T[jt.child_link] = T[jt.parent_link] * local;
  }
}

t4o_manager* buildManager(const t4c_scene_desc* d, int cast)
{
  auto* m = new t4o_manager();
  m->cast = cast != 0;
  m->verts.resize((size_t)d->n_vertices);
  for (int i = 0; i < d->n_vertices; ++i)
    m->verts[i] = V3(d->vertices[3 * i], d->vertices[3 * i + 1], d->vertices[3 * i + 2]);
  m->links.resize((size_t)d->n_links);
  for (int i = 0; i < d->n_links; ++i)
  {
    Link& l = m->links[i];
    l.name = d->links[i].name ? d->links[i].name : "";
    l.type_id = d->links[i].type_id;
    l.enabled = d->links[i].enabled != 0;
    l.active = d->links[i].active != 0;
    l.pose = l.pose1 = Xf::from12(d->links[i].pose);
  }
  m->shapes.resize((size_t)d->n_shapes);
  for (int i = 0; i < d->n_shapes; ++i)
  {
    const t4c_shape& sd = d->shapes[i];
    Shape& s = m->shapes[i];
    s.link = sd.link;
    s.shape_id = sd.shape_id;
    s.subshape_id = sd.subshape_id;
    s.local = Xf::from12(sd.pose);
    s.cv.type = sd.type;
    switch (sd.type)
    {
      case T4C_SPHERE:
        s.cv.r = sd.params[0];
        break;
      case T4C_BOX:
        s.cv.he = V3(sd.params[0], sd.params[1], sd.params[2]);
        break;
      case T4C_CYLINDER:
      case T4C_CONE:
      case T4C_CAPSULE:
        s.cv.r = sd.params[0];
        s.cv.h = sd.params[1];
        break;
      default:
        s.vert_offset = sd.vertex_offset;  // pointer resolved in fixPointers()
        s.cv.nverts = sd.type == T4C_TRIANGLE ? 3 : sd.vertex_count;
        break;
    }
    s.cast_cv.is_cast = true;
    s.cast_cv.type = sd.type;
    m->links[sd.link].shapes.push_back(i);
  }
  m->fixPointers();
  for (int i = 0; i < d->n_acm_pairs; ++i)
    m->acm.insert(pairKey(d->acm_pairs[2 * i], d->acm_pairs[2 * i + 1]));
  m->default_margin = d->default_margin;
  for (int i = 0; i < d->n_margin_pairs; ++i)
    m->pair_margins[pairKey(d->margin_pairs[2 * i], d->margin_pairs[2 * i + 1])] = d->margin_values[i];
  m->updateMaxMargin();
  m->refreshLists();
  return m;
}

t4o_manager* cloneManager(const t4o_manager* src)
{
  auto* m = new t4o_manager(*src);
  m->fixPointers();  // re-base vertex pointers onto the clone's own pool
  m->res.clear();
  m->cnt = Counters();
  return m;
}

void linspaced(int64_t cnt, double low, double high, int64_t i, double& out)
{
  // Eigen::VectorXd::LinSpaced (Eigen 3.3 linspaced_op_impl, non-integer)
  int64_t size1 = cnt == 1 ? 1 : cnt - 1;
  double step = cnt == 1 ? 0.0 : (high - low) / double(cnt - 1);
  bool flip = std::fabs(high) < std::fabs(low);
  if (flip)
    out = (i == 0) ? low : (high - double(size1 - i) * step);
  else
    out = (i == size1) ? high : (low + double(i) * step);
}
}  // namespace

extern "C" {
t4o_manager* t4o_manager_create(const t4c_scene_desc* desc, int cast) { return buildManager(desc, cast); }
t4o_manager* t4o_manager_clone(const t4o_manager* m) { return cloneManager(m); }
void t4o_manager_destroy(t4o_manager* m) { delete m; }
void t4o_manager_set_active(t4o_manager* m, const int32_t* flags)
{
  for (size_t i = 0; i < m->links.size(); ++i)
    m->links[i].active = flags[i] != 0;
  m->refreshLists();
}
void t4o_manager_set_enabled(t4o_manager* m, int32_t link, int32_t enabled) { m->links[link].enabled = enabled != 0; }
void t4o_manager_set_acm(t4o_manager* m, int32_t n, const int32_t* pairs)
{
  m->acm.clear();
  for (int i = 0; i < n; ++i)
    m->acm.insert(pairKey(pairs[2 * i], pairs[2 * i + 1]));
}
void t4o_manager_set_margins(t4o_manager* m, double def, int32_t n, const int32_t* pairs, const double* values)
{
  m->default_margin = def;
  m->pair_margins.clear();
  for (int i = 0; i < n; ++i)
    m->pair_margins[pairKey(pairs[2 * i], pairs[2 * i + 1])] = values[i];
  m->updateMaxMargin();
}
void t4o_manager_set_pose(t4o_manager* m, int32_t link, const double* p)
{
  m->links[link].pose = m->links[link].pose1 = Xf::from12(p);
  if (!m->links[link].active)
    m->bvh_dirty = true;
}
void t4o_manager_set_pose_cast(t4o_manager* m, int32_t link, const double* p0, const double* p1)
{
  m->links[link].pose = Xf::from12(p0);
  m->links[link].pose1 = Xf::from12(p1);
  if (!m->links[link].active)
    m->bvh_dirty = true;
}
int64_t t4o_manager_contact_test(t4o_manager* m, const t4c_request* req, int64_t tag, t4o_contact* out, int64_t cap)
{
  return m->contactTest(*req, tag, out, cap);
}
void t4o_manager_get_counters(const t4o_manager* m, uint64_t* o)
{
  o[0] = m->cnt.candidates;
  o[1] = m->cnt.gjk_pairs;
  o[2] = m->cnt.gjk_iterations;
  o[3] = m->cnt.epa_pairs;
  o[4] = m->cnt.epa_iterations;
  o[5] = m->cnt.contacts;
}

void t4o_fk(const t4c_scene_desc* scene, const t4c_fk_model_desc* model, const double* q, int64_t n, double* out)
{
  std::vector<Xf> T;
  for (int64_t i = 0; i < n; ++i)
  {
    fkOne(scene, model, q + i * model->n_q, T);
    for (int l = 0; l < model->n_links; ++l)
      T[l].to12(out + (i * model->n_links + l) * 12);
  }
}

static int64_t runBatch(const t4c_scene_desc* scene, const t4c_fk_model_desc* model, const double* q0,
                        const double* q1, int64_t n, const t4c_request* req, uint8_t* hit, t4o_contact* out,
                        int64_t capacity, int n_threads, uint64_t* counters6, int cast)
{
  if (n_threads < 1)
    n_threads = 1;
  t4o_manager* base = buildManager(scene, cast);
  std::vector<std::vector<t4o_contact>> per_thread((size_t)n_threads);
  std::vector<Counters> cnts((size_t)n_threads);
  auto work = [&](int t) {
    t4o_manager* m = cloneManager(base);  // one clone per thread (discrete_contact_manager.h:59-65)
    int64_t lo = n * t / n_threads, hi = n * (t + 1) / n_threads;
    std::vector<Xf> T0, T1;
    std::vector<t4o_contact> buf(4096);
    for (int64_t i = lo; i < hi; ++i)
    {
      fkOne(scene, model, q0 + i * model->n_q, T0);
      if (cast)
        fkOne(scene, model, q1 + i * model->n_q, T1);
      for (int li : m->active_links)
      {
        m->links[li].pose = T0[li];
        m->links[li].pose1 = cast ? T1[li] : T0[li];
      }
      int64_t k = m->contactTest(*req, i, buf.data(), (int64_t)buf.size());
      if (k > (int64_t)buf.size())
      {
        buf.resize((size_t)k);
        k = m->contactTest(*req, i, buf.data(), (int64_t)buf.size());
      }
      if (hit)
        hit[i] = k > 0;
      if (out)
        per_thread[t].insert(per_thread[t].end(), buf.begin(), buf.begin() + k);
      else
        cnts[t].contacts += 0;
    }
    cnts[t] = m->cnt;
    delete m;
  };
  std::vector<std::thread> th;
  for (int t = 1; t < n_threads; ++t)
    th.emplace_back(work, t);
  work(0);
  for (auto& x : th)
    x.join();
  int64_t total = 0;
  Counters sum;
  for (int t = 0; t < n_threads; ++t)
  {
    for (auto& c : per_thread[t])
    {
      if (out && total < capacity)
        out[total] = c;
      ++total;
    }
    sum.candidates += cnts[t].candidates;
    sum.gjk_pairs += cnts[t].gjk_pairs;
    sum.gjk_iterations += cnts[t].gjk_iterations;
    sum.epa_pairs += cnts[t].epa_pairs;
    sum.epa_iterations += cnts[t].epa_iterations;
    sum.contacts += cnts[t].contacts;
  }
  if (!out)
    total = (int64_t)sum.contacts;
  if (counters6)
  {
    counters6[0] = sum.candidates;
    counters6[1] = sum.gjk_pairs;
    counters6[2] = sum.gjk_iterations;
    counters6[3] = sum.epa_pairs;
    counters6[4] = sum.epa_iterations;
    counters6[5] = sum.contacts;
  }
  delete base;
  return total;
}

int64_t t4o_check_discrete(const t4c_scene_desc* scene, const t4c_fk_model_desc* model, const double* q, int64_t n,
                           const t4c_request* req, uint8_t* hit, t4o_contact* out, int64_t capacity, int n_threads,
                           uint64_t* counters6)
{
  return runBatch(scene, model, q, nullptr, n, req, hit, out, capacity, n_threads, counters6, 0);
}

int64_t t4o_check_cast(const t4c_scene_desc* scene, const t4c_fk_model_desc* model, const double* q0,
                       const double* q1, int64_t n, const t4c_request* req, uint8_t* hit, t4o_contact* out,
                       int64_t capacity, int n_threads, uint64_t* counters6)
{
  return runBatch(scene, model, q0, q1, n, req, hit, out, capacity, n_threads, counters6, 1);
}

int64_t t4o_lvs_expand(const double* traj, int64_t n_traj, int32_t n_wp, int32_t n_q, double lvs, int cast, double* q0,
                       double* q1, int64_t* item_traj, int32_t* item_step, int64_t capacity)
{
  int64_t n = 0;
  auto emit = [&](int64_t t, int step, const double* a, const double* b) {
    if (n < capacity)
    {
      if (q0)
        std::memcpy(q0 + n * n_q, a, sizeof(double) * (size_t)n_q);
      if (q1 && b)
        std::memcpy(q1 + n * n_q, b, sizeof(double) * (size_t)n_q);
      if (item_traj)
        item_traj[n] = t;
      if (item_step)
        item_step[n] = step;
    }
    ++n;
  };
  std::vector<double> sa((size_t)n_q), sb((size_t)n_q);
  for (int64_t t = 0; t < n_traj; ++t)
  {
    const double* tr = traj + t * n_wp * n_q;
    int last = cast ? n_wp - 1 : n_wp;
    for (int s = 0; s < last; ++s)
    {
      const double* a = tr + (int64_t)s * n_q;
      const double* b = tr + (int64_t)(s + 1) * n_q;
      double dist = -1;
      if (s < n_wp - 1)
      {
        double acc = 0;
        for (int k = 0; k < n_q; ++k)
          acc += (b[k] - a[k]) * (b[k] - a[k]);
        dist = std::sqrt(acc);
      }
      if (dist > 0 && dist > lvs)
      {
        int64_t cnt = (int64_t)std::ceil(dist / lvs) + 1;
        for (int64_t i = 0; i < cnt - 1; ++i)
        {
          for (int k = 0; k < n_q; ++k)
          {
            linspaced(cnt, a[k], b[k], i, sa[k]);
            linspaced(cnt, a[k], b[k], i + 1, sb[k]);
          }
          emit(t, s, sa.data(), cast ? sb.data() : nullptr);
        }
      }
      else
        emit(t, s, a, cast ? b : nullptr);
    }
  }
  return n;
}

int t4o_pair_distance(const t4c_shape* sa, const double* pose_a, const t4c_shape* sb, const double* pose_b,
                      const double* vertices, double threshold, int calc_pen, int calc_dist, double* out_dist,
                      double* out_normal_on_b, double* out_point_on_b)
{
  auto mk = [&](const t4c_shape* sd, std::vector<V3>& store) {
    Convex c;
    c.type = sd->type;
    switch (sd->type)
    {
      case T4C_SPHERE:
        c.r = sd->params[0];
        break;
      case T4C_BOX:
        c.he = V3(sd->params[0], sd->params[1], sd->params[2]);
        break;
      case T4C_CYLINDER:
      case T4C_CONE:
      case T4C_CAPSULE:
        c.r = sd->params[0];
        c.h = sd->params[1];
        break;
      default:
      {
        int nv = sd->type == T4C_TRIANGLE ? 3 : sd->vertex_count;
        store.resize((size_t)nv);
        for (int i = 0; i < nv; ++i)
          store[i] = V3(vertices[3 * (sd->vertex_offset + i)], vertices[3 * (sd->vertex_offset + i) + 1],
                        vertices[3 * (sd->vertex_offset + i) + 2]);
        c.verts = store.data();
        c.nverts = nv;
      }
    }
    return c;
  };
  std::vector<V3> va, vb;
  Convex A = mk(sa, va), B = mk(sb, vb);
  PairRequest req;
  req.calculate_penetration = calc_pen != 0;
  req.calculate_distance = calc_dist != 0;
  PairResult pr;
  Counters cnt;
  collideConvexConvex(A, Xf::from12(pose_a), B, Xf::from12(pose_b), threshold, req, pr, cnt);
  if (!pr.has)
    return 0;
  *out_dist = pr.distance;
  for (int k = 0; k < 3; ++k)
  {
    out_normal_on_b[k] = pr.normalOnB[k];
    out_point_on_b[k] = pr.pointOnB[k];
  }
  return 1;
}
}
