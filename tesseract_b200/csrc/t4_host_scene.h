/*
 * t4_host_scene.h -- host-side scene compiler: t4c_scene_desc -> flat device records.
 *
 * Mirrors what the reference does once per manager (cold path):
 *   createShapePrimitive family              tesseract_collision/src/bullet/bullet_utils.cpp:59-365
 *   updateCollisionObjectFilters             bullet_utils.h:900-917 (active = kinematic, static x static never tested)
 *   needsCollisionCheck / isContactAllowed   bullet_utils.h:381-386 ; core/common.h:79-104
 *   CollisionMarginData (default + pairs)    tesseract_common/collision_margin_data.h:66-270
 * and replaces btDbvtBroadphase's static tree by a linear BVH (Morton order, split at the highest differing bit)
 * over the world boxes of all STATIC convex shapes, built once per scene revision and uploaded to the GPU.
 */
#ifndef T4_HOST_SCENE_H_
#define T4_HOST_SCENE_H_

#include <algorithm>
#include <cmath>
#include <cstring>
#include <memory>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "t4_kernels.h"

namespace t4
{
struct HostLink
{
  std::string name;
  int type_id = 0;
  bool enabled = true;
  bool active = false;
  double pose[12];
};

inline uint64_t hostPairKey(int a, int b)
{
  if (a > b)
    std::swap(a, b);
  return ((uint64_t)(uint32_t)a << 32) | (uint32_t)b;
}

struct CompiledScene
{
  std::vector<DShape> shapes;
  std::vector<DStaticW> statics;
  std::vector<double> shape_d;    // [shape][16]
  std::vector<double> statics_d;  // [shape][12]
  std::vector<f4> verts;
  std::vector<uint32_t> sup_cells;  // hull support tables: per hull 6*R*R+1 CSR offsets into sup_cand
  std::vector<uint16_t> sup_cand;
  std::vector<DBvhNode> bvh;
  std::vector<int32_t> active_shapes;
  // kinematic-kinematic broadphase (a9): per kinematic link a child tree over its shapes in the LINK frame, and one item
  // per (child shape of the smaller link, other link) of every admissible link pair
  std::vector<int32_t> aa_pairs;   // [3 * items]: shape, active ordinal of the other link, 1 if the shape's link is the higher one
  std::vector<DBvhNode> link_bvh;  // child trees of all kinematic links, boxes in the link frame
  std::vector<int32_t> link_root;  // [n_alinks] root node in link_bvh, -1: the link has no shapes
  std::vector<int32_t> link_group; // [n_links] dense id of enabled links with more than one shape (CLOSEST pruning), else -1
  int n_groups = 0;
  std::vector<uint32_t> acm_bits;
  int acm_words = 0;
  std::vector<double> pair_margin;  // empty: default only
  std::vector<int32_t> alinks;     // link index of every active link, ascending
  std::vector<int32_t> link_alink; // per link: active ordinal or -1
  float default_margin = 0, max_margin = 0;
  double default_margin_d = 0, max_margin_d = 0;
  bool cast_unsupported = false;
  int bvh_depth = 0;
};

// Geometry (shape records + vertex pool) is immutable once a scene exists and is shared by its clones -- the
// reference shares shared_ptr<const Geometry> between manager clones the same way (types.h:52, bullet_utils.h:183-198);
// adding / removing an object makes a private copy first (t4c.cu).
struct HostGeometry
{
  std::vector<t4c_shape> shapes;
  std::vector<double> vertices;
};

struct HostScene
{
  std::vector<HostLink> links;
  std::shared_ptr<HostGeometry> geom = std::make_shared<HostGeometry>();
  std::unordered_set<uint64_t> acm;
  double default_margin = 0;
  std::unordered_map<uint64_t, double> pair_margins;

  double maxMargin() const
  {
    double m = default_margin;
    for (auto& kv : pair_margins)
      m = std::max(m, kv.second);
    return m;
  }
};

inline uint32_t expandBits10(uint32_t v)
{
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}

struct LbvhBuilder
{
  struct Box
  {
    float mn[3], mx[3];
  };
  const std::vector<Box>& boxes;  // per leaf (sorted order)
  const std::vector<uint32_t>& codes;
  const std::vector<int32_t>& leaf_shape;
  std::vector<DBvhNode>& nodes;
  int max_depth = 0;  // deepest inner node (root = 1)

  // returns child reference (>=0 node, <0 leaf) and its box
  int32_t build(int lo, int hi, Box& out, int depth = 1)
  {
    if (hi - lo == 1)
    {
      out = boxes[lo];
      return ~leaf_shape[lo];
    }
    // split at the highest differing Morton bit (Karras 2012 hierarchy), middle when all codes are equal
    int split = (lo + hi) / 2;
    const uint32_t first = codes[lo], last = codes[hi - 1];
    if (first != last)
    {
      const int prefix = __builtin_clz(first ^ last);
      int s = lo, step = hi - 1 - lo;
      do
      {
        step = (step + 1) >> 1;
        const int ns = s + step;
        if (ns < hi - 1 && __builtin_clz(first ^ codes[ns]) > prefix)
          s = ns;
      } while (step > 1);
      split = s + 1;
    }
    const int idx = (int)nodes.size();
    nodes.emplace_back();
    max_depth = std::max(max_depth, depth);
    Box l, r;
    const int32_t lc = build(lo, split, l, depth + 1);
    const int32_t rc = build(split, hi, r, depth + 1);
    DBvhNode& n = nodes[idx];
    for (int k = 0; k < 3; ++k)
    {
      n.lmn[k] = l.mn[k], n.lmx[k] = l.mx[k], n.rmn[k] = r.mn[k], n.rmx[k] = r.mx[k];
      out.mn[k] = std::min(l.mn[k], r.mn[k]);
      out.mx[k] = std::max(l.mx[k], r.mx[k]);
    }
    n.left = lc, n.right = rc;
    n.pad[0] = n.pad[1] = 0;
    return idx;
  }
};

inline float roundDown(double x) { return std::nextafter((float)x, -INFINITY); }
inline float roundUp(double x) { return std::nextafter((float)x, INFINITY); }

// shape-frame bounding box (centre, half extents incl. margins) of one convex shape
inline void shapeLocalBox(const t4c_shape& s, const double* vertices, double c[3], double h[3])
{
  c[0] = c[1] = c[2] = 0;
  switch (s.type)
  {
    case T4C_SPHERE:
      h[0] = h[1] = h[2] = s.params[0];
      return;
    case T4C_BOX:
      h[0] = s.params[0], h[1] = s.params[1], h[2] = s.params[2];
      return;
    case T4C_CYLINDER:
      h[0] = h[1] = s.params[0], h[2] = s.params[1];
      return;
    case T4C_CONE:
      h[0] = h[1] = s.params[0], h[2] = s.params[1] * 0.5;
      return;
    case T4C_CAPSULE:
      h[0] = h[1] = s.params[0], h[2] = s.params[0] + s.params[1];
      return;
    default:
    {
      const int nv = s.type == T4C_TRIANGLE ? 3 : s.vertex_count;
      double mn[3] = { 1e300, 1e300, 1e300 }, mx[3] = { -1e300, -1e300, -1e300 };
      for (int i = 0; i < nv; ++i)
        for (int k = 0; k < 3; ++k)
        {
          const double v = vertices[3 * (size_t)(s.vertex_offset + i) + k];
          mn[k] = std::min(mn[k], v);
          mx[k] = std::max(mx[k], v);
        }
      if (nv == 0)
        mn[0] = mn[1] = mn[2] = mx[0] = mx[1] = mx[2] = 0;
      for (int k = 0; k < 3; ++k)
      {
        c[k] = 0.5 * (mn[k] + mx[k]);
        h[k] = 0.5 * (mx[k] - mn[k]);
      }
      return;
    }
  }
}

// world-frame data of one STATIC shape (FP64, rounded once): pose, bounding sphere, padded box
inline void staticShapeWorld(const t4c_shape& s, const double* vertices, const double* link_pose, DStaticW& w,
                             double* pose_d12, float* box_mn, float* box_mx)
{
  double c[3], h[3];
  shapeLocalBox(s, vertices, c, h);
  const double rad = s.type == T4C_SPHERE ? s.params[0] : std::sqrt(h[0] * h[0] + h[1] * h[1] + h[2] * h[2]);
  const Xf<double> xw = Xf<double>::from12(link_pose) * Xf<double>::from12(s.pose);
  xw.to12(pose_d12);
  Xf<float> xf = xw.cast<float>();
  xf.to12(w.pose);
  V3<double> mn, mx;
  boxAabb(xw, V3<double>(c[0], c[1], c[2]), V3<double>(h[0], h[1], h[2]), mn, mx);
  const V3<double> cw = xw(V3<double>(c[0], c[1], c[2]));
  w.bs[0] = (float)cw.x, w.bs[1] = (float)cw.y, w.bs[2] = (float)cw.z;
  w.bs[3] = roundUp(rad * (1.0 + 1e-6) + 1e-6);
  const double pad = 1e-6;
  box_mn[0] = roundDown(mn.x - pad), box_mn[1] = roundDown(mn.y - pad), box_mn[2] = roundDown(mn.z - pad);
  box_mx[0] = roundUp(mx.x + pad), box_mx[1] = roundUp(mx.y + pad), box_mx[2] = roundUp(mx.z + pad);
}

inline bool isIdentity12(const double* p)
{
  static const double I[12] = { 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0 };
  return std::memcmp(p, I, sizeof(I)) == 0;
}

/*
 * Support table of one hull (t4_shapes.h coreSupport, SH_HULL).  Directions are bucketed on a cube map: face = major
 * axis and its sign, (u, v) = the other two components divided by |major|, T4_SUP_R x T4_SUP_R cells per face.  A cell
 * lists, in ascending index order, every vertex w that can win the arg-max for SOME direction d of the cell.
 * Necessary condition used (conservative): with u* the winner at the cell centre d_c, a winner w at d = d_c + delta
 * satisfies d.(w - u*) >= 0, and d.(w - u*) <= d_c.(w - u*) + hu |(w-u*)_a| + hv |(w-u*)_b|.  Half widths are padded
 * for the FP32 / FP64 rounding of the device's cell computation, the comparison by eps for the rounding of its dots.
 * Vertices are the FP32-rounded pool values the device reads.
 */
inline void buildSupportTable(const f4* v, int nv, std::vector<uint32_t>& cells, std::vector<uint16_t>& cand)
{
  const int R = T4_SUP_R;
  double vmax = 0;
  for (int i = 0; i < nv; ++i)
    vmax = std::max(vmax, std::max(std::fabs((double)v[i].x), std::max(std::fabs((double)v[i].y), std::fabs((double)v[i].z))));
  const double eps = 1e-9 * vmax + 1e-300;
  const double half = 1.0 / R + 1e-5;
  std::vector<double> p((size_t)nv);
  for (int face = 0; face < 6; ++face)
  {
    const int m = face >> 1, a = m == 0 ? 1 : 0, b = m == 2 ? 1 : 2;  // (u, v) axes as in coreSupport
    const double sgn = (face & 1) ? -1.0 : 1.0;
    for (int iu = 0; iu < R; ++iu)
      for (int iv = 0; iv < R; ++iv)
      {
        double dc[3];
        dc[m] = sgn;
        dc[a] = (iu + 0.5) * (2.0 / R) - 1.0;
        dc[b] = (iv + 0.5) * (2.0 / R) - 1.0;
        int best = 0;
        for (int i = 0; i < nv; ++i)
        {
          p[i] = dc[0] * (double)v[i].x + dc[1] * (double)v[i].y + dc[2] * (double)v[i].z;
          if (p[i] > p[best])
            best = i;
        }
        const double vb[3] = { (double)v[best].x, (double)v[best].y, (double)v[best].z };
        cells.push_back((uint32_t)cand.size());
        for (int i = 0; i < nv; ++i)
        {
          const double vi[3] = { (double)v[i].x, (double)v[i].y, (double)v[i].z };
          if (p[i] - p[best] + half * (std::fabs(vi[a] - vb[a]) + std::fabs(vi[b] - vb[b])) >= -eps)
            cand.push_back((uint16_t)i);
        }
      }
  }
  cells.push_back((uint32_t)cand.size());
}

inline bool compileScene(const HostScene& H, CompiledScene& C, std::string& err)
{
  const HostGeometry& G = *H.geom;
  const int nl = (int)H.links.size(), ns = (int)G.shapes.size();
  C = CompiledScene();
  C.default_margin_d = H.default_margin;
  C.max_margin_d = H.maxMargin();
  C.default_margin = roundUp(H.default_margin);
  C.max_margin = roundUp(H.maxMargin());
  C.link_alink.assign((size_t)nl, -1);
  for (int i = 0; i < nl; ++i)
    if (H.links[i].active)
    {
      C.link_alink[i] = (int)C.alinks.size();
      C.alinks.push_back(i);
    }
  const int na = (int)C.alinks.size();
  C.verts.resize(G.vertices.size() / 3);
  for (size_t i = 0; i < C.verts.size(); ++i)
  {
    C.verts[i].x = (float)G.vertices[3 * i], C.verts[i].y = (float)G.vertices[3 * i + 1];
    C.verts[i].z = (float)G.vertices[3 * i + 2], C.verts[i].w = 0.f;
  }
  C.shapes.resize((size_t)ns);
  C.statics.resize((size_t)ns);
  C.shape_d.assign((size_t)std::max(ns, 1) * 16, 0.0);
  C.statics_d.assign((size_t)std::max(ns, 1) * 12, 0.0);
  std::vector<LbvhBuilder::Box> sbox;
  std::vector<int32_t> sshape;
  std::vector<std::vector<int32_t>> link_active_shapes((size_t)nl);
  for (int i = 0; i < ns; ++i)
  {
    const t4c_shape& s = G.shapes[i];
    if (s.link < 0 || s.link >= nl)
    {
      err = "shape refers to an unknown link";
      return false;
    }
    if ((s.type == T4C_HULL || s.type == T4C_TRIANGLE) &&
        (s.vertex_offset < 0 ||
         (size_t)s.vertex_offset + (size_t)(s.type == T4C_TRIANGLE ? 3 : s.vertex_count) > C.verts.size()))
    {
      err = "shape vertex range out of bounds";
      return false;
    }
    DShape& d = C.shapes[i];
    std::memset(&d, 0, sizeof(d));
    for (int k = 0; k < 12; ++k)
      d.lpose[k] = (float)s.pose[k];
    d.params[0] = (float)s.params[0], d.params[1] = (float)s.params[1], d.params[2] = (float)s.params[2];
    for (int k = 0; k < 3; ++k)
      C.shape_d[(size_t)i * 16 + k] = s.params[k];
    for (int k = 0; k < 12; ++k)
      C.shape_d[(size_t)i * 16 + 4 + k] = s.pose[k];
    double c[3], h[3];
    shapeLocalBox(s, G.vertices.data(), c, h);
    const double rad = s.type == T4C_SPHERE ? s.params[0] : std::sqrt(h[0] * h[0] + h[1] * h[1] + h[2] * h[2]);
    for (int k = 0; k < 3; ++k)
    {
      d.bc[k] = (float)c[k];
      d.bh[k] = roundUp(h[k] + std::fabs(c[k] - (double)(float)c[k]));
    }
    d.bc[3] = roundUp(rad * (1.0 + 1e-6));
    d.type = s.type, d.link = s.link, d.vert_off = s.vertex_offset, d.vert_cnt = s.vertex_count;
    d.shape_id = s.shape_id, d.subshape_id = s.subshape_id;
    d.alink = C.link_alink[s.link];
    d.identity_lpose = isIdentity12(s.pose) ? 1 : 0;
    const HostLink& L = H.links[s.link];
    if (!L.enabled)
      continue;  // disabled objects take no part in any test (bullet_utils.h:381-386)
    if (L.active)
    {
      link_active_shapes[s.link].push_back(i);
      C.active_shapes.push_back(i);
      continue;
    }
    // static shape: world pose + bounding data in FP64, rounded once
    LbvhBuilder::Box b;
    staticShapeWorld(s, G.vertices.data(), L.pose, C.statics[i], &C.statics_d[(size_t)i * 12], b.mn, b.mx);
    sbox.push_back(b);
    sshape.push_back(i);
  }
  // ---- hull support tables (one per distinct vertex range, shared by every shape that uses it)
  {
    std::unordered_map<uint64_t, int32_t> table_of;
    for (int i = 0; i < ns; ++i)
    {
      DShape& d = C.shapes[i];
      d.sup_off = -1;
      if (d.type != SH_HULL || d.vert_cnt < T4_SUP_MIN_VERTS || d.vert_cnt > 65535 || !H.links[d.link].enabled)
        continue;
      const uint64_t key = ((uint64_t)(uint32_t)d.vert_off << 32) | (uint32_t)d.vert_cnt;
      auto it = table_of.find(key);
      if (it == table_of.end())
      {
        if (C.sup_cand.size() > ((size_t)256 << 20))
          continue;  // table budget exhausted: remaining hulls keep the plain scan
        it = table_of.emplace(key, (int32_t)C.sup_cells.size()).first;
        buildSupportTable(C.verts.data() + d.vert_off, d.vert_cnt, C.sup_cells, C.sup_cand);
      }
      d.sup_off = it->second;
    }
  }
  // ---- LBVH over the static shapes
  const int nst = (int)sbox.size();
  if (nst > 0)
  {
    float smn[3] = { 1e30f, 1e30f, 1e30f }, smx[3] = { -1e30f, -1e30f, -1e30f };
    for (auto& b : sbox)
      for (int k = 0; k < 3; ++k)
      {
        smn[k] = std::min(smn[k], 0.5f * (b.mn[k] + b.mx[k]));
        smx[k] = std::max(smx[k], 0.5f * (b.mn[k] + b.mx[k]));
      }
    std::vector<uint32_t> code((size_t)nst);
    std::vector<int32_t> order((size_t)nst);
    for (int i = 0; i < nst; ++i)
    {
      uint32_t m = 0;
      for (int k = 0; k < 3; ++k)
      {
        const float ext = std::max(smx[k] - smn[k], 1e-9f);
        float t = (0.5f * (sbox[i].mn[k] + sbox[i].mx[k]) - smn[k]) / ext;
        t = std::min(std::max(t * 1024.f, 0.f), 1023.f);
        m |= expandBits10((uint32_t)t) << (2 - k);
      }
      code[i] = m;
      order[i] = i;
    }
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return code[a] < code[b]; });
    std::vector<LbvhBuilder::Box> boxes((size_t)nst);
    std::vector<uint32_t> codes((size_t)nst);
    std::vector<int32_t> leaf((size_t)nst);
    for (int i = 0; i < nst; ++i)
    {
      boxes[i] = sbox[order[i]];
      codes[i] = code[order[i]];
      leaf[i] = sshape[order[i]];
    }
    C.bvh.reserve((size_t)nst);
    LbvhBuilder B{ boxes, codes, leaf, C.bvh, 0 };
    if (nst == 1)
    {
      DBvhNode n;
      std::memset(&n, 0, sizeof(n));
      for (int k = 0; k < 3; ++k)
      {
        n.lmn[k] = boxes[0].mn[k], n.lmx[k] = boxes[0].mx[k];
        n.rmn[k] = 1e30f, n.rmx[k] = -1e30f;  // empty box: never overlaps
      }
      n.left = ~leaf[0];
      n.right = ~leaf[0];
      C.bvh.push_back(n);
    }
    else
    {
      LbvhBuilder::Box root;
      B.build(0, nst, root);
      // 30 Morton bits + log2(longest run of equal codes): 50 for a million coincident shapes.  The device walk keeps
      // one pending sibling per level in a T4_BVH_STACK-entry stack; a deeper tree is refused here, never truncated
      // silently on the device.
      if (B.max_depth > T4_BVH_STACK - 1)
      {
        err = "environment BVH too deep for the device traversal stack (" + std::to_string(B.max_depth) + " levels)";
        return false;
      }
      C.bvh_depth = B.max_depth;
    }
  }
  // ---- ACM bit-matrix [n_alinks][n_links]: bit set = pair never tested
  C.acm_words = (nl + 31) / 32;
  C.acm_bits.assign((size_t)std::max(na, 1) * (size_t)std::max(C.acm_words, 1), 0u);
  for (uint64_t key : H.acm)
  {
    const int a = (int)(key >> 32), b = (int)(key & 0xffffffffu);
    if (a >= nl || b >= nl)
      continue;
    if (C.link_alink[a] >= 0)
      C.acm_bits[(size_t)C.link_alink[a] * C.acm_words + (b >> 5)] |= 1u << (b & 31);
    if (C.link_alink[b] >= 0)
      C.acm_bits[(size_t)C.link_alink[b] * C.acm_words + (a >> 5)] |= 1u << (a & 31);
  }
  // ---- per-pair margins (dense table only when overrides exist)
  if (!H.pair_margins.empty())
  {
    if ((size_t)na * (size_t)nl > (size_t)256 << 20)
    {
      err = "pair-margin table too large";
      return false;
    }
    C.pair_margin.assign((size_t)std::max(na, 1) * (size_t)nl, H.default_margin);
    for (auto& kv : H.pair_margins)
    {
      const int a = (int)(kv.first >> 32), b = (int)(kv.first & 0xffffffffu);
      if (a >= nl || b >= nl)
        continue;
      if (C.link_alink[a] >= 0)
        C.pair_margin[(size_t)C.link_alink[a] * nl + b] = kv.second;
      if (C.link_alink[b] >= 0)
        C.pair_margin[(size_t)C.link_alink[b] * nl + a] = kv.second;
    }
  }
  // ---- links that are compounds of several shapes (octomaps, meshes): the keys CLOSEST pruning works on
  {
    std::vector<int> count((size_t)std::max(nl, 1), 0);
    for (int i = 0; i < ns; ++i)
      if (H.links[G.shapes[i].link].enabled)
        ++count[G.shapes[i].link];
    C.link_group.assign((size_t)std::max(nl, 1), -1);
    for (int l = 0; l < nl; ++l)
      if (count[l] > 1)
        C.link_group[l] = C.n_groups++;
  }
  // ---- child trees of the kinematic links (TesseractCompoundCollisionAlgorithm walks the compound's btDbvt with the
  //      other shape's box, tesseract_compound_collision_algorithm.cpp:323-338; compound x compound is a dual-tree walk,
  //      tesseract_compound_compound_collision_algorithm.cpp:256-325).  Boxes are the children's boxes in the LINK frame.
  C.link_root.assign((size_t)std::max(na, 1), -1);
  for (int a = 0; a < na; ++a)
  {
    const std::vector<int32_t>& kids = link_active_shapes[C.alinks[a]];
    if (kids.empty())
      continue;
    std::vector<LbvhBuilder::Box> kb(kids.size());
    float lo[3] = { 1e30f, 1e30f, 1e30f }, hi[3] = { -1e30f, -1e30f, -1e30f };
    for (size_t k = 0; k < kids.size(); ++k)
    {
      const t4c_shape& sh = G.shapes[kids[k]];
      double c[3], h[3];
      shapeLocalBox(sh, G.vertices.data(), c, h);
      V3<double> mn, mx;
      boxAabb(Xf<double>::from12(sh.pose), V3<double>(c[0], c[1], c[2]), V3<double>(h[0], h[1], h[2]), mn, mx);
      const double pad = 1e-6;
      kb[k].mn[0] = roundDown(mn.x - pad), kb[k].mn[1] = roundDown(mn.y - pad), kb[k].mn[2] = roundDown(mn.z - pad);
      kb[k].mx[0] = roundUp(mx.x + pad), kb[k].mx[1] = roundUp(mx.y + pad), kb[k].mx[2] = roundUp(mx.z + pad);
      for (int d = 0; d < 3; ++d)
      {
        lo[d] = std::min(lo[d], 0.5f * (kb[k].mn[d] + kb[k].mx[d]));
        hi[d] = std::max(hi[d], 0.5f * (kb[k].mn[d] + kb[k].mx[d]));
      }
    }
    std::vector<uint32_t> code(kids.size());
    std::vector<int32_t> order(kids.size());
    for (size_t k = 0; k < kids.size(); ++k)
    {
      uint32_t m = 0;
      for (int d = 0; d < 3; ++d)
      {
        const float ext = std::max(hi[d] - lo[d], 1e-9f);
        float t = (0.5f * (kb[k].mn[d] + kb[k].mx[d]) - lo[d]) / ext;
        t = std::min(std::max(t * 1024.f, 0.f), 1023.f);
        m |= expandBits10((uint32_t)t) << (2 - d);
      }
      code[k] = m;
      order[k] = (int32_t)k;
    }
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return code[x] < code[y]; });
    std::vector<LbvhBuilder::Box> boxes(kids.size());
    std::vector<uint32_t> codes(kids.size());
    std::vector<int32_t> leaf(kids.size());
    for (size_t k = 0; k < kids.size(); ++k)
    {
      boxes[k] = kb[order[k]];
      codes[k] = code[order[k]];
      leaf[k] = kids[order[k]];
    }
    C.link_root[a] = (int32_t)C.link_bvh.size();
    if (kids.size() == 1)
    {
      DBvhNode n;
      std::memset(&n, 0, sizeof(n));
      for (int d = 0; d < 3; ++d)
      {
        n.lmn[d] = boxes[0].mn[d], n.lmx[d] = boxes[0].mx[d];
        n.rmn[d] = 1e30f, n.rmx[d] = -1e30f;  // empty box: never overlaps
      }
      n.left = ~leaf[0];
      n.right = ~leaf[0];
      C.link_bvh.push_back(n);
      continue;
    }
    // the builder numbers nodes from the size of the array it appends to: child references stay valid in the shared array
    LbvhBuilder B{ boxes, codes, leaf, C.link_bvh, 0 };
    LbvhBuilder::Box root;
    B.build(0, (int)kids.size(), root);
    if (B.max_depth > T4_BVH_STACK - 1)
    {
      err = "child tree of link '" + H.links[C.alinks[a]].name + "' too deep for the device traversal stack";
      return false;
    }
  }
  // ---- kinematic x kinematic items (links enabled, different, not allowed by the ACM): every child of the link with
  //      FEWER shapes walks the child tree of the other link.  No cap on the number of shape pairs.
  for (int i = 0; i < na; ++i)
    for (int j = i + 1; j < na; ++j)
    {
      const int la = C.alinks[i], lb = C.alinks[j];
      if (link_active_shapes[la].empty() || link_active_shapes[lb].empty())
        continue;
      if (H.acm.count(hostPairKey(la, lb)))
        continue;
      const bool walk_j = link_active_shapes[la].size() <= link_active_shapes[lb].size();
      const std::vector<int32_t>& kids = walk_j ? link_active_shapes[la] : link_active_shapes[lb];
      for (int sh : kids)
      {
        C.aa_pairs.push_back(sh);
        C.aa_pairs.push_back(walk_j ? j : i);
        C.aa_pairs.push_back(walk_j ? 0 : 1);  // candidates are reported as (shape of the lower link, shape of the higher)
      }
    }
  return true;
}

// Refit of the environment BVH after static links moved (the topology stays, the boxes follow): recompute the world data
// of the shapes of the moved links and every node box bottom-up.  Nodes are numbered in pre-order (a parent before its
// children), so one sweep from the last node to the first sees every child before its parent.  The tree stays correct
// whatever the motion; it only gets looser when objects travel far from where they were built -- a scene edit that adds
// or removes shapes rebuilds it (compileScene).
inline void refitStaticLinks(const HostScene& H, CompiledScene& C, const std::vector<char>& link_moved)
{
  const HostGeometry& G = *H.geom;
  const int ns = (int)G.shapes.size();
  std::vector<LbvhBuilder::Box> box((size_t)ns);
  for (int i = 0; i < ns; ++i)
  {
    const t4c_shape& s = G.shapes[i];
    const HostLink& L = H.links[s.link];
    if (!L.enabled || L.active)
      continue;
    if (link_moved[s.link])
      staticShapeWorld(s, G.vertices.data(), L.pose, C.statics[i], &C.statics_d[(size_t)i * 12], box[i].mn, box[i].mx);
    else
    {
      // unchanged: its box is rebuilt from the stored world pose (cheap and keeps this function free of extra state)
      DStaticW w;
      double pd[12];
      staticShapeWorld(s, G.vertices.data(), L.pose, w, pd, box[i].mn, box[i].mx);
    }
  }
  for (int i = (int)C.bvh.size() - 1; i >= 0; --i)
  {
    DBvhNode& n = C.bvh[i];
    for (int side = 0; side < 2; ++side)
    {
      const int child = side ? n.right : n.left;
      float* mn = side ? n.rmn : n.lmn;
      float* mx = side ? n.rmx : n.lmx;
      if (side == 1 && C.bvh.size() == 1 && n.right == n.left)
        continue;  // one-leaf tree: the right slot is the empty box
      if (child < 0)
      {
        for (int k = 0; k < 3; ++k)
          mn[k] = box[~child].mn[k], mx[k] = box[~child].mx[k];
      }
      else
      {
        const DBvhNode& c = C.bvh[child];
        for (int k = 0; k < 3; ++k)
        {
          mn[k] = std::min(c.lmn[k], c.rmn[k]);
          mx[k] = std::max(c.lmx[k], c.rmx[k]);
        }
      }
    }
  }
}

}  // namespace t4
#endif
