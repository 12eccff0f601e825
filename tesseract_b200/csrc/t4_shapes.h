/*
 * t4_shapes.h -- device scene records and convex support mappings.
 *
 * Support / margin / AABB conventions follow the reference's shape conversion
 * (tesseract_collision/src/bullet/bullet_utils.cpp:59-365: sphere = point core + margin r, capsule = segment core
 * + margin r, every other shape margin BULLET_MARGIN = 0, bullet_utils.h:63) and CastHullShape
 * (tesseract_collision/include/tesseract_collision/bullet/bullet_utils.h:217-304: support = arg max over the
 * shape at pose 0 and pose 1 of the WITH-margin support, reported margin 0).
 */
#ifndef T4_SHAPES_H_
#define T4_SHAPES_H_

#include "t4_math.h"

// resolution of the hull support tables: 6 cube-map faces x T4_SUP_R x T4_SUP_R cells (even)
#define T4_SUP_R 16
// hulls with fewer vertices than this keep the plain scan
#define T4_SUP_MIN_VERTS 24

namespace t4
{
enum
{
  SH_SPHERE = 0,
  SH_BOX = 1,
  SH_CYLINDER = 2,
  SH_CONE = 3,
  SH_CAPSULE = 4,
  SH_HULL = 5,
  SH_TRIANGLE = 6
};

// One convex shape (128 B).  Mesh faces / octree leaves are already expanded into one record each.
struct alignas(16) DShape
{
  float lpose[12];  // pose in the link frame, 3x4 row-major
  float params[4];  // sphere: r | box: hx hy hz | cylinder: r, half length | cone: r, height | capsule: r, half length
  float bc[4];      // shape-frame bounding box centre xyz, [3] = bounding-sphere radius about that centre
  float bh[3];      // shape-frame bounding box half extents (margins included)
  int32_t sup_off;  // hulls: first cell of the support table in KParams::sup_cells, or -1 (brute-force scan)
  int32_t type, link, vert_off, vert_cnt;
  int32_t shape_id, subshape_id, alink, identity_lpose;  // alink: active-link ordinal or -1 (static)
};
static_assert(sizeof(DShape) == 128, "DShape layout");

// world-frame data of a STATIC shape, precomputed on the host in FP64 and rounded once
struct alignas(16) DStaticW
{
  float pose[12];  // world pose of the shape
  float bs[4];     // world bounding-sphere centre + radius
};
static_assert(sizeof(DStaticW) == 64, "DStaticW layout");

// binary BVH node over static shapes: both children's boxes in the parent (one 64 B fetch tests two boxes)
struct alignas(16) DBvhNode
{
  float lmn[3], lmx[3], rmn[3], rmx[3];
  int32_t left, right;  // >= 0: node index; < 0: leaf, shape = ~index
  int32_t pad[2];
};
static_assert(sizeof(DBvhNode) == 64, "DBvhNode layout");

template <class T>
struct ShapeW
{
  int type;
  T r, h;  // radius ; half length (cylinder, capsule) or height (cone)
  V3<T> he;
  const f4* verts;
  int nv;
  const uint32_t* sup_cells;  // hull support table (T4_SUP_R^2 * 6 + 1 CSR offsets into sup_cand) or null
  const uint16_t* sup_cand;
  bool cast;
  Xf<T> x0, x1;  // world pose at t0 (and t1 for cast shapes), recentred by the caller
  Xf<T> t01;     // cast shapes: pose at t1 relative to t0, (T0 L)^-1 (T1 L)  (bullet_cast_bvh_manager.cpp:281-305)
  // hull vertices whose dot product the support mapping has evaluated so far (t4c_stats::support_vertices: the work
  // the roofline figure may claim)
  mutable uint32_t evals;
  T4_HD T innerMargin() const { return (type == SH_SPHERE || type == SH_CAPSULE) ? r : T(0); }
  // margin the detector subtracts analytically.  A cast shape is conv(shape@t0 U shape@t1) built from the WITH-margin
  // support and reports margin 0 (CastHullShape::getMargin, bullet_utils.h:230-235,286): the swept ball is part of
  // the support, exactly as in the reference, so that the GJK iteration (and where it stops) is the same.
  T4_HD T margin() const { return cast ? T(0) : innerMargin(); }
  T4_HD bool polyhedral() const { return type == SH_BOX || type == SH_HULL || type == SH_TRIANGLE; }
  T4_HD bool curvedCore() const { return type == SH_CYLINDER || type == SH_CONE; }
};

template <class T>
T4_HD V3<T> vtx(const f4* verts, int i)
{
  f4 v = verts[i];
  return V3<T>((T)v.x, (T)v.y, (T)v.z);
}

// btConvexShape::localGetSupportVertexWithoutMarginNonVirtual and the virtual cone / capsule variants
template <class T>
T4_HD V3<T> coreSupportInline(const ShapeW<T>& s, const V3<T>& d)
{
  switch (s.type)
  {
    case SH_SPHERE:
      return V3<T>(0, 0, 0);
    case SH_BOX:
      return V3<T>(d.x >= 0 ? s.he.x : -s.he.x, d.y >= 0 ? s.he.y : -s.he.y, d.z >= 0 ? s.he.z : -s.he.z);
    case SH_CYLINDER:
    {
      T sl = t4sqrt(d.x * d.x + d.y * d.y);
      if (sl != T(0))
      {
        T k = s.r / sl;
        return V3<T>(d.x * k, d.y * k, d.z < T(0) ? -s.h : s.h);
      }
      return V3<T>(s.r, 0, d.z < T(0) ? -s.h : s.h);
    }
    case SH_CONE:
    {
      T half = s.h * T(0.5);
      T sin_a = s.r / t4sqrt(s.r * s.r + s.h * s.h);
      if (d.z > d.len() * sin_a)
        return V3<T>(0, 0, half);
      T sl = t4sqrt(d.x * d.x + d.y * d.y);
      if (sl > Eps<T>::eps())
      {
        T k = s.r / sl;
        return V3<T>(d.x * k, d.y * k, -half);
      }
      return V3<T>(0, 0, -half);
    }
    case SH_CAPSULE:
      // segment end with the larger projection, ties -> +z (first candidate wins on strict >)
      return V3<T>(0, 0, (d.z >= T(0)) ? s.h : -s.h);
    case SH_HULL:
    {
      int best = 0;
      T mx = -Eps<T>::large();
      if (s.sup_cells)
      {
        // cube-map cell of d -> the (index-ordered) vertices that can be the maximiser anywhere in that cell; the scan
        // below evaluates the same dot products with the same strict '>' as the full scan: identical result.
        const T ax = t4abs(d.x), ay = t4abs(d.y), az = t4abs(d.z);
        const T m = t4max(ax, t4max(ay, az));
        if (m > T(1e-30) && m < T(1e30))  // 1/m and the dots stay finite; anything else takes the plain scan
        {
          const T inv = T(1) / m;
          int face;
          T u, v;
          if (ax >= ay && ax >= az)
            face = d.x < T(0) ? 1 : 0, u = d.y * inv, v = d.z * inv;
          else if (ay >= az)
            face = d.y < T(0) ? 3 : 2, u = d.x * inv, v = d.z * inv;
          else
            face = d.z < T(0) ? 5 : 4, u = d.x * inv, v = d.y * inv;
          int iu = (int)((u + T(1)) * T(T4_SUP_R / 2));
          int iv = (int)((v + T(1)) * T(T4_SUP_R / 2));
          iu = iu < 0 ? 0 : (iu > T4_SUP_R - 1 ? T4_SUP_R - 1 : iu);
          iv = iv < 0 ? 0 : (iv > T4_SUP_R - 1 ? T4_SUP_R - 1 : iv);
          const int cell = (face * T4_SUP_R + iu) * T4_SUP_R + iv;
          const uint32_t beg = s.sup_cells[cell], end = s.sup_cells[cell + 1];
          s.evals += end - beg;
          for (uint32_t k = beg; k < end; ++k)
          {
            const int i = (int)s.sup_cand[k];
            f4 w = s.verts[i];
            T dt = d.x * (T)w.x + d.y * (T)w.y + d.z * (T)w.z;
            if (dt > mx)
            {
              mx = dt;
              best = i;
            }
          }
          return vtx<T>(s.verts, best);
        }
      }
      s.evals += (uint32_t)s.nv;
      for (int i = 0; i < s.nv; ++i)
      {
        f4 v = s.verts[i];
        T dt = d.x * (T)v.x + d.y * (T)v.y + d.z * (T)v.z;
        if (dt > mx)
        {
          mx = dt;
          best = i;
        }
      }
      return vtx<T>(s.verts, best);
    }
    default:  // SH_TRIANGLE: btVector3::maxAxis over the three projections
    {
      V3<T> a = vtx<T>(s.verts, 0), b = vtx<T>(s.verts, 1), c = vtx<T>(s.verts, 2);
      T d0 = d.dot(a), d1 = d.dot(b), d2 = d.dot(c);
      int ax = d0 < d1 ? (d1 < d2 ? 2 : 1) : (d0 < d2 ? 2 : 0);
      return ax == 0 ? a : (ax == 1 ? b : c);
    }
  }
}

template <class T>
T4_HDN V3<T> coreSupportCall(const ShapeW<T>& s, const V3<T>& d)
{
  return coreSupportInline(s, d);
}
template <class T>
T4_HD V3<T> coreSupport(const ShapeW<T>& s, const V3<T>& d)
{
  if constexpr (sizeof(T) == 8)
    return coreSupportCall(s, d);
  else
    return coreSupportInline(s, d);
}

// support of the shape placed by `x`, world direction d; with_margin adds margin * d/|d|
template <class T>
T4_HD V3<T> supportAt(const ShapeW<T>& s, const Xf<T>& x, const V3<T>& d, T margin)
{
  V3<T> dl = x.rotT(d);
  V3<T> sv = coreSupportInline(s, dl);  // the hot, non-swept path: one inlined copy inside supportWCall
  if (margin != T(0))
    sv += normalizedOrDefault(dl) * margin;
  return x(sv);
}

// World support of the core (with_margin=false, what GJK / EPA use) or of the inflated shape (with_margin=true).
// Cast shapes always carry their inner margin (CastHullShape::localGetSupportingVertex, bullet_utils.h:230-235).
template <class T>
T4_HD V3<T> supportWInline(const ShapeW<T>& s, const V3<T>& d, bool with_margin)
{
  if (!s.cast)
    return supportAt(s, s.x0, d, with_margin ? s.innerMargin() : T(0));
  // CastHullShape::localGetSupportingVertex, evaluated in the frame of pose 0 exactly like the reference
  const T m = s.innerMargin();
  const V3<T> dl = s.x0.rotT(d);
  V3<T> sv0 = coreSupport(s, dl);
  if (m != T(0))
    sv0 += normalizedOrDefault(dl) * m;
  const V3<T> dl1 = s.t01.rotT(dl);
  V3<T> sv1 = coreSupport(s, dl1);
  if (m != T(0))
    sv1 += normalizedOrDefault(dl1) * m;
  sv1 = s.t01(sv1);
  return s.x0((dl.dot(sv0) > dl.dot(sv1)) ? sv0 : sv1);
}

// FP64 (exact pass): a real call, so that the exact kernels stay small enough for the instruction cache (the fully
// inlined GJK kernel was 63k SASS instructions, the EPA kernel 216k; ncu showed 21 "no instruction" stall cycles per
// issued instruction).  FP32 (cull): inlined, the kernel is small and the call overhead costs more than it saves.
template <class T>
T4_HDN V3<T> supportWCall(const ShapeW<T>& s, const V3<T>& d, bool with_margin)
{
  return supportWInline(s, d, with_margin);
}
template <class T>
T4_HD V3<T> supportW(const ShapeW<T>& s, const V3<T>& d, bool with_margin)
{
  if constexpr (sizeof(T) == 8)
    return supportWCall(s, d, with_margin);
  else
    return supportWInline(s, d, with_margin);
}

// world AABB of the shape-frame box (centre c, half h) under pose x
template <class T>
T4_HD void boxAabb(const Xf<T>& x, const V3<T>& c, const V3<T>& h, V3<T>& mn, V3<T>& mx)
{
  V3<T> wc = x(c);
  V3<T> e(t4abs(x.r[0]) * h.x + t4abs(x.r[1]) * h.y + t4abs(x.r[2]) * h.z,
          t4abs(x.r[3]) * h.x + t4abs(x.r[4]) * h.y + t4abs(x.r[5]) * h.z,
          t4abs(x.r[6]) * h.x + t4abs(x.r[7]) * h.y + t4abs(x.r[8]) * h.z);
  mn = wc - e;
  mx = wc + e;
}

}  // namespace t4
#endif
