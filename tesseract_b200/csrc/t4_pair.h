/*
 * t4_pair.h -- one convex-convex query: algorithm dispatch, analytic pairs, GJK + EPA, continuous (cast) data.
 *
 * Dispatch matrix = what TesseractCollisionConfiguration + Bullet's closest-point table select
 * (tesseract_collision/src/bullet/tesseract_collision_configuration.cpp:57-138):
 *   sphere x sphere                 -> analytic (Bullet btSphereSphereCollisionAlgorithm)
 *   sphere x triangle               -> Bullet SphereTriangleDetector
 *   capsule x capsule|sphere        -> tesseract_convex_convex_algorithm.cpp:67-181, 311-399
 *   anything else convex x convex   -> GJK / EPA (tesseract_gjk_pair_detector.cpp:683-1185)
 * Cast (swept) shapes always take the GJK / EPA path (bullet_utils.h:217-304).
 */
#ifndef T4_PAIR_H_
#define T4_PAIR_H_

#include "t4_bullet_detector.h"

namespace t4
{
// Bullet gContactBreakingThreshold (the relative-threshold flag is cleared at bullet_discrete_bvh_manager.cpp:65-66)
#define T4_CONTACT_BREAKING_THRESHOLD 0.02
#define T4_BULLET_SUPPORT_FUNC_TOLERANCE 0.01  // bullet_utils.h:64
#define T4_BULLET_LENGTH_TOLERANCE 0.001      // bullet_utils.h:65
#define T4_BULLET_EPSILON 1e-3                // bullet_utils.h:66

template <class T>
struct PairResult
{
  bool has;
  V3<T> normalOnB;  // world, from B towards A
  V3<T> pointOnB;   // world
  T distance;
  int gjk_iters, epa_iters;
  bool used_epa;
};

template <class T>
T4_HD void makeShapeW(ShapeW<T>& w, const DShape& s, const f4* verts, const uint32_t* sup_cells,
                      const uint16_t* sup_cand, const Xf<T>& x0, const Xf<T>& x1, bool cast)
{
  w.type = s.type;
  w.r = 0, w.h = 0;
  w.nv = 0;
  w.evals = 0;
  w.verts = nullptr;
  w.sup_cells = nullptr;
  w.sup_cand = nullptr;
  switch (s.type)
  {
    case SH_SPHERE:
      w.r = (T)s.params[0];
      break;
    case SH_BOX:
      w.he = V3<T>((T)s.params[0], (T)s.params[1], (T)s.params[2]);
      break;
    case SH_CYLINDER:
    case SH_CONE:
    case SH_CAPSULE:
      w.r = (T)s.params[0];
      w.h = (T)s.params[1];
      break;
    default:
      w.verts = verts + s.vert_off;
      w.nv = s.type == SH_TRIANGLE ? 3 : s.vert_cnt;
      if (s.type == SH_HULL && s.sup_off >= 0 && sup_cells)
      {
        w.sup_cells = sup_cells + s.sup_off;
        w.sup_cand = sup_cand;
      }
      break;
  }
  w.cast = cast;
  w.x0 = x0;
  if (cast)  // x1 / t01 are only read for swept shapes: discrete queries never touch (or store) them
  {
    w.t01 = x0.invTimes(x1);
    w.x1 = x0 * w.t01;  // tfWorld1 as the reference forms it (bullet_utils.h:466)
  }
}

// ---- capsule x capsule|sphere (tesseract_convex_convex_algorithm.cpp:67-181) -------------------------------------
template <class T>
T4_HD void segmentsClosestPoints(V3<T>& ptsVector, V3<T>& offsetA, V3<T>& offsetB, T& tA, T& tB, const V3<T>& tr,
                                 const V3<T>& dirA, T hlenA, const V3<T>& dirB, T hlenB)
{
  const T dAdB = dirA.dot(dirB), dAt = dirA.dot(tr), dBt = dirB.dot(tr);
  const T denom = T(1) - dAdB * dAdB;
  if (denom == T(0))
    tA = 0;
  else
  {
    tA = (dAt - dBt * dAdB) / denom;
    tA = t4min(t4max(tA, -hlenA), hlenA);
  }
  tB = tA * dAdB - dBt;
  if (tB < -hlenB)
  {
    tB = -hlenB;
    tA = t4min(t4max(tB * dAdB + dAt, -hlenA), hlenA);
  }
  else if (tB > hlenB)
  {
    tB = hlenB;
    tA = t4min(t4max(tB * dAdB + dAt, -hlenA), hlenA);
  }
  offsetA = dirA * tA;
  offsetB = dirB * tB;
  ptsVector = tr - offsetA + offsetB;
}

template <class T>
T4_HD void planeSpace1(const V3<T>& n, V3<T>& p)
{
  if (t4abs(n.z) > T(0.70710678118654752440))
  {
    const T k = T(1) / t4sqrt(n.y * n.y + n.z * n.z);
    p = V3<T>(0, -n.z * k, n.y * k);
  }
  else
  {
    const T k = T(1) / t4sqrt(n.x * n.x + n.y * n.y);
    p = V3<T>(-n.y * k, n.x * k, 0);
  }
}

template <class T>
T4_HD T capsuleCapsuleDistance(V3<T>& normalOnB, V3<T>& pointOnB, T lenA, T rA, T lenB, T rB, int axisA, int axisB,
                               const Xf<T>& ta, const Xf<T>& tb, T thr)
{
  const V3<T> dA = ta.col(axisA), dB = tb.col(axisB);
  V3<T> pv, oA, oB;
  T tA, tB;
  segmentsClosestPoints(pv, oA, oB, tA, tB, tb.p - ta.p, dA, lenA, dB, lenB);
  const T distance = pv.len() - rA - rB;
  if (distance > thr)
    return distance;
  const T l2 = pv.len2();
  if (l2 <= Eps<T>::eps() * Eps<T>::eps())
    planeSpace1(dA, normalOnB);
  else
    normalOnB = pv * (T(-1) / t4sqrt(l2));
  pointOnB = tb.p + oB + normalOnB * rB;
  return distance;
}

// ---- Bullet SphereTriangleDetector::collide (sphere centre given in the triangle's frame) ------------------------
template <class T>
T4_HD T segmentSqrDistance(const V3<T>& from, const V3<T>& to, const V3<T>& p, V3<T>& nearest)
{
  V3<T> diff = p - from;
  const V3<T> v = to - from;
  T t = v.dot(diff);
  if (t > T(0))
  {
    const T vv = v.dot(v);
    if (t < vv)
    {
      t /= vv;
      diff -= v * t;
    }
    else
    {
      t = 1;
      diff -= v;
    }
  }
  else
    t = 0;
  nearest = from + v * t;
  return diff.dot(diff);
}

template <class T>
T4_HD bool sphereTriangleCollide(const V3<T>& c, T radius, const V3<T>* tv, V3<T>& point, V3<T>& rn, T& depth, T cbt)
{
  const T rwt = radius + cbt;
  V3<T> normal = (tv[1] - tv[0]).cross(tv[2] - tv[0]);
  const T l2 = normal.len2();
  bool hasContact = false;
  V3<T> cp;
  if (l2 >= Eps<T>::eps() * Eps<T>::eps())
  {
    normal = normal * (T(1) / t4sqrt(l2));
    T dfp = (c - tv[0]).dot(normal);
    if (dfp < T(0))
    {
      dfp = -dfp;
      normal = -normal;
    }
    if (dfp < rwt)
    {
      const V3<T> e1 = tv[1] - tv[0], e2 = tv[2] - tv[1], e3 = tv[0] - tv[2];
      const T r1 = e1.cross(normal).dot(c - tv[0]);
      const T r2 = e2.cross(normal).dot(c - tv[1]);
      const T r3 = e3.cross(normal).dot(c - tv[2]);
      if ((r1 > 0 && r2 > 0 && r3 > 0) || (r1 <= 0 && r2 <= 0 && r3 <= 0))
      {
        hasContact = true;
        cp = c - normal * dfp;
      }
      else
      {
        T mind = rwt * rwt;
        V3<T> ne;
        for (int i = 0; i < 3; ++i)
        {
          const T ds = segmentSqrDistance(tv[i], tv[(i + 1) % 3], c, ne);
          if (ds < mind)
          {
            mind = ds;
            hasContact = true;
            cp = ne;
          }
        }
      }
    }
  }
  if (hasContact)
  {
    const V3<T> ctc = c - cp;
    const T ds = ctc.len2();
    if (ds < rwt * rwt)
    {
      if (ds > Eps<T>::eps())
      {
        const T d = t4sqrt(ds);
        rn = ctc * (T(1) / d);
        point = cp;
        depth = -(radius - d);
      }
      else
      {
        rn = normal;
        point = cp;
        depth = -radius;
      }
      return true;
    }
  }
  return false;
}

/*
 * Analytic members of the dispatch matrix (sphere-sphere, sphere-triangle, capsule pairs).  Returns false when the
 * pair belongs to the GJK / EPA detector instead.  A and B carry WORLD poses; threshold = max contact margin.
 */
template <class T>
T4_HDN bool analyticPairQuery(const ShapeW<T>& A0, const ShapeW<T>& B0, T threshold, PairResult<T>& out)
{
  out.has = false;
  out.gjk_iters = 0;
  out.epa_iters = 0;
  out.used_epa = false;
  const bool plain = !A0.cast && !B0.cast;
  if (!plain)
    return false;
  if (A0.type == SH_SPHERE && B0.type == SH_SPHERE)
  {
    const V3<T> diff = A0.x0.p - B0.x0.p;
    const T len = diff.len();
    if (len > A0.r + B0.r + threshold)
      return true;
    V3<T> n(1, 0, 0);
    if (len > Eps<T>::eps())
      n = diff * (T(1) / len);
    out.has = true;
    out.normalOnB = n;
    out.pointOnB = B0.x0.p + n * B0.r;
    out.distance = len - (A0.r + B0.r);
    return true;
  }
  if (A0.type == SH_SPHERE && B0.type == SH_TRIANGLE)
  {
    V3<T> tv[3] = { vtx<T>(B0.verts, 0), vtx<T>(B0.verts, 1), vtx<T>(B0.verts, 2) };
    V3<T> point, normal;
    T depth = 0;
    if (sphereTriangleCollide(B0.x0.invApply(A0.x0.p), A0.r, tv, point, normal, depth,
                              T(T4_CONTACT_BREAKING_THRESHOLD) + threshold))
    {
      out.has = true;
      out.normalOnB = B0.x0.rot(normal);
      out.pointOnB = B0.x0(point);
      out.distance = depth;
    }
    return true;
  }
  if (A0.type == SH_TRIANGLE && B0.type == SH_SPHERE)
  {
    V3<T> tv[3] = { vtx<T>(A0.verts, 0), vtx<T>(A0.verts, 1), vtx<T>(A0.verts, 2) };
    V3<T> point, normal;
    T depth = 0;
    if (sphereTriangleCollide(A0.x0.invApply(B0.x0.p), B0.r, tv, point, normal, depth,
                              T(T4_CONTACT_BREAKING_THRESHOLD) + threshold))
    {
      const V3<T> nt = A0.x0.rot(normal);
      out.has = true;
      out.normalOnB = -nt;
      out.pointOnB = A0.x0(point) + nt * depth;
      out.distance = depth;
    }
    return true;
  }
  if ((A0.type == SH_CAPSULE && (B0.type == SH_CAPSULE || B0.type == SH_SPHERE)) ||
      (A0.type == SH_SPHERE && B0.type == SH_CAPSULE))
  {
    const T thr = T(T4_CONTACT_BREAKING_THRESHOLD) + threshold;
    V3<T> n, p;
    const T lenA = A0.type == SH_CAPSULE ? A0.h : T(0);
    const T lenB = B0.type == SH_CAPSULE ? B0.h : T(0);
    const T dist = capsuleCapsuleDistance(n, p, lenA, A0.r, lenB, B0.r, A0.type == SH_CAPSULE ? 2 : 1,
                                          B0.type == SH_CAPSULE ? 2 : 1, A0.x0, B0.x0, thr);
    if (dist < thr)
    {
      out.has = true;
      out.normalOnB = n;
      out.pointOnB = p;
      out.distance = dist;
    }
    return true;
  }
  return false;
}

template <class T>
T4_HD T detectorMaxDist2(const ShapeW<T>& A0, const ShapeW<T>& B0, T threshold)
{
  const T maxDist = A0.margin() + B0.margin() + T(T4_CONTACT_BREAKING_THRESHOLD) + threshold;
  return maxDist * maxDist;
}

template <class T>
T4_HD void copyOut(const bt::DetectorOut<T>& d, PairResult<T>& out)
{
  out.has = d.has;
  out.normalOnB = d.normalOnB;
  out.pointOnB = d.pointOnB;
  out.distance = d.distance;
  out.gjk_iters = d.gjk_iters;
  out.epa_iters = d.epa_iters;
  out.used_epa = d.used_epa;
}

// ---- continuous collision data (bullet_utils.h:306-349 GetAverageSupport, :452-525 calculateContinuousData) ------
template <class T>
T4_HDN void averageSupport(const ShapeW<T>& s, const V3<T>& n, T& outsupport, V3<T>& outpt)
{
  if (s.polyhedral())
  {
    V3<T> sum(0, 0, 0);
    T cnt = 0;
    T mx = -1000;
    const int npts = s.type == SH_BOX ? 8 : s.nv;
    for (int i = 0; i < npts; ++i)
    {
      V3<T> pt;
      if (s.type == SH_BOX)
        pt = V3<T>((i & 1) ? -s.he.x : s.he.x, (i & 2) ? -s.he.y : s.he.y, (i & 4) ? -s.he.z : s.he.z);
      else
        pt = vtx<T>(s.verts, i);
      const T sup = pt.dot(n);
      if (sup > mx + T(T4_BULLET_EPSILON))
      {
        cnt = 1;
        sum = pt;
        mx = sup;
      }
      else if (sup < mx - T(T4_BULLET_EPSILON))
      {
      }
      else
      {
        cnt += 1;
        sum += pt;
      }
    }
    outsupport = mx;
    outpt = sum * (T(1) / cnt);
  }
  else
  {
    V3<T> sv = coreSupport(s, n);
    const T m = s.innerMargin();
    if (m != T(0))
      sv += normalizedOrDefault(n) * m;
    outpt = sv;
    outsupport = n.dot(sv);
  }
}

// s: the UN-swept shape with x0 / x1 = its world pose at t0 / t1; link0_inv: inverse of the link pose at t0
template <class T>
T4_HDN void continuousData(const ShapeW<T>& s, const Xf<T>& link0, const V3<T>& pt_world, const V3<T>& normal_world,
                          T& cc_time, int& cc_type, V3<T>& point_local)
{
  const V3<T> nl0 = s.x0.rotT(normal_world), nl1 = s.x1.rotT(normal_world);
  V3<T> pl0, pl1;
  T ls0, ls1;
  averageSupport(s, nl0, ls0, pl0);
  averageSupport(s, nl1, ls1, pl1);
  const V3<T> pw0 = s.x0(pl0), pw1 = s.x1(pl1);
  const T sup0 = normal_world.dot(pw0), sup1 = normal_world.dot(pw1);
  if (sup0 - sup1 > T(T4_BULLET_SUPPORT_FUNC_TOLERANCE))
  {
    cc_time = 0;
    cc_type = 1;  // T4C_CC_TIME0
  }
  else if (sup1 - sup0 > T(T4_BULLET_SUPPORT_FUNC_TOLERANCE))
  {
    cc_time = 1;
    cc_type = 2;  // T4C_CC_TIME1
  }
  else
  {
    const T l0c = (pt_world - pw0).len(), l1c = (pt_world - pw1).len();
    point_local = link0.invApply(s.x0((pl0 + pl1) * T(0.5)));
    cc_type = 3;  // T4C_CC_BETWEEN
    cc_time = (l0c + l1c < T(T4_BULLET_LENGTH_TOLERANCE)) ? T(0.5) : l0c / (l0c + l1c);
  }
}

}  // namespace t4
#endif
