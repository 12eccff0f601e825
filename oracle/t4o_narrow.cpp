/*
 * t4o_narrow.cpp -- CPU ORACLE narrowphase (test infrastructure, not product code).
 *
 * Restates, in FP64:
 *   - shape support / margin / AABB rules           Bullet btConvexShape.cpp, btBoxShape.h, btCylinderShape.h,
 *                                                   btConeShape.cpp, btCapsuleShape.h (un-vendored)
 *   - CastHullShape                                 tesseract_collision/include/tesseract_collision/bullet/bullet_utils.h:217-251
 *   - btVoronoiSimplexSolver                        Bullet; used at tesseract_gjk_pair_detector.cpp:836-975
 *   - btGjkEpaSolver2 (GJK, EPA, Penetration, Distance) + btGjkEpaPenetrationDepthSolver::calcPenDepth
 *                                                   Bullet; called at tesseract_gjk_pair_detector.cpp:1024-1032
 *   - TesseractGjkPairDetector::getClosestPointsNonVirtual
 *                                                   tesseract_collision/src/bullet/tesseract_gjk_pair_detector.cpp:683-1185
 *                                                   (libccd-style simplex helpers :266-676)
 *   - TesseractConvexConvexAlgorithm::processCollision (capsule analytic + GJK dispatch)
 *                                                   tesseract_collision/src/bullet/tesseract_convex_convex_algorithm.cpp:67-181,289-444
 *   - btSphereSphereCollisionAlgorithm, SphereTriangleDetector   Bullet
 * Constants are the BT_USE_DOUBLE_PRECISION ones (the reference's intended build).
 */
#include "t4o.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>

namespace t4o
{
// ------------------------------------------------------------------------------------------------
// Shapes
// ------------------------------------------------------------------------------------------------
double Convex::margin() const
{
  if (is_cast)
    return 0.0;  // CastHullShape::getMargin (bullet_utils.h:286)
  if (type == T4C_SPHERE || type == T4C_CAPSULE)
    return r;
  return 0.0;  // BULLET_MARGIN (bullet_utils.h:63)
}

static V3 normalizedOrDefault(const V3& d)
{
  V3 n = d;
  if (n.length2() < SIMD_EPS * SIMD_EPS)
    n = V3(-1, -1, -1);
  return n / n.length();
}

static V3 primitiveSupportNoMargin(const Convex& s, const V3& d)
{
  switch (s.type)
  {
    case T4C_SPHERE:
      return V3(0, 0, 0);
    case T4C_BOX:
      return V3(d.x >= 0 ? s.he.x : -s.he.x, d.y >= 0 ? s.he.y : -s.he.y, d.z >= 0 ? s.he.z : -s.he.z);
    case T4C_CYLINDER:
    {
      double sl = std::sqrt(d.x * d.x + d.y * d.y);
      if (sl != 0.0)
      {
        double k = s.r / sl;
        return V3(d.x * k, d.y * k, d.z < 0.0 ? -s.h : s.h);
      }
      return V3(s.r, 0, d.z < 0.0 ? -s.h : s.h);
    }
    case T4C_CONE:
    {
      double half = s.h * 0.5;
      double sin_a = s.r / std::sqrt(s.r * s.r + s.h * s.h);
      if (d.z > d.length() * sin_a)
        return V3(0, 0, half);
      double sl = std::sqrt(d.x * d.x + d.y * d.y);
      if (sl > SIMD_EPS)
      {
        double k = s.r / sl;
        return V3(d.x * k, d.y * k, -half);
      }
      return V3(0, 0, -half);
    }
    case T4C_CAPSULE:
    {
      V3 vec = d;
      double l2 = vec.length2();
      if (l2 < SIMD_EPS * SIMD_EPS)
        vec = V3(1, 0, 0);
      else
        vec *= 1.0 / std::sqrt(l2);
      V3 sup(0, 0, 0);
      double max_dot = -LARGE;
      V3 p(0, 0, s.h);
      double nd = vec.dot(p);
      if (nd > max_dot)
      {
        max_dot = nd;
        sup = p;
      }
      p = V3(0, 0, -s.h);
      nd = vec.dot(p);
      if (nd > max_dot)
      {
        max_dot = nd;
        sup = p;
      }
      return sup;
    }
    case T4C_HULL:
    {
      int best = 0;
      double max_dot = -LARGE;
      for (int i = 0; i < s.nverts; ++i)
      {
        double dt = d.dot(s.verts[i]);
        if (dt > max_dot)
        {
          max_dot = dt;
          best = i;
        }
      }
      return s.verts[best];
    }
    case T4C_TRIANGLE:
    {
      double d0 = d.dot(s.verts[0]), d1 = d.dot(s.verts[1]), d2 = d.dot(s.verts[2]);
      int ax = d0 < d1 ? (d1 < d2 ? 2 : 1) : (d0 < d2 ? 2 : 0);  // btVector3::maxAxis
      return s.verts[ax];
    }
  }
  return V3();
}

// btConvexInternalShape::localGetSupportingVertex (virtual path: direction is NOT normalised for the core)
static V3 virtualSupportWithMargin(const Convex& s, const V3& vec)
{
  V3 sv = primitiveSupportNoMargin(s, vec);
  double m = s.margin();
  if (m != 0.0)
    sv += normalizedOrDefault(vec) * m;
  return sv;
}

V3 Convex::supportNoMargin(const V3& d) const
{
  if (!is_cast)
    return primitiveSupportNoMargin(*this, d);
  // CastHullShape::localGetSupportingVertex (bullet_utils.h:230-235)
  V3 sv0 = virtualSupportWithMargin(*inner, d);
  V3 sv1 = t01(virtualSupportWithMargin(*inner, t01.R.tmul(d)));
  return (d.dot(sv0) > d.dot(sv1)) ? sv0 : sv1;
}

V3 Convex::supportWithMargin(const V3& d) const
{
  V3 n = normalizedOrDefault(d);
  return supportNoMargin(n) + n * margin();
}

static void transformAabb(const V3& half, const V3& center_local, const Xf& t, V3& mn, V3& mx)
{
  M3 ab = t.R.absolute();
  V3 c = t(center_local);
  V3 e(ab.row(0).dot(half), ab.row(1).dot(half), ab.row(2).dot(half));
  mn = c - e;
  mx = c + e;
}

void Convex::aabb(const Xf& t, V3& mn, V3& mx) const
{
  if (is_cast)
  {
    // CastHullShape::getAabb (bullet_utils.h:238-245)
    inner->aabb(t, mn, mx);
    V3 mn1, mx1;
    inner->aabb(t * t01, mn1, mx1);
    for (int i = 0; i < 3; ++i)
    {
      mn[i] = std::min(mn[i], mn1[i]);
      mx[i] = std::max(mx[i], mx1[i]);
    }
    return;
  }
  switch (type)
  {
    case T4C_SPHERE:
      mn = t.p - V3(r, r, r);
      mx = t.p + V3(r, r, r);
      return;
    case T4C_BOX:
      transformAabb(he, V3(), t, mn, mx);
      return;
    case T4C_CYLINDER:
      transformAabb(V3(r, r, h), V3(), t, mn, mx);
      return;
    case T4C_CAPSULE:
      transformAabb(V3(r, r, r + h), V3(), t, mn, mx);
      return;
    case T4C_HULL:
    {
      // btPolyhedralConvexAabbCachingShape: OBB of the cached local AABB (Bullet pads the cached box by the
      // construction-time margin 0.04; that only changes candidate counts, never results -- not reproduced)
      V3 lmn(LARGE, LARGE, LARGE), lmx(-LARGE, -LARGE, -LARGE);
      for (int i = 0; i < nverts; ++i)
        for (int k = 0; k < 3; ++k)
        {
          lmn[k] = std::min(lmn[k], verts[i][k]);
          lmx[k] = std::max(lmx[k], verts[i][k]);
        }
      transformAabb((lmx - lmn) * 0.5, (lmx + lmn) * 0.5, t, mn, mx);
      return;
    }
    default:
    {
      // getAabbSlow: supports along the six world axes (cone, triangle)
      for (int i = 0; i < 3; ++i)
      {
        V3 vec(0, 0, 0);
        vec[i] = 1.0;
        V3 sv = virtualSupportWithMargin(*this, t.R.tmul(vec));
        mx[i] = t(sv)[i];
        vec[i] = -1.0;
        sv = virtualSupportWithMargin(*this, t.R.tmul(vec));
        mn[i] = t(sv)[i];
      }
      return;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// btVoronoiSimplexSolver (double precision thresholds)
// ------------------------------------------------------------------------------------------------
namespace
{
constexpr double VORONOI_EQUAL_VERTEX_THRESHOLD = 1e-12;

struct SubSimplexResult
{
  V3 closest;
  bool used[4] = { false, false, false, false };
  double bary[4] = { 0, 0, 0, 0 };
  bool degenerate = false;
  void reset()
  {
    degenerate = false;
    setBary(0, 0, 0, 0);
    used[0] = used[1] = used[2] = used[3] = false;
  }
  bool isValid() const { return bary[0] >= 0 && bary[1] >= 0 && bary[2] >= 0 && bary[3] >= 0; }
  void setBary(double a, double b, double c, double d)
  {
    bary[0] = a;
    bary[1] = b;
    bary[2] = c;
    bary[3] = d;
  }
};

struct VoronoiSimplexSolver
{
  int n = 0;
  V3 W[5], P[5], Q[5];
  V3 cachedP1, cachedP2, cachedV, lastW;
  bool cachedValid = false;
  SubSimplexResult bc;
  bool needsUpdate = true;

  void reset()
  {
    cachedValid = false;
    n = 0;
    needsUpdate = true;
    lastW = V3(LARGE, LARGE, LARGE);
    bc.reset();
  }
  void addVertex(const V3& w, const V3& p, const V3& q)
  {
    lastW = w;
    needsUpdate = true;
    W[n] = w;
    P[n] = p;
    Q[n] = q;
    ++n;
  }
  void removeVertex(int i)
  {
    --n;
    W[i] = W[n];
    P[i] = P[n];
    Q[i] = Q[n];
  }
  void reduceVertices(const bool* used)
  {
    if (n >= 4 && !used[3])
      removeVertex(3);
    if (n >= 3 && !used[2])
      removeVertex(2);
    if (n >= 2 && !used[1])
      removeVertex(1);
    if (n >= 1 && !used[0])
      removeVertex(0);
  }
  bool fullSimplex() const { return n == 4; }
  bool inSimplex(const V3& w) const
  {
    bool found = false;
    for (int i = 0; i < n; ++i)
      if ((W[i] - w).length2() <= VORONOI_EQUAL_VERTEX_THRESHOLD)
      {
        found = true;
        break;
      }
    if (w == lastW)
      return true;
    return found;
  }

  static bool closestPtPointTriangle(const V3& p, const V3& a, const V3& b, const V3& c, SubSimplexResult& r)
  {
    r.used[0] = r.used[1] = r.used[2] = r.used[3] = false;
    V3 ab = b - a, ac = c - a, ap = p - a;
    double d1 = ab.dot(ap), d2 = ac.dot(ap);
    if (d1 <= 0.0 && d2 <= 0.0)
    {
      r.closest = a;
      r.used[0] = true;
      r.setBary(1, 0, 0, 0);
      return true;
    }
    V3 bp = p - b;
    double d3 = ab.dot(bp), d4 = ac.dot(bp);
    if (d3 >= 0.0 && d4 <= d3)
    {
      r.closest = b;
      r.used[1] = true;
      r.setBary(0, 1, 0, 0);
      return true;
    }
    double vc = d1 * d4 - d3 * d2;
    if (vc <= 0.0 && d1 >= 0.0 && d3 <= 0.0)
    {
      double v = d1 / (d1 - d3);
      r.closest = a + ab * v;
      r.used[0] = r.used[1] = true;
      r.setBary(1 - v, v, 0, 0);
      return true;
    }
    V3 cp = p - c;
    double d5 = ab.dot(cp), d6 = ac.dot(cp);
    if (d6 >= 0.0 && d5 <= d6)
    {
      r.closest = c;
      r.used[2] = true;
      r.setBary(0, 0, 1, 0);
      return true;
    }
    double vb = d5 * d2 - d1 * d6;
    if (vb <= 0.0 && d2 >= 0.0 && d6 <= 0.0)
    {
      double w = d2 / (d2 - d6);
      r.closest = a + ac * w;
      r.used[0] = r.used[2] = true;
      r.setBary(1 - w, 0, w, 0);
      return true;
    }
    double va = d3 * d6 - d5 * d4;
    if (va <= 0.0 && (d4 - d3) >= 0.0 && (d5 - d6) >= 0.0)
    {
      double w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
      r.closest = b + (c - b) * w;
      r.used[1] = r.used[2] = true;
      r.setBary(0, 1 - w, w, 0);
      return true;
    }
    double denom = 1.0 / (va + vb + vc);
    double v = vb * denom, w = vc * denom;
    r.closest = a + ab * v + ac * w;
    r.used[0] = r.used[1] = r.used[2] = true;
    r.setBary(1 - v - w, v, w, 0);
    return true;
  }

  static int pointOutsideOfPlane(const V3& p, const V3& a, const V3& b, const V3& c, const V3& d)
  {
    V3 normal = (b - a).cross(c - a);
    double signp = (p - a).dot(normal);
    double signd = (d - a).dot(normal);
    if (signd * signd < (1e-8 * 1e-8))
      return -1;
    return signp * signd < 0.0 ? 1 : 0;
  }

  static bool closestPtPointTetrahedron(const V3& p, const V3& a, const V3& b, const V3& c, const V3& d,
                                        SubSimplexResult& fr)
  {
    SubSimplexResult tmp;
    fr.closest = p;
    fr.used[0] = fr.used[1] = fr.used[2] = fr.used[3] = true;
    int oABC = pointOutsideOfPlane(p, a, b, c, d);
    int oACD = pointOutsideOfPlane(p, a, c, d, b);
    int oADB = pointOutsideOfPlane(p, a, d, b, c);
    int oBDC = pointOutsideOfPlane(p, b, d, c, a);
    if (oABC < 0 || oACD < 0 || oADB < 0 || oBDC < 0)
    {
      fr.degenerate = true;
      return false;
    }
    if (!oABC && !oACD && !oADB && !oBDC)
      return false;
    double best = std::numeric_limits<double>::max();
    if (oABC)
    {
      closestPtPointTriangle(p, a, b, c, tmp);
      V3 q = tmp.closest;
      double sq = (q - p).dot(q - p);
      if (sq < best)
      {
        best = sq;
        fr.closest = q;
        fr.used[0] = tmp.used[0];
        fr.used[1] = tmp.used[1];
        fr.used[2] = tmp.used[2];
        fr.used[3] = false;
        fr.setBary(tmp.bary[0], tmp.bary[1], tmp.bary[2], 0);
      }
    }
    if (oACD)
    {
      closestPtPointTriangle(p, a, c, d, tmp);
      V3 q = tmp.closest;
      double sq = (q - p).dot(q - p);
      if (sq < best)
      {
        best = sq;
        fr.closest = q;
        fr.used[0] = tmp.used[0];
        fr.used[1] = false;
        fr.used[2] = tmp.used[1];
        fr.used[3] = tmp.used[2];
        fr.setBary(tmp.bary[0], 0, tmp.bary[1], tmp.bary[2]);
      }
    }
    if (oADB)
    {
      closestPtPointTriangle(p, a, d, b, tmp);
      V3 q = tmp.closest;
      double sq = (q - p).dot(q - p);
      if (sq < best)
      {
        best = sq;
        fr.closest = q;
        fr.used[0] = tmp.used[0];
        fr.used[1] = tmp.used[2];
        fr.used[2] = false;
        fr.used[3] = tmp.used[1];
        fr.setBary(tmp.bary[0], tmp.bary[2], 0, tmp.bary[1]);
      }
    }
    if (oBDC)
    {
      closestPtPointTriangle(p, b, d, c, tmp);
      V3 q = tmp.closest;
      double sq = (q - p).dot(q - p);
      if (sq < best)
      {
        best = sq;
        fr.closest = q;
        fr.used[0] = false;
        fr.used[1] = tmp.used[0];
        fr.used[2] = tmp.used[2];
        fr.used[3] = tmp.used[1];
        fr.setBary(0, tmp.bary[0], tmp.bary[2], tmp.bary[1]);
      }
    }
    return true;
  }

  bool updateClosestVectorAndPoints()
  {
    if (needsUpdate)
    {
      bc.reset();
      needsUpdate = false;
      switch (n)
      {
        case 0:
          cachedValid = false;
          break;
        case 1:
          cachedP1 = P[0];
          cachedP2 = Q[0];
          cachedV = cachedP1 - cachedP2;
          bc.reset();
          bc.setBary(1, 0, 0, 0);
          cachedValid = bc.isValid();
          break;
        case 2:
        {
          const V3& from = W[0];
          const V3& to = W[1];
          V3 diff = V3(0, 0, 0) - from;
          V3 v = to - from;
          double t = v.dot(diff);
          if (t > 0)
          {
            double dotVV = v.dot(v);
            if (t < dotVV)
            {
              t /= dotVV;
              diff -= v * t;
              bc.used[0] = true;
              bc.used[1] = true;
            }
            else
            {
              t = 1;
              diff -= v;
              bc.used[1] = true;
            }
          }
          else
          {
            t = 0;
            bc.used[0] = true;
          }
          bc.setBary(1 - t, t, 0, 0);
          cachedP1 = P[0] + (P[1] - P[0]) * t;
          cachedP2 = Q[0] + (Q[1] - Q[0]) * t;
          cachedV = cachedP1 - cachedP2;
          reduceVertices(bc.used);
          cachedValid = bc.isValid();
          break;
        }
        case 3:
        {
          V3 p(0, 0, 0);
          closestPtPointTriangle(p, W[0], W[1], W[2], bc);
          cachedP1 = P[0] * bc.bary[0] + P[1] * bc.bary[1] + P[2] * bc.bary[2];
          cachedP2 = Q[0] * bc.bary[0] + Q[1] * bc.bary[1] + Q[2] * bc.bary[2];
          cachedV = cachedP1 - cachedP2;
          reduceVertices(bc.used);
          cachedValid = bc.isValid();
          break;
        }
        case 4:
        {
          V3 p(0, 0, 0);
          bool has_sep = closestPtPointTetrahedron(p, W[0], W[1], W[2], W[3], bc);
          if (has_sep)
          {
            cachedP1 = P[0] * bc.bary[0] + P[1] * bc.bary[1] + P[2] * bc.bary[2] + P[3] * bc.bary[3];
            cachedP2 = Q[0] * bc.bary[0] + Q[1] * bc.bary[1] + Q[2] * bc.bary[2] + Q[3] * bc.bary[3];
            cachedV = cachedP1 - cachedP2;
            reduceVertices(bc.used);
          }
          else
          {
            if (bc.degenerate)
              cachedValid = false;
            else
            {
              cachedValid = true;
              cachedV = V3(0, 0, 0);
            }
            break;
          }
          cachedValid = bc.isValid();
          break;
        }
        default:
          cachedValid = false;
      }
    }
    return cachedValid;
  }
  bool closest(V3& v)
  {
    bool ok = updateClosestVectorAndPoints();
    v = cachedV;
    return ok;
  }
  void compute_points(V3& p1, V3& p2)
  {
    updateClosestVectorAndPoints();
    p1 = cachedP1;
    p2 = cachedP2;
  }
};

// ------------------------------------------------------------------------------------------------
// btGjkEpa2 (double precision configuration)
// ------------------------------------------------------------------------------------------------
constexpr int GJK_MAX_ITERATIONS = 128;
constexpr double GJK_ACCURACY = 1e-12;
constexpr double GJK_MIN_DISTANCE = 1e-12;
constexpr double GJK_DUPLICATED_EPS = 1e-12;
constexpr double GJK_SIMPLEX2_EPS = 0.0;
constexpr double GJK_SIMPLEX3_EPS = 0.0;
constexpr double GJK_SIMPLEX4_EPS = 0.0;
constexpr int EPA_MAX_VERTICES = 128;
constexpr int EPA_MAX_ITERATIONS = 255;
constexpr double EPA_ACCURACY = 1e-12;
constexpr double EPA_PLANE_EPS = 1e-14;
constexpr int EPA_MAX_FACES = EPA_MAX_VERTICES * 2;

struct MinkowskiDiff
{
  const Convex* shapes[2];
  M3 toshape1;
  Xf toshape0;
  bool margin = false;
  V3 ls(const Convex* s, const V3& d) const { return margin ? s->supportWithMargin(d) : s->supportNoMargin(d); }
  V3 Support0(const V3& d) const { return ls(shapes[0], d); }
  V3 Support1(const V3& d) const { return toshape0(ls(shapes[1], toshape1 * d)); }
  V3 Support(const V3& d) const { return Support0(d) - Support1(-d); }
  V3 Support(const V3& d, unsigned index) const { return index ? Support1(d) : Support0(d); }
};

struct GJK
{
  struct sSV
  {
    V3 d, w;
  };
  struct sSimplex
  {
    sSV* c[4];
    double p[4];
    unsigned rank;
  };
  enum eStatus
  {
    Valid,
    Inside,
    Failed
  };
  MinkowskiDiff m_shape;
  V3 m_ray;
  double m_distance = 0;
  sSimplex m_simplices[2];
  sSV m_store[4];
  sSV* m_free[4];
  unsigned m_nfree = 0;
  unsigned m_current = 0;
  sSimplex* m_simplex = nullptr;
  eStatus m_status = Failed;

  eStatus Evaluate(const MinkowskiDiff& shapearg, const V3& guess)
  {
    unsigned iterations = 0;
    double sqdist = 0;
    double alpha = 0;
    V3 lastw[4];
    unsigned clastw = 0;
    m_free[0] = &m_store[0];
    m_free[1] = &m_store[1];
    m_free[2] = &m_store[2];
    m_free[3] = &m_store[3];
    m_nfree = 4;
    m_current = 0;
    m_status = Valid;
    m_shape = shapearg;
    m_distance = 0;
    m_simplices[0].rank = 0;
    m_ray = guess;
    const double sqrl = m_ray.length2();
    appendvertice(m_simplices[0], sqrl > 0 ? -m_ray : V3(1, 0, 0));
    m_simplices[0].p[0] = 1;
    m_ray = m_simplices[0].c[0]->w;
    sqdist = sqrl;
    lastw[0] = lastw[1] = lastw[2] = lastw[3] = m_ray;
    do
    {
      const unsigned next = 1 - m_current;
      sSimplex& cs = m_simplices[m_current];
      sSimplex& ns = m_simplices[next];
      const double rl = m_ray.length();
      if (rl < GJK_MIN_DISTANCE)
      {
        m_status = Inside;
        break;
      }
      appendvertice(cs, -m_ray);
      const V3& w = cs.c[cs.rank - 1]->w;
      bool found = false;
      for (unsigned i = 0; i < 4; ++i)
        if ((w - lastw[i]).length2() < GJK_DUPLICATED_EPS)
        {
          found = true;
          break;
        }
      if (found)
      {
        removevertice(m_simplices[m_current]);
        break;
      }
      lastw[clastw = (clastw + 1) & 3] = w;
      const double omega = m_ray.dot(w) / rl;
      alpha = std::max(omega, alpha);
      if (((rl - alpha) - (GJK_ACCURACY * rl)) <= 0)
      {
        removevertice(m_simplices[m_current]);
        break;
      }
      double weights[4];
      unsigned mask = 0;
      switch (cs.rank)
      {
        case 2:
          sqdist = projectorigin(cs.c[0]->w, cs.c[1]->w, weights, mask);
          break;
        case 3:
          sqdist = projectorigin(cs.c[0]->w, cs.c[1]->w, cs.c[2]->w, weights, mask);
          break;
        case 4:
          sqdist = projectorigin(cs.c[0]->w, cs.c[1]->w, cs.c[2]->w, cs.c[3]->w, weights, mask);
          break;
      }
      if (sqdist >= 0)
      {
        ns.rank = 0;
        m_ray = V3(0, 0, 0);
        m_current = next;
        for (unsigned i = 0, ni = cs.rank; i < ni; ++i)
        {
          if (mask & (1u << i))
          {
            ns.c[ns.rank] = cs.c[i];
            ns.p[ns.rank++] = weights[i];
            m_ray += cs.c[i]->w * weights[i];
          }
          else
            m_free[m_nfree++] = cs.c[i];
        }
        if (mask == 15)
          m_status = Inside;
      }
      else
      {
        removevertice(m_simplices[m_current]);
        break;
      }
      m_status = ((++iterations) < (unsigned)GJK_MAX_ITERATIONS) ? m_status : Failed;
    } while (m_status == Valid);
    m_simplex = &m_simplices[m_current];
    switch (m_status)
    {
      case Valid:
        m_distance = m_ray.length();
        break;
      case Inside:
        m_distance = 0;
        break;
      default:
        break;
    }
    return m_status;
  }

  bool EncloseOrigin()
  {
    switch (m_simplex->rank)
    {
      case 1:
        for (unsigned i = 0; i < 3; ++i)
        {
          V3 axis(0, 0, 0);
          axis[i] = 1;
          appendvertice(*m_simplex, axis);
          if (EncloseOrigin())
            return true;
          removevertice(*m_simplex);
          appendvertice(*m_simplex, -axis);
          if (EncloseOrigin())
            return true;
          removevertice(*m_simplex);
        }
        break;
      case 2:
      {
        const V3 d = m_simplex->c[1]->w - m_simplex->c[0]->w;
        for (unsigned i = 0; i < 3; ++i)
        {
          V3 axis(0, 0, 0);
          axis[i] = 1;
          const V3 p = d.cross(axis);
          if (p.length2() > 0)
          {
            appendvertice(*m_simplex, p);
            if (EncloseOrigin())
              return true;
            removevertice(*m_simplex);
            appendvertice(*m_simplex, -p);
            if (EncloseOrigin())
              return true;
            removevertice(*m_simplex);
          }
        }
        break;
      }
      case 3:
      {
        const V3 nrm = (m_simplex->c[1]->w - m_simplex->c[0]->w).cross(m_simplex->c[2]->w - m_simplex->c[0]->w);
        if (nrm.length2() > 0)
        {
          appendvertice(*m_simplex, nrm);
          if (EncloseOrigin())
            return true;
          removevertice(*m_simplex);
          appendvertice(*m_simplex, -nrm);
          if (EncloseOrigin())
            return true;
          removevertice(*m_simplex);
        }
        break;
      }
      case 4:
        if (std::fabs(det(m_simplex->c[0]->w - m_simplex->c[3]->w, m_simplex->c[1]->w - m_simplex->c[3]->w,
                          m_simplex->c[2]->w - m_simplex->c[3]->w)) > 0)
          return true;
        break;
    }
    return false;
  }

  void getsupport(const V3& d, sSV& sv) const
  {
    sv.d = d / d.length();
    sv.w = m_shape.Support(sv.d);
  }
  void removevertice(sSimplex& s) { m_free[m_nfree++] = s.c[--s.rank]; }
  void appendvertice(sSimplex& s, const V3& v)
  {
    s.p[s.rank] = 0;
    s.c[s.rank] = m_free[--m_nfree];
    getsupport(v, *s.c[s.rank++]);
  }
  static double det(const V3& a, const V3& b, const V3& c)
  {
    return a.y * b.z * c.x + a.z * b.x * c.y - a.x * b.z * c.y - a.y * b.x * c.z + a.x * b.y * c.z -
           a.z * b.y * c.x;
  }
  static double projectorigin(const V3& a, const V3& b, double* w, unsigned& m)
  {
    const V3 d = b - a;
    const double l = d.length2();
    if (l > GJK_SIMPLEX2_EPS)
    {
      const double t = l > 0 ? -a.dot(d) / l : 0;
      if (t >= 1)
      {
        w[0] = 0;
        w[1] = 1;
        m = 2;
        return b.length2();
      }
      else if (t <= 0)
      {
        w[0] = 1;
        w[1] = 0;
        m = 1;
        return a.length2();
      }
      else
      {
        w[0] = 1 - (w[1] = t);
        m = 3;
        return (a + d * t).length2();
      }
    }
    return -1;
  }
  static double projectorigin(const V3& a, const V3& b, const V3& c, double* w, unsigned& m)
  {
    static const unsigned imd3[] = { 1, 2, 0 };
    const V3* vt[] = { &a, &b, &c };
    const V3 dl[] = { a - b, b - c, c - a };
    const V3 n = dl[0].cross(dl[1]);
    const double l = n.length2();
    if (l > GJK_SIMPLEX3_EPS)
    {
      double mindist = -1;
      double subw[2] = { 0, 0 };
      unsigned subm = 0;
      for (unsigned i = 0; i < 3; ++i)
      {
        if (vt[i]->dot(dl[i].cross(n)) > 0)
        {
          const unsigned j = imd3[i];
          const double subd = projectorigin(*vt[i], *vt[j], subw, subm);
          if ((mindist < 0) || (subd < mindist))
          {
            mindist = subd;
            m = ((subm & 1) ? 1u << i : 0) + ((subm & 2) ? 1u << j : 0);
            w[i] = subw[0];
            w[j] = subw[1];
            w[imd3[j]] = 0;
          }
        }
      }
      if (mindist < 0)
      {
        const double d = a.dot(n);
        const double s = std::sqrt(l);
        const V3 p = n * (d / l);
        mindist = p.length2();
        m = 7;
        w[0] = (dl[1].cross(b - p)).length() / s;
        w[1] = (dl[2].cross(c - p)).length() / s;
        w[2] = 1 - (w[0] + w[1]);
      }
      return mindist;
    }
    return -1;
  }
  static double projectorigin(const V3& a, const V3& b, const V3& c, const V3& d, double* w, unsigned& m)
  {
    static const unsigned imd3[] = { 1, 2, 0 };
    const V3* vt[] = { &a, &b, &c, &d };
    const V3 dl[] = { a - d, b - d, c - d };
    const double vl = det(dl[0], dl[1], dl[2]);
    const bool ng = (vl * a.dot((b - c).cross(a - b))) <= 0;
    if (ng && (std::fabs(vl) > GJK_SIMPLEX4_EPS))
    {
      double mindist = -1;
      double subw[3] = { 0, 0, 0 };
      unsigned subm = 0;
      for (unsigned i = 0; i < 3; ++i)
      {
        const unsigned j = imd3[i];
        const double s = vl * d.dot(dl[i].cross(dl[j]));
        if (s > 0)
        {
          const double subd = projectorigin(*vt[i], *vt[j], d, subw, subm);
          if ((mindist < 0) || (subd < mindist))
          {
            mindist = subd;
            m = ((subm & 1) ? 1u << i : 0) + ((subm & 2) ? 1u << j : 0) + ((subm & 4) ? 8 : 0);
            w[i] = subw[0];
            w[j] = subw[1];
            w[imd3[j]] = 0;
            w[3] = subw[2];
          }
        }
      }
      if (mindist < 0)
      {
        mindist = 0;
        m = 15;
        w[0] = det(c, b, d) / vl;
        w[1] = det(a, c, d) / vl;
        w[2] = det(b, a, d) / vl;
        w[3] = 1 - (w[0] + w[1] + w[2]);
      }
      return mindist;
    }
    return -1;
  }
};

struct EPA
{
  typedef GJK::sSV sSV;
  struct sFace
  {
    V3 n;
    double d;
    sSV* c[3];
    sFace* f[3];
    sFace* l[2];
    unsigned char e[3];
    unsigned char pass;
  };
  struct sList
  {
    sFace* root = nullptr;
    unsigned count = 0;
  };
  struct sHorizon
  {
    sFace* cf = nullptr;
    sFace* ff = nullptr;
    unsigned nf = 0;
  };
  enum eStatus
  {
    Valid,
    Touching,
    Degenerated,
    NonConvex,
    InvalidHull,
    OutOfFaces,
    OutOfVertices,
    AccuraryReached,
    FallBack,
    Failed
  };
  eStatus m_status = Failed;
  GJK::sSimplex m_result;
  V3 m_normal;
  double m_depth = 0;
  sSV m_sv_store[EPA_MAX_VERTICES];
  sFace m_fc_store[EPA_MAX_FACES];
  unsigned m_nextsv = 0;
  sList m_hull;
  sList m_stock;
  int iterations_done = 0;

  EPA()
  {
    for (unsigned i = 0; i < (unsigned)EPA_MAX_FACES; ++i)
      append(m_stock, &m_fc_store[EPA_MAX_FACES - i - 1]);
  }
  static void bind(sFace* fa, unsigned ea, sFace* fb, unsigned eb)
  {
    fa->e[ea] = (unsigned char)eb;
    fa->f[ea] = fb;
    fb->e[eb] = (unsigned char)ea;
    fb->f[eb] = fa;
  }
  static void append(sList& list, sFace* face)
  {
    face->l[0] = nullptr;
    face->l[1] = list.root;
    if (list.root)
      list.root->l[0] = face;
    list.root = face;
    ++list.count;
  }
  static void remove(sList& list, sFace* face)
  {
    if (face->l[1])
      face->l[1]->l[0] = face->l[0];
    if (face->l[0])
      face->l[0]->l[1] = face->l[1];
    if (face == list.root)
      list.root = face->l[1];
    --list.count;
  }

  eStatus Evaluate(GJK& gjk, const V3& guess)
  {
    GJK::sSimplex& simplex = *gjk.m_simplex;
    if ((simplex.rank > 1) && gjk.EncloseOrigin())
    {
      while (m_hull.root)
      {
        sFace* f = m_hull.root;
        remove(m_hull, f);
        append(m_stock, f);
      }
      m_status = Valid;
      m_nextsv = 0;
      if (GJK::det(simplex.c[0]->w - simplex.c[3]->w, simplex.c[1]->w - simplex.c[3]->w,
                   simplex.c[2]->w - simplex.c[3]->w) < 0)
      {
        std::swap(simplex.c[0], simplex.c[1]);
        std::swap(simplex.p[0], simplex.p[1]);
      }
      sFace* tetra[] = { newface(simplex.c[0], simplex.c[1], simplex.c[2], true),
                         newface(simplex.c[1], simplex.c[0], simplex.c[3], true),
                         newface(simplex.c[2], simplex.c[1], simplex.c[3], true),
                         newface(simplex.c[0], simplex.c[2], simplex.c[3], true) };
      if (m_hull.count == 4)
      {
        sFace* best = findbest();
        sFace outer = *best;
        unsigned pass = 0;
        unsigned iterations = 0;
        bind(tetra[0], 0, tetra[1], 0);
        bind(tetra[0], 1, tetra[2], 0);
        bind(tetra[0], 2, tetra[3], 0);
        bind(tetra[1], 1, tetra[3], 2);
        bind(tetra[1], 2, tetra[2], 1);
        bind(tetra[2], 2, tetra[3], 1);
        m_status = Valid;
        for (; iterations < (unsigned)EPA_MAX_ITERATIONS; ++iterations)
        {
          if (m_nextsv < (unsigned)EPA_MAX_VERTICES)
          {
            sHorizon horizon;
            sSV* w = &m_sv_store[m_nextsv++];
            bool valid = true;
            best->pass = (unsigned char)(++pass);
            gjk.getsupport(best->n, *w);
            const double wdist = best->n.dot(w->w) - best->d;
            if (wdist > EPA_ACCURACY)
            {
              for (unsigned j = 0; (j < 3) && valid; ++j)
                valid &= expand(pass, w, best->f[j], best->e[j], horizon);
              if (valid && (horizon.nf >= 3))
              {
                bind(horizon.cf, 1, horizon.ff, 2);
                remove(m_hull, best);
                append(m_stock, best);
                best = findbest();
                outer = *best;
              }
              else
              {
                m_status = InvalidHull;
                break;
              }
            }
            else
            {
              m_status = AccuraryReached;
              break;
            }
          }
          else
          {
            m_status = OutOfVertices;
            break;
          }
        }
        iterations_done = (int)iterations;
        const V3 projection = outer.n * outer.d;
        m_normal = outer.n;
        m_depth = outer.d;
        m_result.rank = 3;
        m_result.c[0] = outer.c[0];
        m_result.c[1] = outer.c[1];
        m_result.c[2] = outer.c[2];
        m_result.p[0] = (outer.c[1]->w - projection).cross(outer.c[2]->w - projection).length();
        m_result.p[1] = (outer.c[2]->w - projection).cross(outer.c[0]->w - projection).length();
        m_result.p[2] = (outer.c[0]->w - projection).cross(outer.c[1]->w - projection).length();
        const double sum = m_result.p[0] + m_result.p[1] + m_result.p[2];
        m_result.p[0] /= sum;
        m_result.p[1] /= sum;
        m_result.p[2] /= sum;
        return m_status;
      }
    }
    m_status = FallBack;
    m_normal = -guess;
    const double nl = m_normal.length();
    if (nl > 0)
      m_normal = m_normal / nl;
    else
      m_normal = V3(1, 0, 0);
    m_depth = 0;
    m_result.rank = 1;
    m_result.c[0] = simplex.c[0];
    m_result.p[0] = 1;
    return m_status;
  }

  static bool getedgedist(sFace* face, sSV* a, sSV* b, double& dist)
  {
    const V3 ba = b->w - a->w;
    const V3 n_ab = ba.cross(face->n);
    const double a_dot_nab = a->w.dot(n_ab);
    if (a_dot_nab < 0)
    {
      const double ba_l2 = ba.length2();
      const double a_dot_ba = a->w.dot(ba);
      const double b_dot_ba = b->w.dot(ba);
      if (a_dot_ba > 0)
        dist = a->w.length();
      else if (b_dot_ba < 0)
        dist = b->w.length();
      else
      {
        const double a_dot_b = a->w.dot(b->w);
        dist = std::sqrt(std::max((a->w.length2() * b->w.length2() - a_dot_b * a_dot_b) / ba_l2, 0.0));
      }
      return true;
    }
    return false;
  }

  sFace* newface(sSV* a, sSV* b, sSV* c, bool forced)
  {
    if (m_stock.root)
    {
      sFace* face = m_stock.root;
      remove(m_stock, face);
      append(m_hull, face);
      face->pass = 0;
      face->c[0] = a;
      face->c[1] = b;
      face->c[2] = c;
      face->n = (b->w - a->w).cross(c->w - a->w);
      const double l = face->n.length();
      const bool v = l > EPA_ACCURACY;
      if (v)
      {
        if (!(getedgedist(face, a, b, face->d) || getedgedist(face, b, c, face->d) ||
              getedgedist(face, c, a, face->d)))
          face->d = a->w.dot(face->n) / l;
        face->n /= l;
        if (forced || (face->d >= -EPA_PLANE_EPS))
          return face;
        else
        {
          m_status = NonConvex;
          if (getenv("T4O_DEBUG"))
            fprintf(stderr, "[t4o] newface NonConvex d=%.6g planed=%.6g\n", face->d, a->w.dot(face->n));
        }
      }
      else
      {
        m_status = Degenerated;
        if (getenv("T4O_DEBUG"))
          fprintf(stderr, "[t4o] newface Degenerated l=%.6g\n", l);
      }
      remove(m_hull, face);
      append(m_stock, face);
      return nullptr;
    }
    m_status = m_stock.root ? OutOfVertices : OutOfFaces;
    return nullptr;
  }

  sFace* findbest()
  {
    sFace* minf = m_hull.root;
    double mind = minf->d * minf->d;
    for (sFace* f = minf->l[1]; f; f = f->l[1])
    {
      const double sqd = f->d * f->d;
      if (sqd < mind)
      {
        minf = f;
        mind = sqd;
      }
    }
    return minf;
  }

  bool expand(unsigned pass, sSV* w, sFace* f, unsigned e, sHorizon& horizon)
  {
    static const unsigned i1m3[] = { 1, 2, 0 };
    static const unsigned i2m3[] = { 2, 0, 1 };
    if (f->pass != pass)
    {
      const unsigned e1 = i1m3[e];
      if ((f->n.dot(w->w) - f->d) < -EPA_PLANE_EPS)
      {
        sFace* nf = newface(f->c[e1], f->c[e], w, false);
        if (nf)
        {
          bind(nf, 0, f, e);
          if (horizon.cf)
            bind(horizon.cf, 1, nf, 2);
          else
            horizon.ff = nf;
          horizon.cf = nf;
          ++horizon.nf;
          return true;
        }
      }
      else
      {
        const unsigned e2 = i2m3[e];
        f->pass = (unsigned char)pass;
        if (expand(pass, w, f->f[e1], f->e[e1], horizon) && expand(pass, w, f->f[e2], f->e[e2], horizon))
        {
          remove(m_hull, f);
          append(m_stock, f);
          return true;
        }
      }
    }
    return false;
  }
};


// ------------------------------------------------------------------------------------------------
// Robust EPA used ONLY when the faithful btGjkEpa2 restatement above stops unconverged (statuses
// Degenerated / NonConvex / InvalidHull / OutOfFaces / OutOfVertices).  Those exits are artefacts of
// exactly degenerate input (e.g. the perfectly symmetric sphere_p25m.ply hull pair, where Bullet itself
// only converges because btConvexHullComputer snaps the vertices) or of Bullet's 128-vertex
// cap; the reference would return the current, unconverged polytope face.  The oracle instead returns the
// converged minimum-translation answer -- the value the reference's own golden tests pin.  DEVIATION,
// documented in DESIGN.md.  Algorithm: classic expanding polytope with breadth-first visible-set marking and
// an explicit horizon (no recursive silhouette walk), same support mapping / margins as btGjkEpa2.
// ------------------------------------------------------------------------------------------------
struct RobustEPA
{
  struct Vert
  {
    V3 w, d;
  };
  struct Face
  {
    int v[3];
    int adj[3];
    V3 n;
    double d;
    bool alive;
    int mark;
  };
  std::vector<Vert> verts;
  std::vector<Face> faces;
  int iterations = 0;
  V3 normal;
  double depth = 0;
  int rv[3];
  double rp[3];
  bool ok = false;
  double saved_depth = -1e300;
  Face saved_face;
  Vert saved_verts[3];
  bool have_saved = false, deteriorated = false;

  bool makeFace(int a, int b, int c, Face& f) const
  {
    f.v[0] = a;
    f.v[1] = b;
    f.v[2] = c;
    f.adj[0] = f.adj[1] = f.adj[2] = -1;
    V3 n = (verts[b].w - verts[a].w).cross(verts[c].w - verts[a].w);
    double l = n.length();
    if (!(l > 1e-20))
      return false;
    f.n = n / l;
    f.d = f.n.dot(verts[a].w);
    f.alive = true;
    f.mark = 0;
    return true;
  }

  static int edgeIndex(const Face& g, int from, int to)
  {
    for (int e = 0; e < 3; ++e)
      if (g.v[e] == from && g.v[(e + 1) % 3] == to)
        return e;
    return -1;
  }

  bool run(const MinkowskiDiff& shape, const GJK::sSimplex& simplex, int max_iter = 2000, double tol = 1e-12)
  {
    if (simplex.rank != 4)
      return false;
    verts.clear();
    faces.clear();
    for (int i = 0; i < 4; ++i)
      verts.push_back({ simplex.c[i]->w, simplex.c[i]->d });
    if (GJK::det(verts[0].w - verts[3].w, verts[1].w - verts[3].w, verts[2].w - verts[3].w) < 0)
      std::swap(verts[0], verts[1]);
    const int tet[4][3] = { { 0, 1, 2 }, { 1, 0, 3 }, { 2, 1, 3 }, { 0, 2, 3 } };
    for (int i = 0; i < 4; ++i)
    {
      Face f;
      if (!makeFace(tet[i][0], tet[i][1], tet[i][2], f))
        return false;
      faces.push_back(f);
    }
    for (int i = 0; i < 4; ++i)
      for (int e = 0; e < 3; ++e)
        for (int j = 0; j < 4; ++j)
          if (j != i && edgeIndex(faces[j], faces[i].v[(e + 1) % 3], faces[i].v[e]) >= 0)
            faces[i].adj[e] = j;
    int mark = 0;
    int best = -1;
    for (iterations = 0; iterations < max_iter; ++iterations)
    {
      best = -1;
      for (int i = 0; i < (int)faces.size(); ++i)
        if (faces[i].alive && (best < 0 || faces[i].d < faces[best].d))
          best = i;
      if (best < 0)
        return false;
      // The closest-face distance of an expanding polytope is a LOWER bound of the penetration depth and can only
      // grow.  On curved supports (cylinder rims) sliver faces eventually corrupt the hull and the bound drops:
      // stop there and report the tightest bound seen.
      if (faces[best].d + 1e-12 < saved_depth)
      {
        deteriorated = true;
        break;
      }
      if (faces[best].d >= saved_depth)
      {
        saved_depth = faces[best].d;
        saved_face = faces[best];
        for (int k = 0; k < 3; ++k)
          saved_verts[k] = verts[faces[best].v[k]];
        have_saved = true;
      }
      Vert nv;
      nv.d = faces[best].n;
      nv.w = shape.Support(nv.d);
      double wdist = faces[best].n.dot(nv.w) - faces[best].d;
      if (wdist <= tol)
        break;
      int wi = (int)verts.size();
      verts.push_back(nv);
      // visible set: breadth-first from the closest face
      ++mark;
      std::vector<int> stack{ best };
      std::vector<int> visible;
      faces[best].mark = mark;
      while (!stack.empty())
      {
        int f = stack.back();
        stack.pop_back();
        visible.push_back(f);
        for (int e = 0; e < 3; ++e)
        {
          int g = faces[f].adj[e];
          if (g < 0 || faces[g].mark == mark)
            continue;
          if (faces[g].n.dot(nv.w) - faces[g].d > 1e-14)
          {
            faces[g].mark = mark;
            stack.push_back(g);
          }
        }
      }
      // horizon; absorb outside neighbours that would give a zero-area new face
      struct HEdge
      {
        int a, b, g;
      };
      std::vector<HEdge> horizon;
      for (int guard = 0; guard < 64; ++guard)
      {
        horizon.clear();
        bool grew = false;
        for (int f : visible)
          for (int e = 0; e < 3 && !grew; ++e)
          {
            int g = faces[f].adj[e];
            if (g >= 0 && faces[g].mark == mark)
              continue;
            int a = faces[f].v[e], b = faces[f].v[(e + 1) % 3];
            V3 n = (verts[b].w - verts[a].w).cross(nv.w - verts[a].w);
            if (g >= 0 && n.length() < 1e-14 * (1.0 + nv.w.length2()))
            {
              faces[g].mark = mark;
              visible.push_back(g);
              grew = true;
              break;
            }
            horizon.push_back({ a, b, g });
          }
        if (!grew)
          break;
      }
      if (horizon.size() < 3)
        break;
      for (int f : visible)
        faces[f].alive = false;
      int first_new = (int)faces.size();
      bool bad = false;
      for (const HEdge& h : horizon)
      {
        Face nf;
        if (!makeFace(h.a, h.b, wi, nf))
        {
          bad = true;
          break;
        }
        nf.adj[0] = h.g;
        faces.push_back(nf);
        if (h.g >= 0)
        {
          int ge = edgeIndex(faces[h.g], h.b, h.a);
          if (ge >= 0)
            faces[h.g].adj[ge] = (int)faces.size() - 1;
        }
      }
      if (bad)
        break;
      for (int i = first_new; i < (int)faces.size(); ++i)
        for (int j = first_new; j < (int)faces.size(); ++j)
        {
          if (faces[j].v[0] == faces[i].v[1])
            faces[i].adj[1] = j;  // edge (b, w) twin is (w, b) of the face starting at b
          if (faces[j].v[1] == faces[i].v[0])
            faces[i].adj[2] = j;  // edge (w, a) twin is (a, w) of the face ending at a
        }
    }
    if (best < 0)
      return false;
    if (have_saved && (deteriorated || faces[best].d < saved_depth))
    {
      // report the tightest lower bound (see above); its three vertices were copied when it was recorded
      for (int k = 0; k < 3; ++k)
      {
        verts.push_back(saved_verts[k]);
        saved_face.v[k] = (int)verts.size() - 1;
      }
      faces.push_back(saved_face);
      best = (int)faces.size() - 1;
    }
    const Face& o = faces[best];
    normal = o.n;
    depth = o.d;
    V3 proj = o.n * o.d;
    for (int k = 0; k < 3; ++k)
      rv[k] = o.v[k];
    rp[0] = (verts[o.v[1]].w - proj).cross(verts[o.v[2]].w - proj).length();
    rp[1] = (verts[o.v[2]].w - proj).cross(verts[o.v[0]].w - proj).length();
    rp[2] = (verts[o.v[0]].w - proj).cross(verts[o.v[1]].w - proj).length();
    double sum = rp[0] + rp[1] + rp[2];
    if (!(sum > 0))
      return false;
    for (int k = 0; k < 3; ++k)
      rp[k] /= sum;
    ok = true;
    return true;
  }
};

void initMinkowski(const Convex& a, const Xf& wtrs0, const Convex& b, const Xf& wtrs1, GjkEpaResults& res,
                   MinkowskiDiff& shape, bool withmargins)
{
  res.witnesses[0] = res.witnesses[1] = V3(0, 0, 0);
  res.status = 0;
  shape.shapes[0] = &a;
  shape.shapes[1] = &b;
  shape.toshape1 = wtrs1.R.transpose() * wtrs0.R;  // transposeTimes
  shape.toshape0 = wtrs0.inverseTimes(wtrs1);
  shape.margin = withmargins;
}
}  // namespace

bool gjkEpaPenetration(const Convex& a, const Xf& wtrs0, const Convex& b, const Xf& wtrs1, const V3& guess,
                       GjkEpaResults& results, bool usemargins)
{
  MinkowskiDiff shape;
  initMinkowski(a, wtrs0, b, wtrs1, results, shape, usemargins);
  GJK gjk;
  GJK::eStatus gjk_status = gjk.Evaluate(shape, -guess);
  switch (gjk_status)
  {
    case GJK::Inside:
    {
      EPA epa;
      EPA::eStatus epa_status = epa.Evaluate(gjk, -guess);
      results.epa_iterations = epa.iterations_done;
      if (getenv("T4O_DEBUG"))
        fprintf(stderr, "[t4o] EPA status=%d iters=%d depth=%.9g nextsv=%u hull=%u\n", (int)epa_status,
                epa.iterations_done, epa.m_depth, epa.m_nextsv, epa.m_hull.count);
      const bool unconverged = epa_status == EPA::Degenerated || epa_status == EPA::NonConvex ||
                               epa_status == EPA::InvalidHull || epa_status == EPA::OutOfFaces ||
                               epa_status == EPA::OutOfVertices;
      if (unconverged && gjk.m_simplex->rank == 4)
      {
        RobustEPA repa;
        // both the faithful polytope's closest face and the robust one are lower bounds of the depth: keep the
        // tighter (the faithful result is reported by the generic branch below)
        if (repa.run(shape, *gjk.m_simplex) && (epa_status == EPA::Failed || repa.depth >= epa.m_depth))
        {
          results.epa_iterations += repa.iterations;
          V3 w0(0, 0, 0);
          for (int i = 0; i < 3; ++i)
            w0 += shape.Support(repa.verts[repa.rv[i]].d, 0) * repa.rp[i];
          results.status = 1;
          results.witnesses[0] = wtrs0 * w0;
          results.witnesses[1] = wtrs0 * (w0 - repa.normal * repa.depth);
          results.normal = -repa.normal;
          results.distance = -repa.depth;
          if (getenv("T4O_DEBUG"))
            fprintf(stderr, "[t4o] robust EPA iters=%d depth=%.9g\n", repa.iterations, repa.depth);
          return true;
        }
      }
      if (epa_status != EPA::Failed)
      {
        V3 w0(0, 0, 0);
        for (unsigned i = 0; i < epa.m_result.rank; ++i)
          w0 += shape.Support(epa.m_result.c[i]->d, 0) * epa.m_result.p[i];
        results.status = 1;
        results.witnesses[0] = wtrs0 * w0;
        results.witnesses[1] = wtrs0 * (w0 - epa.m_normal * epa.m_depth);
        results.normal = -epa.m_normal;
        results.distance = -epa.m_depth;
        return true;
      }
      results.status = 3;
      break;
    }
    case GJK::Failed:
      results.status = 2;
      break;
    default:
      break;
  }
  return false;
}

bool gjkEpaDistance(const Convex& a, const Xf& wtrs0, const Convex& b, const Xf& wtrs1, const V3& guess,
                    GjkEpaResults& results)
{
  MinkowskiDiff shape;
  initMinkowski(a, wtrs0, b, wtrs1, results, shape, false);
  GJK gjk;
  GJK::eStatus gjk_status = gjk.Evaluate(shape, guess);
  if (gjk_status == GJK::Valid)
  {
    V3 w0(0, 0, 0), w1(0, 0, 0);
    for (unsigned i = 0; i < gjk.m_simplex->rank; ++i)
    {
      const double p = gjk.m_simplex->p[i];
      w0 += shape.Support(gjk.m_simplex->c[i]->d, 0) * p;
      w1 += shape.Support(-gjk.m_simplex->c[i]->d, 1) * p;
    }
    results.witnesses[0] = wtrs0 * w0;
    results.witnesses[1] = wtrs0 * w1;
    results.normal = w0 - w1;
    results.distance = results.normal.length();
    results.normal /= results.distance > GJK_MIN_DISTANCE ? results.distance : 1;
    return true;
  }
  results.status = gjk_status == GJK::Inside ? 1 : 2;
  return false;
}

namespace
{
V3 safeNormalize(V3 v)
{
  double l2 = v.length2();
  if (l2 >= SIMD_EPS * SIMD_EPS)
    return v / std::sqrt(l2);
  return V3(1, 0, 0);
}

// btGjkEpaPenetrationDepthSolver::calcPenDepth
bool calcPenDepth(const Convex& a, const Convex& b, const Xf& ta, const Xf& tb, V3& v, V3& wa, V3& wb,
                  Counters& cnt)
{
  V3 guess[] = { safeNormalize(tb.p - ta.p), safeNormalize(ta.p - tb.p), V3(0, 0, 1), V3(0, 1, 0), V3(1, 0, 0),
                 V3(1, 1, 0), V3(1, 1, 1), V3(0, 1, 1), V3(1, 0, 1) };
  for (const V3& g : guess)
  {
    GjkEpaResults results;
    if (gjkEpaPenetration(a, ta, b, tb, g, results))
    {
      cnt.epa_iterations += (uint64_t)results.epa_iterations;
      wa = results.witnesses[0];
      wb = results.witnesses[1];
      v = results.normal;
      return true;
    }
    cnt.epa_iterations += (uint64_t)results.epa_iterations;
    if (gjkEpaDistance(a, ta, b, tb, g, results))
    {
      wa = results.witnesses[0];
      wb = results.witnesses[1];
      v = results.normal;
      return false;
    }
  }
  wa = wb = v = V3(0, 0, 0);
  return false;
}

// ------------------------------------------------------------------------------------------------
// libccd-style simplex helpers of TesseractGjkPairDetector (tesseract_gjk_pair_detector.cpp:136-676)
// ------------------------------------------------------------------------------------------------
struct SupportVec
{
  V3 v, v1, v2;
};
struct CcdSimplex
{
  SupportVec ps[4];
  int last = -1;
  int size() const { return last + 1; }
  void add(const SupportVec& s) { ps[++last] = s; }
  void set(int pos, const SupportVec& s) { ps[pos] = s; }
  void setSize(int n) { last = n - 1; }
};

int ccdEq(double a_, double b_)
{
  double ab = std::fabs(a_ - b_);
  if (std::fabs(ab) < SIMD_EPS)
    return 1;
  double a = std::fabs(a_), b = std::fabs(b_);
  if (b > a)
    return ab < SIMD_EPS * b;
  return ab < SIMD_EPS * a;
}
int vec3Eq(const V3& a, const V3& b) { return ccdEq(a.x, b.x) && ccdEq(a.y, b.y) && ccdEq(a.z, b.z); }
int ccdSign(double val)
{
  if (fuzzyZero(val))
    return 0;
  return val < 0 ? -1 : 1;
}
V3 tripleCross(const V3& a, const V3& b, const V3& c) { return a.cross(b).cross(c); }

double pointSegmentDist2(const V3& P, const V3& x0, const V3& b, V3* witness)
{
  double dist, t;
  V3 d = b - x0;
  V3 a = x0 - P;
  t = -1.0 * a.dot(d);
  t /= d.dot(d);
  if (t < 0 || fuzzyZero(t))
  {
    dist = (x0 - P).length2();
    if (witness)
      *witness = x0;
  }
  else if (t > 1 || ccdEq(t, 1))
  {
    dist = (b - P).length2();
    if (witness)
      *witness = b;
  }
  else
  {
    if (witness)
    {
      *witness = d * t + x0;
      dist = (*witness - P).length2();
    }
    else
    {
      d = d * t + a;
      dist = d.dot(d);
    }
  }
  return dist;
}

double pointTriDist2(const V3& P, const V3& x0, const V3& B, const V3& C, V3* witness)
{
  V3 d1 = B - x0, d2 = C - x0, a = x0 - P;
  double u = a.dot(a), v = d1.dot(d1), w = d2.dot(d2), p = a.dot(d1), q = a.dot(d2), r = d1.dot(d2);
  double s = (q * r - w * p) / (w * v - r * r);
  double t = (-s * r - q) / w;
  double dist, dist2;
  V3 witness2;
  if ((fuzzyZero(s) || s > 0) && (ccdEq(s, 1) || s < 1) && (fuzzyZero(t) || t > 0) && (ccdEq(t, 1) || t < 1) &&
      (ccdEq(t + s, 1) || (t + s) < 1))
  {
    if (witness)
    {
      *witness = x0 + d1 * s + d2 * t;
      dist = (*witness - P).length2();
    }
    else
    {
      dist = s * s * v;
      dist += t * t * w;
      dist += 2.0 * s * t * r;
      dist += 2.0 * s * p;
      dist += 2.0 * t * q;
      dist += u;
    }
  }
  else
  {
    dist = pointSegmentDist2(P, x0, B, witness);
    dist2 = pointSegmentDist2(P, x0, C, &witness2);
    if (dist2 < dist)
    {
      dist = dist2;
      if (witness)
        *witness = witness2;
    }
    dist2 = pointSegmentDist2(P, B, C, &witness2);
    if (dist2 < dist)
    {
      dist = dist2;
      if (witness)
        *witness = witness2;
    }
  }
  return dist;
}

int doSimplex2(CcdSimplex& s, V3& dir)
{
  const SupportVec A = s.ps[s.last];
  const SupportVec B = s.ps[0];
  V3 AB = B.v - A.v;
  V3 AO = -A.v;
  double dot = AB.dot(AO);
  V3 tmp = AB.cross(AO);
  if (fuzzyZero(tmp.dot(tmp)) && dot > 0)
    return 1;
  if (fuzzyZero(dot) || dot < 0)
  {
    s.set(0, A);
    s.setSize(1);
    dir = AO;
  }
  else
    dir = tripleCross(AB, AO, AB);
  return 0;
}

int doSimplex3(CcdSimplex& s, V3& dir)
{
  const SupportVec A = s.ps[s.last];
  const SupportVec B = s.ps[1];
  const SupportVec C = s.ps[0];
  double dist = pointTriDist2(V3(0, 0, 0), A.v, B.v, C.v, nullptr);
  if (fuzzyZero(dist))
    return 1;
  if (vec3Eq(A.v, B.v) || vec3Eq(A.v, C.v))
    return -1;
  V3 AO = -A.v;
  V3 AB = B.v - A.v, AC = C.v - A.v;
  V3 ABC = AB.cross(AC);
  V3 tmp = ABC.cross(AC);
  double dot = tmp.dot(AO);
  if (fuzzyZero(dot) || dot > 0)
  {
    dot = AC.dot(AO);
    if (fuzzyZero(dot) || dot > 0)
    {
      s.set(1, A);
      s.setSize(2);
      dir = tripleCross(AC, AO, AC);
    }
    else
    {
      dot = AB.dot(AO);
      if (fuzzyZero(dot) || dot > 0)
      {
        s.set(0, B);
        s.set(1, A);
        s.setSize(2);
        dir = tripleCross(AB, AO, AB);
      }
      else
      {
        s.set(0, A);
        s.setSize(1);
        dir = AO;
      }
    }
  }
  else
  {
    tmp = AB.cross(ABC);
    dot = tmp.dot(AO);
    if (fuzzyZero(dot) || dot > 0)
    {
      dot = AB.dot(AO);
      if (fuzzyZero(dot) || dot > 0)
      {
        s.set(0, B);
        s.set(1, A);
        s.setSize(2);
        dir = tripleCross(AB, AO, AB);
      }
      else
      {
        s.set(0, A);
        s.setSize(1);
        dir = AO;
      }
    }
    else
    {
      dot = ABC.dot(AO);
      if (fuzzyZero(dot) || dot > 0)
        dir = ABC;
      else
      {
        SupportVec t = C;
        s.set(0, B);
        s.set(1, t);
        dir = -ABC;
      }
    }
  }
  return 0;
}

int doSimplex4(CcdSimplex& s, V3& dir)
{
  const SupportVec A = s.ps[s.last];
  const SupportVec B = s.ps[2];
  const SupportVec C = s.ps[1];
  const SupportVec D = s.ps[0];
  double dist = pointTriDist2(A.v, B.v, C.v, D.v, nullptr);
  if (fuzzyZero(dist))
    return -1;
  const V3 O(0, 0, 0);
  if (fuzzyZero(pointTriDist2(O, A.v, B.v, C.v, nullptr)))
    return 1;
  if (fuzzyZero(pointTriDist2(O, A.v, C.v, D.v, nullptr)))
    return 1;
  if (fuzzyZero(pointTriDist2(O, A.v, B.v, D.v, nullptr)))
    return 1;
  if (fuzzyZero(pointTriDist2(O, B.v, C.v, D.v, nullptr)))
    return 1;
  V3 AO = -A.v;
  V3 AB = B.v - A.v, AC = C.v - A.v, AD = D.v - A.v;
  V3 ABC = AB.cross(AC), ACD = AC.cross(AD), ADB = AD.cross(AB);
  int B_on_ACD = ccdSign(ACD.dot(AB));
  int C_on_ADB = ccdSign(ADB.dot(AC));
  int D_on_ABC = ccdSign(ABC.dot(AD));
  int AB_O = ccdSign(ACD.dot(AO)) == B_on_ACD;
  int AC_O = ccdSign(ADB.dot(AO)) == C_on_ADB;
  int AD_O = ccdSign(ABC.dot(AO)) == D_on_ABC;
  if (AB_O && AC_O && AD_O)
    return 1;
  else if (!AB_O)
  {
    s.set(2, A);
    s.setSize(3);
  }
  else if (!AC_O)
  {
    s.set(1, D);
    s.set(0, B);
    s.set(2, A);
    s.setSize(3);
  }
  else
  {
    s.set(0, C);
    s.set(1, B);
    s.set(2, A);
    s.setSize(3);
  }
  return doSimplex3(s, dir);
}

int doSimplex(CcdSimplex& s, V3& dir)
{
  if (s.size() == 2)
    return doSimplex2(s, dir);
  if (s.size() == 3)
    return doSimplex3(s, dir);
  return doSimplex4(s, dir);
}

void computeSupport(const Convex& a, const Xf& ta, const Convex& b, const Xf& tb, const V3& dir, V3& supA, V3& supB,
                    V3& aMinb)
{
  V3 dirA = ta.R.tmul(dir);
  V3 dirB = tb.R.tmul(-dir);
  supA = ta(a.supportNoMargin(dirA));
  supB = tb(b.supportNoMargin(dirB));
  aMinb = supA - supB;
}

constexpr double REL_ERROR2 = 1.0e-12;
constexpr double GJK_EPA_PENETRATION_TOLERANCE = 1.0e-12;

// TesseractGjkPairDetector::getClosestPointsNonVirtual (tesseract_gjk_pair_detector.cpp:683-1185)
void gjkPairDetector(const Convex& A, const Xf& transA, const Convex& B, const Xf& transB, double maxDist2,
                     const PairRequest& req, PairResult& out, Counters& cnt)
{
  double distance = 0.0;
  V3 normalInB(0, 0, 0);
  V3 pointOnA, pointOnB;
  Xf localTransA = transA, localTransB = transB;
  V3 positionOffset = (localTransA.p + localTransB.p) * 0.5;
  localTransA.p -= positionOffset;
  localTransB.p -= positionOffset;

  double marginA = A.margin(), marginB = B.margin();
  int curIter = 0;
  const int gGjkMaxIter = 1000;
  V3 cachedSeparatingAxis(0, 1, 0);
  bool isValid = false, checkSimplex = false;
  const bool checkPenetration = true;
  int degenerateSimplex = 0;
  int status = -2;
  V3 orgNormalInB(0, 0, 0);
  double margin = marginA + marginB;
  cnt.gjk_pairs++;

  {
    double squaredDistance = LARGE;
    double delta = 0.0;
    CcdSimplex simplex;
    V3 dir(1, 0, 0);
    {
      V3 lastSupV, supAworld, supBworld;
      computeSupport(A, localTransA, B, localTransB, dir, supAworld, supBworld, lastSupV);
      cnt.gjk_iterations++;
      SupportVec last;
      last.v = lastSupV;
      last.v1 = supAworld;
      last.v2 = supBworld;
      simplex.add(last);
      dir = -lastSupV;
      for (int iterations = 0; iterations < gGjkMaxIter; iterations++)
      {
        computeSupport(A, localTransA, B, localTransB, dir, supAworld, supBworld, lastSupV);
        cnt.gjk_iterations++;
        double dlt = lastSupV.dot(dir);
        if (dlt < 0)
        {
          status = -1;
          break;
        }
        last.v = lastSupV;
        last.v1 = supAworld;
        last.v2 = supBworld;
        simplex.add(last);
        int res = doSimplex(simplex, dir);
        if (res == 1)
        {
          status = 0;
          break;
        }
        else if (res == -1)
        {
          status = -1;
          break;
        }
        if (fuzzyZero(dir.dot(dir)))
          status = -1;
        if (dir.length2() < SIMD_EPS)
        {
          status = -1;
          break;
        }
        if (dir.fuzzyZero())
        {
          status = -1;
          break;
        }
      }
    }

    if (status == 0 && !req.calculate_penetration)
    {
      out.has = true;
      out.normalOnB = normalInB;
      out.pointOnB = pointOnB + positionOffset;  // pointOnB is uninitialised in the reference; zero here
      out.distance = -std::numeric_limits<double>::min();
      return;
    }
    if (status == -1 && !req.calculate_distance)
      return;

    VoronoiSimplexSolver solver;
    solver.reset();
    for (;;)
    {
      V3 sepAxisInA = localTransA.R.tmul(-cachedSeparatingAxis);
      V3 sepAxisInB = localTransB.R.tmul(cachedSeparatingAxis);
      V3 pInA = A.supportNoMargin(sepAxisInA);
      V3 qInB = B.supportNoMargin(sepAxisInB);
      V3 pWorld = localTransA(pInA);
      V3 qWorld = localTransB(qInB);
      cnt.gjk_iterations++;
      V3 w = pWorld - qWorld;
      delta = cachedSeparatingAxis.dot(w);
      if ((delta > 0.0) && (delta * delta > squaredDistance * maxDist2))
      {
        degenerateSimplex = 10;
        checkSimplex = true;
        break;
      }
      if (solver.inSimplex(w))
      {
        degenerateSimplex = 1;
        checkSimplex = true;
        break;
      }
      double f0 = squaredDistance - delta;
      double f1 = squaredDistance * REL_ERROR2;
      if (f0 <= f1)
      {
        degenerateSimplex = (f0 <= 0.0) ? 2 : 11;
        checkSimplex = true;
        break;
      }
      solver.addVertex(w, pWorld, qWorld);
      V3 newAxis;
      if (!solver.closest(newAxis))
      {
        degenerateSimplex = 3;
        checkSimplex = true;
        break;
      }
      if (newAxis.length2() < REL_ERROR2)
      {
        cachedSeparatingAxis = newAxis;
        degenerateSimplex = 6;
        checkSimplex = true;
        break;
      }
      double previousSquaredDistance = squaredDistance;
      squaredDistance = newAxis.length2();
      if (previousSquaredDistance - squaredDistance <= SIMD_EPS * previousSquaredDistance)
      {
        checkSimplex = true;
        degenerateSimplex = 12;
        break;
      }
      cachedSeparatingAxis = newAxis;
      if (curIter++ > gGjkMaxIter)
        break;
      if (solver.fullSimplex())
      {
        degenerateSimplex = 13;
        break;
      }
    }

    if (checkSimplex)
    {
      solver.compute_points(pointOnA, pointOnB);
      normalInB = cachedSeparatingAxis;
      double lenSqr = cachedSeparatingAxis.length2();
      if (lenSqr < REL_ERROR2)
        degenerateSimplex = 5;
      if (lenSqr > SIMD_EPS * SIMD_EPS)
      {
        double rlen = 1.0 / std::sqrt(lenSqr);
        normalInB *= rlen;
        double s = std::sqrt(squaredDistance);
        pointOnA -= cachedSeparatingAxis * (marginA / s);
        pointOnB += cachedSeparatingAxis * (marginB / s);
        distance = (1.0 / rlen) - margin;
        isValid = true;
        orgNormalInB = normalInB;
      }
    }

    bool catchDegeneratePenetrationCase =
        (degenerateSimplex != 0) && ((distance + margin) < GJK_EPA_PENETRATION_TOLERANCE);

    if ((checkPenetration && (!isValid || catchDegeneratePenetrationCase)) || (status == 0))
    {
      V3 tmpPointOnA, tmpPointOnB;
      cachedSeparatingAxis = V3(0, 0, 0);
      cnt.epa_pairs++;
      bool isValid2 = calcPenDepth(A, B, localTransA, localTransB, cachedSeparatingAxis, tmpPointOnA, tmpPointOnB, cnt);
      if (cachedSeparatingAxis.length2() > 0)
      {
        if (isValid2)
        {
          V3 tmpNormalInB = tmpPointOnB - tmpPointOnA;
          double lenSqr = tmpNormalInB.length2();
          if (lenSqr <= (SIMD_EPS * SIMD_EPS))
          {
            tmpNormalInB = cachedSeparatingAxis;
            lenSqr = cachedSeparatingAxis.length2();
          }
          if (lenSqr > (SIMD_EPS * SIMD_EPS))
          {
            tmpNormalInB /= std::sqrt(lenSqr);
            double distance2 = -(tmpPointOnA - tmpPointOnB).length();
            if (!isValid || (distance2 < distance))
            {
              distance = distance2;
              pointOnA = tmpPointOnA;
              pointOnB = tmpPointOnB;
              normalInB = tmpNormalInB;
              isValid = true;
            }
          }
        }
        else
        {
          if (cachedSeparatingAxis.length2() > 0.0)
          {
            double distance2 = (tmpPointOnA - tmpPointOnB).length() - margin;
            if (!isValid || (distance2 < distance))
            {
              distance = distance2;
              pointOnA = tmpPointOnA;
              pointOnB = tmpPointOnB;
              pointOnA -= cachedSeparatingAxis * marginA;
              pointOnB += cachedSeparatingAxis * marginB;
              normalInB = cachedSeparatingAxis;
              normalInB = normalInB / normalInB.length();
              isValid = true;
            }
          }
        }
      }
    }
  }

  if (isValid && ((distance < 0) || (distance * distance < maxDist2)))
  {
    double d2 = 0.0;
    {
      V3 pW = localTransA(A.supportNoMargin(localTransA.R.tmul(-orgNormalInB)));
      V3 qW = localTransB(B.supportNoMargin(localTransB.R.tmul(orgNormalInB)));
      d2 = orgNormalInB.dot(pW - qW) - margin;
    }
    double d1 = 0;
    {
      V3 pW = localTransA(A.supportNoMargin(localTransA.R.tmul(normalInB)));
      V3 qW = localTransB(B.supportNoMargin(localTransB.R.tmul(-normalInB)));
      d1 = (-normalInB).dot(pW - qW) - margin;
    }
    double d0 = 0.0;
    {
      V3 pW = localTransA(A.supportNoMargin(transA.R.tmul(-normalInB)));
      V3 qW = localTransB(B.supportNoMargin(transB.R.tmul(normalInB)));
      d0 = normalInB.dot(pW - qW) - margin;
    }
    cnt.gjk_iterations += 3;
    if (getenv("T4O_DEBUG"))
      fprintf(stderr, "[t4o] phase4 status=%d distance=%.9g d0=%.9g d1=%.9g d2=%.9g n=(%.6f %.6f %.6f) org=(%.6f %.6f %.6f)\n",
              status, distance, d0, d1, d2, normalInB.x, normalInB.y, normalInB.z, orgNormalInB.x, orgNormalInB.y,
              orgNormalInB.z);
    if (d1 > d0)
      normalInB *= -1;
    if (orgNormalInB.length2() > 0)
    {
      if (d2 > d0 && d2 > d1 && d2 > distance)
      {
        normalInB = orgNormalInB;
        distance = d2;
      }
    }
    out.has = true;
    out.normalOnB = normalInB;
    out.pointOnB = pointOnB + positionOffset;
    out.distance = distance;
  }
}

// tesseract_convex_convex_algorithm.cpp:67-181
void segmentsClosestPoints(V3& ptsVector, V3& offsetA, V3& offsetB, double& tA, double& tB, const V3& translation,
                           const V3& dirA, double hlenA, const V3& dirB, double hlenB)
{
  double dAdB = dirA.dot(dirB), dAt = dirA.dot(translation), dBt = dirB.dot(translation);
  double denom = 1.0 - dAdB * dAdB;
  if (denom == 0.0)
    tA = 0.0;
  else
  {
    tA = (dAt - dBt * dAdB) / denom;
    tA = std::min(std::max(tA, -hlenA), hlenA);
  }
  tB = tA * dAdB - dBt;
  if (tB < -hlenB)
  {
    tB = -hlenB;
    tA = tB * dAdB + dAt;
    tA = std::min(std::max(tA, -hlenA), hlenA);
  }
  else if (tB > hlenB)
  {
    tB = hlenB;
    tA = tB * dAdB + dAt;
    tA = std::min(std::max(tA, -hlenA), hlenA);
  }
  offsetA = dirA * tA;
  offsetB = dirB * tB;
  ptsVector = translation - offsetA + offsetB;
}

void planeSpace1(const V3& n, V3& p, V3& q)
{
  const double SQRT12 = 0.7071067811865475244008443621048490;
  if (std::fabs(n.z) > SQRT12)
  {
    double a = n.y * n.y + n.z * n.z;
    double k = 1.0 / std::sqrt(a);
    p = V3(0, -n.z * k, n.y * k);
    q = V3(a * k, -n.x * p.z, n.x * p.y);
  }
  else
  {
    double a = n.x * n.x + n.y * n.y;
    double k = 1.0 / std::sqrt(a);
    p = V3(-n.y * k, n.x * k, 0);
    q = V3(-n.z * p.y, n.z * p.x, a * k);
  }
}

double capsuleCapsuleDistance(V3& normalOnB, V3& pointOnB, double lenA, double rA, double lenB, double rB, int axisA,
                              int axisB, const Xf& ta, const Xf& tb, double distanceThreshold)
{
  V3 directionA = ta.R.col(axisA);
  V3 directionB = tb.R.col(axisB);
  V3 translation = tb.p - ta.p;
  V3 ptsVector, offsetA, offsetB;
  double tA, tB;
  segmentsClosestPoints(ptsVector, offsetA, offsetB, tA, tB, translation, directionA, lenA, directionB, lenB);
  double distance = ptsVector.length() - rA - rB;
  if (distance > distanceThreshold)
    return distance;
  double lenSqr = ptsVector.length2();
  if (lenSqr <= (SIMD_EPS * SIMD_EPS))
  {
    V3 q;
    planeSpace1(directionA, normalOnB, q);
  }
  else
    normalOnB = ptsVector * (-1.0 / std::sqrt(lenSqr));
  pointOnB = tb.p + offsetB + normalOnB * rB;
  return distance;
}

// Bullet SphereTriangleDetector::collide
double segmentSqrDistance(const V3& from, const V3& to, const V3& p, V3& nearest)
{
  V3 diff = p - from;
  V3 v = to - from;
  double t = v.dot(diff);
  if (t > 0)
  {
    double dotVV = v.dot(v);
    if (t < dotVV)
    {
      t /= dotVV;
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

bool pointInTriangle(const V3* vertices, const V3& normal, const V3& p)
{
  const V3 &p1 = vertices[0], &p2 = vertices[1], &p3 = vertices[2];
  V3 e1 = p2 - p1, e2 = p3 - p2, e3 = p1 - p3;
  double r1 = e1.cross(normal).dot(p - p1);
  double r2 = e2.cross(normal).dot(p - p2);
  double r3 = e3.cross(normal).dot(p - p3);
  return (r1 > 0 && r2 > 0 && r3 > 0) || (r1 <= 0 && r2 <= 0 && r3 <= 0);
}

bool sphereTriangleCollide(const V3& sphereCenter, double radius, const V3* vertices, V3& point, V3& resultNormal,
                           double& depth, double contactBreakingThreshold)
{
  double radiusWithThreshold = radius + contactBreakingThreshold;
  V3 normal = (vertices[1] - vertices[0]).cross(vertices[2] - vertices[0]);
  double l2 = normal.length2();
  bool hasContact = false;
  V3 contactPoint;
  if (l2 >= SIMD_EPS * SIMD_EPS)
  {
    normal /= std::sqrt(l2);
    V3 p1ToCentre = sphereCenter - vertices[0];
    double distanceFromPlane = p1ToCentre.dot(normal);
    if (distanceFromPlane < 0.0)
    {
      distanceFromPlane *= -1.0;
      normal *= -1.0;
    }
    if (distanceFromPlane < radiusWithThreshold)
    {
      if (pointInTriangle(vertices, normal, sphereCenter))
      {
        hasContact = true;
        contactPoint = sphereCenter - normal * distanceFromPlane;
      }
      else
      {
        double minDistSqr = radiusWithThreshold * radiusWithThreshold;
        V3 nearestOnEdge;
        for (int i = 0; i < 3; i++)
        {
          const V3& pa = vertices[i];
          const V3& pb = vertices[(i + 1) % 3];
          double distanceSqr = segmentSqrDistance(pa, pb, sphereCenter, nearestOnEdge);
          if (distanceSqr < minDistSqr)
          {
            minDistSqr = distanceSqr;
            hasContact = true;
            contactPoint = nearestOnEdge;
          }
        }
      }
    }
  }
  if (hasContact)
  {
    V3 contactToCentre = sphereCenter - contactPoint;
    double distanceSqr = contactToCentre.length2();
    if (distanceSqr < radiusWithThreshold * radiusWithThreshold)
    {
      if (distanceSqr > SIMD_EPS)
      {
        double distance = std::sqrt(distanceSqr);
        resultNormal = contactToCentre / contactToCentre.length();
        point = contactPoint;
        depth = -(radius - distance);
      }
      else
      {
        resultNormal = normal;
        point = contactPoint;
        depth = -radius;
      }
      return true;
    }
  }
  return false;
}
}  // namespace

constexpr double CONTACT_BREAKING_THRESHOLD = 0.02;  // Bullet gContactBreakingThreshold (flag cleared at
                                                     // bullet_discrete_bvh_manager.cpp:65-66)

void collideConvexConvex(const Convex& a, const Xf& ta, const Convex& b, const Xf& tb, double threshold,
                         const PairRequest& req, PairResult& out, Counters& cnt)
{
  out.has = false;
  cnt.candidates++;
  const bool a_plain = !a.is_cast, b_plain = !b.is_cast;
  // Bullet dispatch matrix for BT_CLOSEST_POINT_ALGORITHMS
  if (a_plain && b_plain && a.type == T4C_SPHERE && b.type == T4C_SPHERE)
  {
    V3 diff = ta.p - tb.p;
    double len = diff.length();
    if (len > (a.r + b.r + threshold))
      return;
    double dist = len - (a.r + b.r);
    V3 n(1, 0, 0);
    if (len > SIMD_EPS)
      n = diff / len;
    out.has = true;
    out.normalOnB = n;
    out.pointOnB = tb.p + n * b.r;
    out.distance = dist;
    return;
  }
  if (a_plain && b_plain && a.type == T4C_SPHERE && b.type == T4C_TRIANGLE)
  {
    Xf sphereInTr = tb.inverseTimes(ta);
    V3 point, normal;
    double depth = 0;
    if (sphereTriangleCollide(sphereInTr.p, a.r, b.verts, point, normal, depth, CONTACT_BREAKING_THRESHOLD + threshold))
    {
      out.has = true;
      out.normalOnB = tb.R * normal;
      out.pointOnB = tb(point);
      out.distance = depth;
    }
    return;
  }
  if (a_plain && b_plain && a.type == T4C_TRIANGLE && b.type == T4C_SPHERE)
  {
    // swapped: sphere is body1 (B)
    Xf sphereInTr = ta.inverseTimes(tb);
    V3 point, normal;
    double depth = 0;
    if (sphereTriangleCollide(sphereInTr.p, b.r, a.verts, point, normal, depth, CONTACT_BREAKING_THRESHOLD + threshold))
    {
      V3 normalOnTri = ta.R * normal;
      out.has = true;
      out.normalOnB = -normalOnTri;
      out.pointOnB = ta(point) + normalOnTri * depth;
      out.distance = depth;
    }
    return;
  }
  // capsule analytic cases (tesseract_convex_convex_algorithm.cpp:311-399)
  if (a_plain && b_plain &&
      ((a.type == T4C_CAPSULE && (b.type == T4C_CAPSULE || b.type == T4C_SPHERE)) ||
       (a.type == T4C_SPHERE && b.type == T4C_CAPSULE)))
  {
    double thr = CONTACT_BREAKING_THRESHOLD + threshold;
    V3 n, p;
    double lenA = a.type == T4C_CAPSULE ? a.h : 0.0;
    double lenB = b.type == T4C_CAPSULE ? b.h : 0.0;
    int axA = a.type == T4C_CAPSULE ? 2 : 1;
    int axB = b.type == T4C_CAPSULE ? 2 : 1;
    double dist = capsuleCapsuleDistance(n, p, lenA, a.r, lenB, b.r, axA, axB, ta, tb, thr);
    if (dist < thr)
    {
      out.has = true;
      out.normalOnB = n;
      out.pointOnB = p;
      out.distance = dist;
    }
    return;
  }
  double maxDist = a.margin() + b.margin() + CONTACT_BREAKING_THRESHOLD + threshold;
  gjkPairDetector(a, ta, b, tb, maxDist * maxDist, req, out, cnt);
}

}  // namespace t4o
