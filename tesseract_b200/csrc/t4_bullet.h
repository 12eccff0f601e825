/*
 * t4_bullet.h -- step-faithful device restatement of the reference's convex-convex distance query.
 *
 * The reported contact must agree with the reference's Bullet manager to 1e-5 m.  Bullet's GJK does NOT always run to
 * the exact closest points: its Voronoi simplex solver flags thin tetrahedra with an ABSOLUTE volume threshold and the
 * detector then stops with the current (upper-bound) distance; on centimetre-sized shapes that is a 1e-5..1e-4 m
 * effect.  Matching the reference therefore means following the same iteration: same start axis, same simplex
 * bookkeeping, same termination codes.  This header restates, for device code,
 *   phase 1  libccd-style boolean GJK            tesseract_collision/src/bullet/tesseract_gjk_pair_detector.cpp:407-676,732-821
 *   phase 2  Bullet GJK + btVoronoiSimplexSolver  tesseract_gjk_pair_detector.cpp:836-1006 (solver: Bullet)
 *   phase 4  normal-direction sanity              tesseract_gjk_pair_detector.cpp:1111-1179
 * with the BT_USE_DOUBLE_PRECISION constants (:36-42).  Phase 3 (penetration depth) is in t4_epa.h.
 */
#ifndef T4_BULLET_H_
#define T4_BULLET_H_

#include "t4_shapes.h"

namespace t4
{
namespace bt
{
template <class T>
struct K
{
  static T4_HD T simdEps() { return Eps<T>::eps(); }
  static T4_HD T relError2() { return T(1.0e-12); }
  static T4_HD T penTol() { return T(1.0e-12); }
  static T4_HD T equalVertex() { return T(1e-12); }
};

template <class T>
T4_HD bool fuzzyZero(T x)
{
  return t4abs(x) < K<T>::simdEps();
}

// ---------------------------------------------------------------------------------------------------------------
// btVoronoiSimplexSolver
// ---------------------------------------------------------------------------------------------------------------
template <class T>
struct SubSimplexResult
{
  V3<T> closest;
  bool used[4];
  T bary[4];
  bool degenerate;
  T4_HD void reset()
  {
    degenerate = false;
    setBary(0, 0, 0, 0);
    used[0] = used[1] = used[2] = used[3] = false;
  }
  T4_HD bool isValid() const { return bary[0] >= 0 && bary[1] >= 0 && bary[2] >= 0 && bary[3] >= 0; }
  T4_HD void setBary(T a, T b, T c, T d)
  {
    bary[0] = a;
    bary[1] = b;
    bary[2] = c;
    bary[3] = d;
  }
};

template <class T>
struct VoronoiSimplexSolver
{
  int n;
  V3<T> W[5], P[5], Q[5];
  V3<T> cachedP1, cachedP2, cachedV, lastW;
  bool cachedValid;
  SubSimplexResult<T> bc;
  bool needsUpdate;

  T4_HD void reset()
  {
    cachedValid = false;
    n = 0;
    needsUpdate = true;
    lastW = V3<T>(Eps<T>::large(), Eps<T>::large(), Eps<T>::large());
    bc.reset();
  }
  T4_HD void addVertex(const V3<T>& w, const V3<T>& p, const V3<T>& q)
  {
    lastW = w;
    needsUpdate = true;
    W[n] = w;
    P[n] = p;
    Q[n] = q;
    ++n;
  }
  T4_HD void removeVertex(int i)
  {
    --n;
    W[i] = W[n];
    P[i] = P[n];
    Q[i] = Q[n];
  }
  T4_HD void reduceVertices(const bool* used)
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
  T4_HD bool fullSimplex() const { return n == 4; }
  T4_HD bool inSimplex(const V3<T>& w) const
  {
    bool found = false;
    for (int i = 0; i < n; ++i)
      if ((W[i] - w).len2() <= K<T>::equalVertex())
      {
        found = true;
        break;
      }
    if (w == lastW)
      return true;
    return found;
  }

  T4_HDN static bool closestPtPointTriangle(const V3<T>& p, const V3<T>& a, const V3<T>& b, const V3<T>& c,
                                           SubSimplexResult<T>& r)
  {
    r.used[0] = r.used[1] = r.used[2] = r.used[3] = false;
    const V3<T> ab = b - a, ac = c - a, ap = p - a;
    const T d1 = ab.dot(ap), d2 = ac.dot(ap);
    if (d1 <= T(0) && d2 <= T(0))
    {
      r.closest = a;
      r.used[0] = true;
      r.setBary(1, 0, 0, 0);
      return true;
    }
    const V3<T> bp = p - b;
    const T d3 = ab.dot(bp), d4 = ac.dot(bp);
    if (d3 >= T(0) && d4 <= d3)
    {
      r.closest = b;
      r.used[1] = true;
      r.setBary(0, 1, 0, 0);
      return true;
    }
    const T vc = d1 * d4 - d3 * d2;
    if (vc <= T(0) && d1 >= T(0) && d3 <= T(0))
    {
      const T v = d1 / (d1 - d3);
      r.closest = a + ab * v;
      r.used[0] = r.used[1] = true;
      r.setBary(1 - v, v, 0, 0);
      return true;
    }
    const V3<T> cp = p - c;
    const T d5 = ab.dot(cp), d6 = ac.dot(cp);
    if (d6 >= T(0) && d5 <= d6)
    {
      r.closest = c;
      r.used[2] = true;
      r.setBary(0, 0, 1, 0);
      return true;
    }
    const T vb = d5 * d2 - d1 * d6;
    if (vb <= T(0) && d2 >= T(0) && d6 <= T(0))
    {
      const T w = d2 / (d2 - d6);
      r.closest = a + ac * w;
      r.used[0] = r.used[2] = true;
      r.setBary(1 - w, 0, w, 0);
      return true;
    }
    const T va = d3 * d6 - d5 * d4;
    if (va <= T(0) && (d4 - d3) >= T(0) && (d5 - d6) >= T(0))
    {
      const T w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
      r.closest = b + (c - b) * w;
      r.used[1] = r.used[2] = true;
      r.setBary(0, 1 - w, w, 0);
      return true;
    }
    const T denom = T(1) / (va + vb + vc);
    const T v = vb * denom, w = vc * denom;
    r.closest = a + ab * v + ac * w;
    r.used[0] = r.used[1] = r.used[2] = true;
    r.setBary(1 - v - w, v, w, 0);
    return true;
  }

  T4_HD static int pointOutsideOfPlane(const V3<T>& p, const V3<T>& a, const V3<T>& b, const V3<T>& c, const V3<T>& d)
  {
    const V3<T> normal = (b - a).cross(c - a);
    const T signp = (p - a).dot(normal);
    const T signd = (d - a).dot(normal);
    if (signd * signd < (T(1e-8) * T(1e-8)))  // Bullet's ABSOLUTE degeneracy threshold
      return -1;
    return signp * signd < T(0) ? 1 : 0;
  }

  T4_HDN static bool closestPtPointTetrahedron(const V3<T>& p, const V3<T>& a, const V3<T>& b, const V3<T>& c,
                                              const V3<T>& d, SubSimplexResult<T>& fr)
  {
    SubSimplexResult<T> tmp;
    fr.closest = p;
    fr.used[0] = fr.used[1] = fr.used[2] = fr.used[3] = true;
    const int oABC = pointOutsideOfPlane(p, a, b, c, d);
    const int oACD = pointOutsideOfPlane(p, a, c, d, b);
    const int oADB = pointOutsideOfPlane(p, a, d, b, c);
    const int oBDC = pointOutsideOfPlane(p, b, d, c, a);
    if (oABC < 0 || oACD < 0 || oADB < 0 || oBDC < 0)
    {
      fr.degenerate = true;
      return false;
    }
    if (!oABC && !oACD && !oADB && !oBDC)
      return false;
    T best = T(1.7976931348623157e308);
    if (sizeof(T) == 4)
      best = T(3.4028234e38);
    if (oABC)
    {
      closestPtPointTriangle(p, a, b, c, tmp);
      const V3<T> q = tmp.closest;
      const T sq = (q - p).dot(q - p);
      if (sq < best)
      {
        best = sq;
        fr.closest = q;
        fr.used[0] = tmp.used[0], fr.used[1] = tmp.used[1], fr.used[2] = tmp.used[2], fr.used[3] = false;
        fr.setBary(tmp.bary[0], tmp.bary[1], tmp.bary[2], 0);
      }
    }
    if (oACD)
    {
      closestPtPointTriangle(p, a, c, d, tmp);
      const V3<T> q = tmp.closest;
      const T sq = (q - p).dot(q - p);
      if (sq < best)
      {
        best = sq;
        fr.closest = q;
        fr.used[0] = tmp.used[0], fr.used[1] = false, fr.used[2] = tmp.used[1], fr.used[3] = tmp.used[2];
        fr.setBary(tmp.bary[0], 0, tmp.bary[1], tmp.bary[2]);
      }
    }
    if (oADB)
    {
      closestPtPointTriangle(p, a, d, b, tmp);
      const V3<T> q = tmp.closest;
      const T sq = (q - p).dot(q - p);
      if (sq < best)
      {
        best = sq;
        fr.closest = q;
        fr.used[0] = tmp.used[0], fr.used[1] = tmp.used[2], fr.used[2] = false, fr.used[3] = tmp.used[1];
        fr.setBary(tmp.bary[0], tmp.bary[2], 0, tmp.bary[1]);
      }
    }
    if (oBDC)
    {
      closestPtPointTriangle(p, b, d, c, tmp);
      const V3<T> q = tmp.closest;
      const T sq = (q - p).dot(q - p);
      if (sq < best)
      {
        best = sq;
        fr.closest = q;
        fr.used[0] = false, fr.used[1] = tmp.used[0], fr.used[2] = tmp.used[2], fr.used[3] = tmp.used[1];
        fr.setBary(0, tmp.bary[0], tmp.bary[2], tmp.bary[1]);
      }
    }
    return true;
  }

  T4_HDN bool updateClosestVectorAndPoints()
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
          const V3<T> from = W[0];
          const V3<T> to = W[1];
          V3<T> diff = V3<T>(0, 0, 0) - from;
          const V3<T> v = to - from;
          T t = v.dot(diff);
          if (t > 0)
          {
            const T dotVV = v.dot(v);
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
          closestPtPointTriangle(V3<T>(0, 0, 0), W[0], W[1], W[2], bc);
          cachedP1 = P[0] * bc.bary[0] + P[1] * bc.bary[1] + P[2] * bc.bary[2];
          cachedP2 = Q[0] * bc.bary[0] + Q[1] * bc.bary[1] + Q[2] * bc.bary[2];
          cachedV = cachedP1 - cachedP2;
          reduceVertices(bc.used);
          cachedValid = bc.isValid();
          break;
        }
        case 4:
        {
          const bool has_sep = closestPtPointTetrahedron(V3<T>(0, 0, 0), W[0], W[1], W[2], W[3], bc);
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
              cachedV = V3<T>(0, 0, 0);
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
  T4_HD bool closest(V3<T>& v)
  {
    const bool ok = updateClosestVectorAndPoints();
    v = cachedV;
    return ok;
  }
  T4_HD void compute_points(V3<T>& p1, V3<T>& p2)
  {
    updateClosestVectorAndPoints();
    p1 = cachedP1;
    p2 = cachedP2;
  }
};

}  // namespace bt
}  // namespace t4
#endif
