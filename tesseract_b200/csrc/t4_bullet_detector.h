/*
 * t4_bullet_detector.h -- TesseractGjkPairDetector::getClosestPointsNonVirtual restated for device code
 * (tesseract_collision/src/bullet/tesseract_gjk_pair_detector.cpp:683-1185).
 *
 *   phase 0  recentre both shapes about the midpoint of their origins            :694-698
 *   phase 1  boolean GJK (t4_bullet_ccd.h)                                        :732-821
 *   early-outs on calculate_penetration / calculate_distance                      :823-834
 *   phase 2  Bullet GJK, start axis (0,1,0), termination codes 1/2/3/6/10/11/12/13 :836-971, witnesses :973-1006
 *   phase 3  penetration depth when invalid / degenerate-near-zero / phase 1 says intersecting   :1008-1108
 *            Bullet's calcPenDepth (un-vendored btGjkEpaSolver2) is replaced by what it converges to:
 *              cores overlap   -> polytope expansion on the cores (t4_epa.h), depth + margins along the face normal
 *              cores disjoint  -> converged core distance minus the margins along the closest direction
 *            (both of Bullet's outcomes, "Penetration" and the "Distance" fall-back, reduce to these formulas)
 *   phase 4  accept iff d < 0 or d^2 < maxDist^2 ; normal-direction sanity with three extra support pairs  :1111-1179
 */
#ifndef T4_BULLET_DETECTOR_H_
#define T4_BULLET_DETECTOR_H_

#include "t4_bullet_ccd.h"
#include "t4_epa.h"

namespace t4
{
namespace bt
{
template <class T>
struct DetectorOut
{
  bool has;
  V3<T> normalOnB, pointOnB;
  T distance;
  int gjk_iters, epa_iters;
  bool used_epa;
};

// running state of one query between the phases (also what the GJK kernel hands to the EPA kernel)
template <class T>
struct DetectorState
{
  T distance;
  V3<T> normalInB, pointOnA, pointOnB, orgNormalInB;
  int isValid, status;
};

enum
{
  DET_DONE = 0,    // finished: `out` is final (possibly out.has == false)
  DET_PHASE4 = 1,  // phases 1-2 produced the answer: run phase 4
  DET_PHASE3 = 2   // penetration / degenerate: run phase 3, then phase 4
};

// recentre both shapes about the midpoint of their origins (:694-698); returns the offset
template <class T>
T4_HD V3<T> recentre(ShapeW<T>& A, ShapeW<T>& B)
{
  const V3<T> positionOffset = (A.x0.p + B.x0.p) * T(0.5);
  A.x0.p -= positionOffset;
  A.x1.p -= positionOffset;
  B.x0.p -= positionOffset;
  B.x1.p -= positionOffset;
  return positionOffset;
}

// phases 1 and 2 on RECENTRED shapes
template <class T>
T4_HDN int detectPhase12(const ShapeW<T>& A, const ShapeW<T>& B, const V3<T>& positionOffset, T maxDist2, bool calc_pen,
                        bool calc_dist, DetectorState<T>& S, DetectorOut<T>& out)
{
  out.has = false;
  out.gjk_iters = 0;
  out.epa_iters = 0;
  out.used_epa = false;
  S.distance = 0;
  S.normalInB = V3<T>(0, 0, 0);
  S.pointOnA = V3<T>(0, 0, 0);
  S.pointOnB = V3<T>(0, 0, 0);
  S.orgNormalInB = V3<T>(0, 0, 0);
  S.isValid = 0;
  const T marginA = A.margin(), marginB = B.margin();
  const T margin = marginA + marginB;
  V3<T> cachedSeparatingAxis(0, 1, 0);
  bool checkSimplex = false;
  int degenerateSimplex = 0;
  int it1 = 0;
  const int status = booleanGjk(A, B, it1);
  S.status = status;
  out.gjk_iters += it1;

  if (status == 0 && !calc_pen)
  {
    out.has = true;
    out.normalOnB = S.normalInB;
    out.pointOnB = positionOffset;
    out.distance = -(T)FLT_MIN;  // reference: -numeric_limits<btScalar>::min(); only the sign is meaningful
    return DET_DONE;
  }
  if (status == -1 && !calc_dist)
    return DET_DONE;
  if (status == 0)
    return DET_PHASE3;  // phase 2 is skipped when the cores intersect: the penetration result always supersedes it

  T squaredDistance = Eps<T>::large();
  T delta = 0;
  VoronoiSimplexSolver<T> solver;
  solver.reset();
  int curIter = 0;
  for (;;)
  {
    const V3<T> pWorld = supportW(A, -cachedSeparatingAxis, false);
    const V3<T> qWorld = supportW(B, cachedSeparatingAxis, false);
    ++out.gjk_iters;
    const V3<T> w = pWorld - qWorld;
    delta = cachedSeparatingAxis.dot(w);
    if ((delta > T(0)) && (delta * delta > squaredDistance * maxDist2))
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
    const T f0 = squaredDistance - delta;
    const T f1 = squaredDistance * K<T>::relError2();
    if (f0 <= f1)
    {
      degenerateSimplex = (f0 <= T(0)) ? 2 : 11;
      checkSimplex = true;
      break;
    }
    solver.addVertex(w, pWorld, qWorld);
    V3<T> newAxis;
    if (!solver.closest(newAxis))
    {
      degenerateSimplex = 3;
      checkSimplex = true;
      break;
    }
    if (newAxis.len2() < K<T>::relError2())
    {
      cachedSeparatingAxis = newAxis;
      degenerateSimplex = 6;
      checkSimplex = true;
      break;
    }
    const T previousSquaredDistance = squaredDistance;
    squaredDistance = newAxis.len2();
    if (previousSquaredDistance - squaredDistance <= K<T>::simdEps() * previousSquaredDistance)
    {
      checkSimplex = true;
      degenerateSimplex = 12;
      break;
    }
    cachedSeparatingAxis = newAxis;
    if (curIter++ > 1000)
      break;
    if (solver.fullSimplex())
    {
      degenerateSimplex = 13;
      break;
    }
  }
  if (checkSimplex)
  {
    solver.compute_points(S.pointOnA, S.pointOnB);
    S.normalInB = cachedSeparatingAxis;
    const T lenSqr = cachedSeparatingAxis.len2();
    if (lenSqr < K<T>::relError2())
      degenerateSimplex = 5;
    if (lenSqr > K<T>::simdEps() * K<T>::simdEps())
    {
      const T rlen = T(1) / t4sqrt(lenSqr);
      S.normalInB = S.normalInB * rlen;
      const T s = t4sqrt(squaredDistance);
      S.pointOnA -= cachedSeparatingAxis * (marginA / s);
      S.pointOnB += cachedSeparatingAxis * (marginB / s);
      S.distance = (T(1) / rlen) - margin;
      S.isValid = 1;
      S.orgNormalInB = S.normalInB;
    }
  }
  const bool catchDegeneratePenetrationCase = (degenerateSimplex != 0) && ((S.distance + margin) < K<T>::penTol());
  if (!S.isValid || catchDegeneratePenetrationCase)
    return DET_PHASE3;
  return DET_PHASE4;
}

// phase 3 on RECENTRED shapes
template <class T, class EPA_T>
T4_HDN void detectPhase3(const ShapeW<T>& A, const ShapeW<T>& B, DetectorState<T>& S, DetectorOut<T>& out, EPA_T& epa)
{
  const T marginA = A.margin(), marginB = B.margin();
  const T margin = marginA + marginB;
  out.used_epa = true;
  GjkOut<T, true> g;
  const V3<T> axis0 = A.x0.p - B.x0.p;
  const int st = gjkDistance<T, true>(A, B, axis0, false, Eps<T>::large(), T(-1), g, 128);
  out.gjk_iters += g.iters;
  bool have2 = false;
  T distance2 = 0;
  V3<T> n2, pA2, pB2;
  if (st == GJK_INTERSECT)
  {
    EpaResult<T> er;
    if (epa.run(A, B, g.sx.W, g.sx.P, g.sx.n, 255, T(1e-12), er))
    {
      out.epa_iters = er.iterations;
      const T depth = er.depth + margin;
      pA2 = er.pA + er.normal * marginA;
      pB2 = pA2 - er.normal * depth;
      n2 = -er.normal;
      distance2 = -depth;
      have2 = true;
    }
  }
  else if (st == GJK_SEPARATED)
  {
    const T dc = g.v.len();
    if (dc > T(0))
    {
      const V3<T> n = g.v * (T(1) / dc);
      pA2 = g.pA - n * marginA;
      pB2 = (g.pA - g.v) + n * marginB;
      n2 = n;
      distance2 = dc - margin;
      have2 = true;
    }
  }
  if (have2 && (!S.isValid || distance2 < S.distance))
  {
    S.distance = distance2;
    S.pointOnA = pA2;
    S.pointOnB = pB2;
    S.normalInB = n2;
    S.isValid = 1;
  }
}

// phase 4 on RECENTRED shapes: acceptance + normal-direction sanity
template <class T>
T4_HDN void detectPhase4(const ShapeW<T>& A, const ShapeW<T>& B, const V3<T>& positionOffset, T maxDist2,
                        DetectorState<T>& S, DetectorOut<T>& out)
{
  const T margin = A.margin() + B.margin();
  T distance = S.distance;
  V3<T> normalInB = S.normalInB;
  const V3<T> orgNormalInB = S.orgNormalInB;
  if (S.isValid && ((distance < T(0)) || (distance * distance < maxDist2)))
  {
    T d2 = 0;
    {
      const V3<T> pW = supportW(A, -orgNormalInB, false);
      const V3<T> qW = supportW(B, orgNormalInB, false);
      d2 = orgNormalInB.dot(pW - qW) - margin;
    }
    T d1 = 0;
    {
      const V3<T> pW = supportW(A, normalInB, false);
      const V3<T> qW = supportW(B, -normalInB, false);
      d1 = (-normalInB).dot(pW - qW) - margin;
    }
    T d0 = 0;
    {
      const V3<T> pW = supportW(A, -normalInB, false);
      const V3<T> qW = supportW(B, normalInB, false);
      d0 = normalInB.dot(pW - qW) - margin;
    }
    out.gjk_iters += 3;
    if (d1 > d0)
      normalInB = -normalInB;
    if (orgNormalInB.len2() > T(0))
    {
      if (d2 > d0 && d2 > d1 && d2 > distance)
      {
        normalInB = orgNormalInB;
        distance = d2;
      }
    }
    out.has = true;
    out.normalOnB = normalInB;
    out.pointOnB = S.pointOnB + positionOffset;
    out.distance = distance;
  }
}

// A, B: shapes with WORLD poses.  maxDist2 = (marginA + marginB + contactBreakingThreshold + max margin)^2  (:435-440 of
// tesseract_convex_convex_algorithm.cpp).  Single-call form (all phases in one thread).
template <class T, class EPA_T>
T4_HD void gjkPairDetector(const ShapeW<T>& A0, const ShapeW<T>& B0, T maxDist2, bool calc_pen, bool calc_dist,
                           DetectorOut<T>& out, EPA_T& epa)
{
  ShapeW<T> A = A0, B = B0;
  const V3<T> positionOffset = recentre(A, B);
  DetectorState<T> S;
  const int r = detectPhase12(A, B, positionOffset, maxDist2, calc_pen, calc_dist, S, out);
  if (r == DET_DONE)
    return;
  if (r == DET_PHASE3)
    detectPhase3(A, B, S, out, epa);
  detectPhase4(A, B, positionOffset, maxDist2, S, out);
}

}  // namespace bt
}  // namespace t4
#endif
