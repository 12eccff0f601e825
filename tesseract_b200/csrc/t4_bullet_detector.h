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
  int gjk_iters, epa_iters;  // support pairs of phases 1 + 2 + 4 (and of the penetration solver's seed GJK); EPA iterations
  int p1_iters;              // ... of which in phase 1
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
  DET_PHASE3 = 2,  // penetration / degenerate: run phase 3, then phase 4
  DET_BUDGET = 3   // the iteration budget ran out before phases 1-2 were decided: nothing is valid, run again without
};

// recentre both shapes about the midpoint of their origins (:694-698); returns the offset
template <class T>
T4_HD V3<T> recentre(ShapeW<T>& A, ShapeW<T>& B)
{
  const V3<T> positionOffset = (A.x0.p + B.x0.p) * T(0.5);
  A.x0.p -= positionOffset;
  B.x0.p -= positionOffset;
  if (A.cast)
    A.x1.p -= positionOffset;
  if (B.cast)
    B.x1.p -= positionOffset;
  return positionOffset;
}

// phase 2 (Bullet GJK with the Voronoi simplex solver, :836-971) as a resumable iteration
template <class T>
struct Phase2State
{
  VoronoiSimplexSolver<T> solver;
  V3<T> axis;  // cachedSeparatingAxis
  T squaredDistance;
  int curIter, degenerateSimplex;
  bool checkSimplex;
  T4_HD void reset()
  {
    solver.reset();
    axis = V3<T>(0, 1, 0);
    squaredDistance = Eps<T>::large();
    curIter = 0;
    degenerateSimplex = 0;
    checkSimplex = false;
  }
};

// consumes pWorld = support_A(-axis), qWorld = support_B(axis); returns true when the loop is over
template <class T>
T4_HDN bool phase2Step(Phase2State<T>& st, const V3<T>& pWorld, const V3<T>& qWorld, T maxDist2)
{
  const V3<T> w = pWorld - qWorld;
  const T delta = st.axis.dot(w);
  if ((delta > T(0)) && (delta * delta > st.squaredDistance * maxDist2))
  {
    st.degenerateSimplex = 10;
    st.checkSimplex = true;
    return true;
  }
  if (st.solver.inSimplex(w))
  {
    st.degenerateSimplex = 1;
    st.checkSimplex = true;
    return true;
  }
  const T f0 = st.squaredDistance - delta;
  const T f1 = st.squaredDistance * K<T>::relError2();
  if (f0 <= f1)
  {
    st.degenerateSimplex = (f0 <= T(0)) ? 2 : 11;
    st.checkSimplex = true;
    return true;
  }
  st.solver.addVertex(w, pWorld, qWorld);
  V3<T> newAxis;
  if (!st.solver.closest(newAxis))
  {
    st.degenerateSimplex = 3;
    st.checkSimplex = true;
    return true;
  }
  if (newAxis.len2() < K<T>::relError2())
  {
    st.axis = newAxis;
    st.degenerateSimplex = 6;
    st.checkSimplex = true;
    return true;
  }
  const T previousSquaredDistance = st.squaredDistance;
  st.squaredDistance = newAxis.len2();
  if (previousSquaredDistance - st.squaredDistance <= K<T>::simdEps() * previousSquaredDistance)
  {
    st.checkSimplex = true;
    st.degenerateSimplex = 12;
    return true;
  }
  st.axis = newAxis;
  if (st.curIter++ > 1000)
    return true;
  if (st.solver.fullSimplex())
  {
    st.degenerateSimplex = 13;
    return true;
  }
  return false;
}

template <class T>
T4_HD void resetDetectorState(DetectorState<T>& S)
{
  S.distance = 0;
  S.normalInB = V3<T>(0, 0, 0);
  S.pointOnA = V3<T>(0, 0, 0);
  S.pointOnB = V3<T>(0, 0, 0);
  S.orgNormalInB = V3<T>(0, 0, 0);
  S.isValid = 0;
  S.status = 0;
}

// witness points, margins, validity after the phase-2 loop (:973-1006); returns DET_PHASE3 or DET_PHASE4
template <class T>
T4_HDN int phase2Finish(Phase2State<T>& st, T marginA, T marginB, DetectorState<T>& S)
{
  const T margin = marginA + marginB;
  if (st.checkSimplex)
  {
    st.solver.compute_points(S.pointOnA, S.pointOnB);
    S.normalInB = st.axis;
    const T lenSqr = st.axis.len2();
    if (lenSqr < K<T>::relError2())
      st.degenerateSimplex = 5;
    if (lenSqr > K<T>::simdEps() * K<T>::simdEps())
    {
      const T rlen = T(1) / t4sqrt(lenSqr);
      S.normalInB = S.normalInB * rlen;
      const T s = t4sqrt(st.squaredDistance);
      S.pointOnA -= st.axis * (marginA / s);
      S.pointOnB += st.axis * (marginB / s);
      S.distance = (T(1) / rlen) - margin;
      S.isValid = 1;
      S.orgNormalInB = S.normalInB;
    }
  }
  const bool catchDegeneratePenetrationCase =
      (st.degenerateSimplex != 0) && ((S.distance + margin) < K<T>::penTol());
  if (!S.isValid || catchDegeneratePenetrationCase)
    return DET_PHASE3;
  return DET_PHASE4;
}

// phases 1 and 2 on RECENTRED shapes
template <class T>
T4_HDN int detectPhase12(const ShapeW<T>& A, const ShapeW<T>& B, const V3<T>& positionOffset, T maxDist2, bool calc_pen,
                        bool calc_dist, DetectorState<T>& S, DetectorOut<T>& out, int budget = 1 << 30,
                        bool skip_phase1 = false)
{
  out.has = false;
  out.gjk_iters = 0;
  out.epa_iters = 0;
  out.p1_iters = 0;
  out.used_epa = false;
  resetDetectorState(S);
  int it1 = 0;
  // skip_phase1: the caller holds a certificate that the cores are separated by more than any tolerance of the boolean
  // GJK (t4_kernels2.h cullBody), i.e. that phase 1 returns "separated"; its only output is that status
  const int status = skip_phase1 ? -1 : booleanGjk(A, B, it1, budget);
  S.status = status;
  out.gjk_iters += it1;
  out.p1_iters = it1;
  if (status == -3)
    return DET_BUDGET;

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
  // Phase 2 runs whatever phase 1 found (:836-1006).  For intersecting cores it usually ends on a degenerate simplex,
  // but it can leave a valid contact (isValid, distance ~ -margins, orgNormalInB): the fall-back when the penetration
  // solver produces nothing, and the candidate of the phase-4 normal override.

  Phase2State<T> st;
  st.reset();
  for (;;)
  {
    const V3<T> pWorld = supportW(A, -st.axis, false);
    const V3<T> qWorld = supportW(B, st.axis, false);
    ++out.gjk_iters;
    if (phase2Step(st, pWorld, qWorld, maxDist2))
      break;
    if (out.gjk_iters >= budget)
      return DET_BUDGET;
  }
  const int next = phase2Finish(st, A.margin(), B.margin(), S);
  return status == 0 ? DET_PHASE3 : next;  // "|| (status == 0)" at :1011
}

// phase 3 on RECENTRED shapes, in three pieces so that the polytope expansion in the middle can be run one iteration
// at a time (lane-refill EPA kernel): seed = FP64 GJK on the cores (simplex for EPA, or the separated answer) ...
template <class T>
T4_HDN int phase3Seed(const ShapeW<T>& A, const ShapeW<T>& B, GjkOut<T, true>& g)
{
  const V3<T> axis0 = A.x0.p - B.x0.p;
  return gjkDistance<T, true>(A, B, axis0, false, Eps<T>::large(), T(-1), g, 128);
}

// ... and the merge of its outcome into the detector state (the deeper result wins, :1024-1108).
// st = outcome of the seed; gv / gpA = its closest vector and witness on A (separated case); epa_ok / er = result of
// the expansion (intersecting case)
template <class T>
T4_HDN void phase3Apply(const ShapeW<T>& A, const ShapeW<T>& B, DetectorState<T>& S, DetectorOut<T>& out, int st,
                       const V3<T>& gv, const V3<T>& gpA, bool epa_ok, const EpaResult<T>& er)
{
  const T marginA = A.margin(), marginB = B.margin();
  const T margin = marginA + marginB;
  bool have2 = false;
  T distance2 = 0;
  V3<T> n2, pA2, pB2;
  if (st == GJK_INTERSECT)
  {
    if (epa_ok)
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
    const T dc = gv.len();
    if (dc > T(0))
    {
      const V3<T> n = gv * (T(1) / dc);
      pA2 = gpA - n * marginA;
      pB2 = (gpA - gv) + n * marginB;
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

template <class T, class EPA_T>
T4_HDN void detectPhase3(const ShapeW<T>& A, const ShapeW<T>& B, DetectorState<T>& S, DetectorOut<T>& out, EPA_T& epa)
{
  out.used_epa = true;
  GjkOut<T, true> g;
  const int st = phase3Seed(A, B, g);
  out.gjk_iters += g.iters;
  EpaResult<T> er;
  er.iterations = 0;
  bool ok = false;
  if (st == GJK_INTERSECT)
    ok = epa.run(A, B, g.sx.W, g.sx.P, g.sx.n, 255, T(1e-12), er);
  phase3Apply(A, B, S, out, st, g.v, g.pA, ok, er);
}

// phase 4 (:1111-1179): acceptance test ...
template <class T>
T4_HD bool phase4Accepts(const DetectorState<T>& S, T maxDist2)
{
  return S.isValid && ((S.distance < T(0)) || (S.distance * S.distance < maxDist2));
}
// ... and the normal-direction sanity decision from the three extra support values
// d2: along +-orgNormalInB, d1: along -+normalInB (flipped), d0: along +-normalInB (all minus the margins)
template <class T>
T4_HD void phase4Decide(const DetectorState<T>& S, T d0, T d1, T d2, const V3<T>& positionOffset, DetectorOut<T>& out)
{
  T distance = S.distance;
  V3<T> normalInB = S.normalInB;
  if (d1 > d0)
    normalInB = -normalInB;
  if (S.orgNormalInB.len2() > T(0))
  {
    if (d2 > d0 && d2 > d1 && d2 > distance)
    {
      normalInB = S.orgNormalInB;
      distance = d2;
    }
  }
  out.has = true;
  out.normalOnB = normalInB;
  out.pointOnB = S.pointOnB + positionOffset;
  out.distance = distance;
}

// phase 4 on RECENTRED shapes: acceptance + normal-direction sanity
template <class T>
T4_HDN void detectPhase4(const ShapeW<T>& A, const ShapeW<T>& B, const V3<T>& positionOffset, T maxDist2,
                        DetectorState<T>& S, DetectorOut<T>& out)
{
  const T margin = A.margin() + B.margin();
  if (!phase4Accepts(S, maxDist2))
    return;
  const V3<T> normalInB = S.normalInB, orgNormalInB = S.orgNormalInB;
  const T d2 = orgNormalInB.dot(supportW(A, -orgNormalInB, false) - supportW(B, orgNormalInB, false)) - margin;
  const T d1 = (-normalInB).dot(supportW(A, normalInB, false) - supportW(B, -normalInB, false)) - margin;
  const T d0 = normalInB.dot(supportW(A, -normalInB, false) - supportW(B, normalInB, false)) - margin;
  out.gjk_iters += 3;
  phase4Decide(S, d0, d1, d2, positionOffset, out);
}

}  // namespace bt
}  // namespace t4
#endif
