/*
 * t4_bullet_ccd.h -- phase 1 of TesseractGjkPairDetector: the libccd-style boolean GJK
 * (tesseract_collision/src/bullet/tesseract_gjk_pair_detector.cpp:136-676 simplex helpers, :732-821 driver).
 * Its only output is "the cores intersect" (0) / "separated" (-1), which decides whether the penetration solver runs.
 */
#ifndef T4_BULLET_CCD_H_
#define T4_BULLET_CCD_H_

#include "t4_bullet.h"

namespace t4
{
namespace bt
{
template <class T>
struct CcdSimplex
{
  V3<T> ps[4];  // Minkowski points only (the witness copies are never read by the driver)
  int last;
  T4_HD int size() const { return last + 1; }
  T4_HD void add(const V3<T>& s) { ps[++last] = s; }
  T4_HD void set(int pos, const V3<T>& s) { ps[pos] = s; }
  T4_HD void setSize(int n) { last = n - 1; }
};

template <class T>
T4_HD int ccdEq(T a_, T b_)
{
  const T ab = t4abs(a_ - b_);
  if (t4abs(ab) < K<T>::simdEps())
    return 1;
  const T a = t4abs(a_), b = t4abs(b_);
  if (b > a)
    return ab < K<T>::simdEps() * b;
  return ab < K<T>::simdEps() * a;
}
template <class T>
T4_HD int vec3Eq(const V3<T>& a, const V3<T>& b)
{
  return ccdEq(a.x, b.x) && ccdEq(a.y, b.y) && ccdEq(a.z, b.z);
}
template <class T>
T4_HD int ccdSign(T val)
{
  if (fuzzyZero(val))
    return 0;
  return val < 0 ? -1 : 1;
}
template <class T>
T4_HD V3<T> tripleCross(const V3<T>& a, const V3<T>& b, const V3<T>& c)
{
  return a.cross(b).cross(c);
}

template <class T>
T4_HDN T pointSegmentDist2(const V3<T>& P, const V3<T>& x0, const V3<T>& b, bool want_witness, V3<T>& witness)
{
  T dist, t;
  V3<T> d = b - x0;
  const V3<T> a = x0 - P;
  t = T(-1) * a.dot(d);
  t /= d.dot(d);
  if (t < 0 || fuzzyZero(t))
  {
    dist = (x0 - P).len2();
    witness = x0;
  }
  else if (t > 1 || ccdEq(t, T(1)))
  {
    dist = (b - P).len2();
    witness = b;
  }
  else
  {
    if (want_witness)
    {
      witness = d * t + x0;
      dist = (witness - P).len2();
    }
    else
    {
      d = d * t + a;
      dist = d.dot(d);
    }
  }
  return dist;
}

// __ccdVec3PointTriDist2 without a witness (tesseract_gjk_pair_detector.cpp:320-405)
template <class T>
T4_HDN T pointTriDist2(const V3<T>& P, const V3<T>& x0, const V3<T>& B, const V3<T>& C)
{
  const V3<T> d1 = B - x0, d2 = C - x0, a = x0 - P;
  const T u = a.dot(a), v = d1.dot(d1), w = d2.dot(d2), p = a.dot(d1), q = a.dot(d2), r = d1.dot(d2);
  const T s = (q * r - w * p) / (w * v - r * r);
  const T t = (-s * r - q) / w;
  T dist, dist2;
  V3<T> wit;
  if ((fuzzyZero(s) || s > 0) && (ccdEq(s, T(1)) || s < 1) && (fuzzyZero(t) || t > 0) && (ccdEq(t, T(1)) || t < 1) &&
      (ccdEq(t + s, T(1)) || (t + s) < 1))
  {
    dist = s * s * v;
    dist += t * t * w;
    dist += T(2) * s * t * r;
    dist += T(2) * s * p;
    dist += T(2) * t * q;
    dist += u;
  }
  else
  {
    dist = pointSegmentDist2(P, x0, B, false, wit);
    dist2 = pointSegmentDist2(P, x0, C, true, wit);
    if (dist2 < dist)
      dist = dist2;
    dist2 = pointSegmentDist2(P, B, C, true, wit);
    if (dist2 < dist)
      dist = dist2;
  }
  return dist;
}

template <class T>
T4_HDN int doSimplex2(CcdSimplex<T>& s, V3<T>& dir)
{
  const V3<T> A = s.ps[s.last];
  const V3<T> B = s.ps[0];
  const V3<T> AB = B - A;
  const V3<T> AO = -A;
  const T dot = AB.dot(AO);
  const V3<T> tmp = AB.cross(AO);
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

template <class T>
T4_HDN int doSimplex3(CcdSimplex<T>& s, V3<T>& dir)
{
  const V3<T> A = s.ps[s.last];
  const V3<T> B = s.ps[1];
  const V3<T> C = s.ps[0];
  const T dist = pointTriDist2(V3<T>(0, 0, 0), A, B, C);
  if (fuzzyZero(dist))
    return 1;
  if (vec3Eq(A, B) || vec3Eq(A, C))
    return -1;
  const V3<T> AO = -A;
  const V3<T> AB = B - A, AC = C - A;
  const V3<T> ABC = AB.cross(AC);
  V3<T> tmp = ABC.cross(AC);
  T dot = tmp.dot(AO);
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
        const V3<T> t = C;
        s.set(0, B);
        s.set(1, t);
        dir = -ABC;
      }
    }
  }
  return 0;
}

template <class T>
T4_HDN int doSimplex4(CcdSimplex<T>& s, V3<T>& dir)
{
  const V3<T> A = s.ps[s.last];
  const V3<T> B = s.ps[2];
  const V3<T> C = s.ps[1];
  const V3<T> D = s.ps[0];
  const T dist = pointTriDist2(A, B, C, D);
  if (fuzzyZero(dist))
    return -1;
  const V3<T> O(0, 0, 0);
  if (fuzzyZero(pointTriDist2(O, A, B, C)))
    return 1;
  if (fuzzyZero(pointTriDist2(O, A, C, D)))
    return 1;
  if (fuzzyZero(pointTriDist2(O, A, B, D)))
    return 1;
  if (fuzzyZero(pointTriDist2(O, B, C, D)))
    return 1;
  const V3<T> AO = -A;
  const V3<T> AB = B - A, AC = C - A, AD = D - A;
  const V3<T> ABC = AB.cross(AC), ACD = AC.cross(AD), ADB = AD.cross(AB);
  const int B_on_ACD = ccdSign(ACD.dot(AB));
  const int C_on_ADB = ccdSign(ADB.dot(AC));
  const int D_on_ABC = ccdSign(ABC.dot(AD));
  const int AB_O = ccdSign(ACD.dot(AO)) == B_on_ACD;
  const int AC_O = ccdSign(ADB.dot(AO)) == C_on_ADB;
  const int AD_O = ccdSign(ABC.dot(AO)) == D_on_ABC;
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

template <class T>
T4_HDN int doSimplex(CcdSimplex<T>& s, V3<T>& dir)
{
  if (s.size() == 2)
    return doSimplex2(s, dir);
  if (s.size() == 3)
    return doSimplex3(s, dir);
  return doSimplex4(s, dir);
}

// phase 1 (tesseract_gjk_pair_detector.cpp:732-821) as a resumable iteration: the state between two support
// evaluations.  status: 0 = cores intersect, -1 = separated, -2 = undecided after 1000 iterations.
template <class T>
struct Phase1State
{
  CcdSimplex<T> simplex;
  V3<T> dir;
  int status, it, started;
  T4_HD void reset()
  {
    simplex.last = -1;
    dir = V3<T>(1, 0, 0);
    status = -2;
    it = 0;
    started = 0;
  }
};

// consumes w = support_A(dir) - support_B(-dir); returns true when phase 1 is over (st.status holds the outcome)
template <class T>
T4_HDN bool phase1Step(Phase1State<T>& st, const V3<T>& w)
{
  if (!st.started)
  {
    st.started = 1;
    st.simplex.add(w);
    st.dir = -w;
    return false;
  }
  if (w.dot(st.dir) < 0)
  {
    st.status = -1;
    return true;
  }
  st.simplex.add(w);
  const int res = doSimplex(st.simplex, st.dir);
  if (res == 1)
  {
    st.status = 0;
    return true;
  }
  else if (res == -1)
  {
    st.status = -1;
    return true;
  }
  if (fuzzyZero(st.dir.dot(st.dir)))
    st.status = -1;
  if (st.dir.len2() < K<T>::simdEps())
  {
    st.status = -1;
    return true;
  }
  return ++st.it >= 1000;
}

// phase 1 driver: 0 = cores intersect, -1 = separated (tesseract_gjk_pair_detector.cpp:732-821)
template <class T>
// budget: give up (return -3, nothing decided) after that many support pairs -- the caller queues the pair for a
// run without a budget (t4_kernels2.h gjkBody); the default never triggers (phase 1 stops at 1000 by itself)
T4_HDN int booleanGjk(const ShapeW<T>& A, const ShapeW<T>& B, int& iters, int budget = 1 << 30)
{
  Phase1State<T> st;
  st.reset();
  iters = 0;
  for (;;)
  {
    const V3<T> w = supportW(A, st.dir, false) - supportW(B, -st.dir, false);
    ++iters;
    if (phase1Step(st, w))
      break;
    if (iters >= budget)
      return -3;
  }
  return st.status;
}

}  // namespace bt
}  // namespace t4
#endif
