/*
 * t4_gjk.h -- GJK distance between two convex shapes (thread-per-pair), templated on the scalar type.
 *
 * What it computes is what TesseractGjkPairDetector::getClosestPointsNonVirtual converges to
 * (tesseract_collision/src/bullet/tesseract_gjk_pair_detector.cpp:836-1006, phase 2: GJK with a Voronoi-region
 * simplex solver, relative duality-gap termination REL_ERROR2 :36-42, witness points with the shape margins
 * subtracted along the axis :973-1006).  The closest-point-on-simplex routine is Ericson's Voronoi-region test,
 * the algorithm behind Bullet's btVoronoiSimplexSolver (un-vendored dependency).
 *
 * Two instantiations are used by the kernels:
 *   - float,  WITNESS=false : conservative culling (lower bound on the distance vs. the contact cut-off)
 *   - double, WITNESS=true  : the reported contact (distance / nearest points / normal within 1e-5 of the reference)
 */
#ifndef T4_GJK_H_
#define T4_GJK_H_

#include "t4_shapes.h"

namespace t4
{
template <class T>
struct SubSimplex
{
  V3<T> closest;
  T bary[4];
  unsigned used;  // bit i: vertex i supports the closest point
  bool degenerate;
};

template <class T>
T4_HD void closestPtTriangle(const V3<T>& a, const V3<T>& b, const V3<T>& c, SubSimplex<T>& r)
{
  // closest point to the ORIGIN on triangle abc (Ericson, Real-Time Collision Detection 5.1.5)
  const V3<T> ab = b - a, ac = c - a, ap = -a;
  const T d1 = ab.dot(ap), d2 = ac.dot(ap);
  r.bary[3] = 0;
  if (d1 <= T(0) && d2 <= T(0))
  {
    r.closest = a;
    r.used = 1u;
    r.bary[0] = 1, r.bary[1] = 0, r.bary[2] = 0;
    return;
  }
  const V3<T> bp = -b;
  const T d3 = ab.dot(bp), d4 = ac.dot(bp);
  if (d3 >= T(0) && d4 <= d3)
  {
    r.closest = b;
    r.used = 2u;
    r.bary[0] = 0, r.bary[1] = 1, r.bary[2] = 0;
    return;
  }
  const T vc = d1 * d4 - d3 * d2;
  if (vc <= T(0) && d1 >= T(0) && d3 <= T(0))
  {
    const T v = d1 / (d1 - d3);
    r.closest = a + ab * v;
    r.used = 3u;
    r.bary[0] = 1 - v, r.bary[1] = v, r.bary[2] = 0;
    return;
  }
  const V3<T> cp = -c;
  const T d5 = ab.dot(cp), d6 = ac.dot(cp);
  if (d6 >= T(0) && d5 <= d6)
  {
    r.closest = c;
    r.used = 4u;
    r.bary[0] = 0, r.bary[1] = 0, r.bary[2] = 1;
    return;
  }
  const T vb = d5 * d2 - d1 * d6;
  if (vb <= T(0) && d2 >= T(0) && d6 <= T(0))
  {
    const T w = d2 / (d2 - d6);
    r.closest = a + ac * w;
    r.used = 5u;
    r.bary[0] = 1 - w, r.bary[1] = 0, r.bary[2] = w;
    return;
  }
  const T va = d3 * d6 - d5 * d4;
  if (va <= T(0) && (d4 - d3) >= T(0) && (d5 - d6) >= T(0))
  {
    const T w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    r.closest = b + (c - b) * w;
    r.used = 6u;
    r.bary[0] = 0, r.bary[1] = 1 - w, r.bary[2] = w;
    return;
  }
  const T denom = T(1) / (va + vb + vc);
  const T v = vb * denom, w = vc * denom;
  r.closest = a + ab * v + ac * w;
  r.used = 7u;
  r.bary[0] = 1 - v - w, r.bary[1] = v, r.bary[2] = w;
}

// 1: origin is outside plane abc (on the side away from d), 0: inside, -1: degenerate tetrahedron
template <class T>
T4_HD int originOutsidePlane(const V3<T>& a, const V3<T>& b, const V3<T>& c, const V3<T>& d)
{
  const V3<T> n = (b - a).cross(c - a);
  const T signp = (-a).dot(n);
  const T signd = (d - a).dot(n);
  // relative degeneracy test (Bullet uses an absolute 1e-8^2 on signd^2, tuned for unit-sized objects)
  // (relative volume below ~1e-9 (FP64) / 1e-4 (FP32): flat -- stopping there only ends GJK one step early with a
  // valid upper bound)
  const T scale = n.len2() * (d - a).len2();
  const T rel = sizeof(T) == 8 ? T(1e-18) : T(1e-8);
  if (signd * signd <= scale * rel)
    return -1;
  return signp * signd < T(0) ? 1 : 0;
}

// returns false if the origin is inside the tetrahedron (or it is degenerate: r.degenerate)
template <class T>
T4_HD bool closestPtTetrahedron(const V3<T>& a, const V3<T>& b, const V3<T>& c, const V3<T>& d, SubSimplex<T>& fr)
{
  SubSimplex<T> tmp;
  fr.closest = V3<T>(0, 0, 0);
  fr.used = 15u;
  fr.degenerate = false;
  const int oABC = originOutsidePlane(a, b, c, d);
  const int oACD = originOutsidePlane(a, c, d, b);
  const int oADB = originOutsidePlane(a, d, b, c);
  const int oBDC = originOutsidePlane(b, d, c, a);
  if (oABC < 0 || oACD < 0 || oADB < 0 || oBDC < 0)
  {
    fr.degenerate = true;
    return false;
  }
  if (!oABC && !oACD && !oADB && !oBDC)
    return false;
  T best = Eps<T>::large();
  if (oABC)
  {
    closestPtTriangle(a, b, c, tmp);
    const T sq = tmp.closest.len2();
    if (sq < best)
    {
      best = sq;
      fr.closest = tmp.closest;
      fr.used = (tmp.used & 1u) | (tmp.used & 2u) | (tmp.used & 4u);
      fr.bary[0] = tmp.bary[0], fr.bary[1] = tmp.bary[1], fr.bary[2] = tmp.bary[2], fr.bary[3] = 0;
    }
  }
  if (oACD)
  {
    closestPtTriangle(a, c, d, tmp);
    const T sq = tmp.closest.len2();
    if (sq < best)
    {
      best = sq;
      fr.closest = tmp.closest;
      fr.used = (tmp.used & 1u) | ((tmp.used & 2u) << 1) | ((tmp.used & 4u) << 1);
      fr.bary[0] = tmp.bary[0], fr.bary[1] = 0, fr.bary[2] = tmp.bary[1], fr.bary[3] = tmp.bary[2];
    }
  }
  if (oADB)
  {
    closestPtTriangle(a, d, b, tmp);
    const T sq = tmp.closest.len2();
    if (sq < best)
    {
      best = sq;
      fr.closest = tmp.closest;
      fr.used = (tmp.used & 1u) | ((tmp.used & 4u) >> 1) | ((tmp.used & 2u) << 2);
      fr.bary[0] = tmp.bary[0], fr.bary[1] = tmp.bary[2], fr.bary[2] = 0, fr.bary[3] = tmp.bary[1];
    }
  }
  if (oBDC)
  {
    closestPtTriangle(b, d, c, tmp);
    const T sq = tmp.closest.len2();
    if (sq < best)
    {
      best = sq;
      fr.closest = tmp.closest;
      fr.used = ((tmp.used & 1u) << 1) | ((tmp.used & 4u)) | ((tmp.used & 2u) << 2);
      fr.bary[0] = 0, fr.bary[1] = tmp.bary[0], fr.bary[2] = tmp.bary[2], fr.bary[3] = tmp.bary[1];
    }
  }
  return true;
}

// Simplex of Minkowski-difference points W = P - Q with the witness points P (on A) kept when WITNESS.
template <class T, bool WITNESS>
struct Simplex
{
  int n;
  V3<T> W[4];
  V3<T> P[WITNESS ? 4 : 1];
  T bary[4];

  T4_HD void reset() { n = 0; }
  T4_HD bool contains(const V3<T>& w) const
  {
    for (int i = 0; i < n; ++i)
      if (W[i] == w)
        return true;
    return false;
  }
  T4_HD void add(const V3<T>& w, const V3<T>& p)
  {
    W[n] = w;
    if (WITNESS)
      P[n] = p;
    ++n;
  }
  T4_HD void removeAt(int i)
  {
    --n;
    W[i] = W[n];
    if (WITNESS)
      P[i] = P[n];
    bary[i] = bary[n];
  }
  T4_HD void reduce(unsigned used)
  {
    if (n >= 4 && !(used & 8u))
      removeAt(3);
    if (n >= 3 && !(used & 4u))
      removeAt(2);
    if (n >= 2 && !(used & 2u))
      removeAt(1);
    if (n >= 1 && !(used & 1u))
      removeAt(0);
  }
  // closest point of the simplex to the origin; drops the vertices that do not support it.
  // returns 0: ok (v = closest), 1: origin enclosed (v = 0, simplex kept with 4 vertices), 2: degenerate
  T4_HDN int closest(V3<T>& v)
  {
    SubSimplex<T> bc;
    bc.degenerate = false;
    switch (n)
    {
      case 1:
        v = W[0];
        bary[0] = 1;
        return 0;
      case 2:
      {
        const V3<T> from = W[0];
        const V3<T> d = W[1] - from;
        T t = d.dot(-from);
        unsigned used;
        if (t > T(0))
        {
          const T dd = d.dot(d);
          if (t < dd)
          {
            t /= dd;
            used = 3u;
          }
          else
          {
            t = 1;
            used = 2u;
          }
        }
        else
        {
          t = 0;
          used = 1u;
        }
        v = from + d * t;
        bary[0] = 1 - t;
        bary[1] = t;
        reduce(used);
        return 0;
      }
      case 3:
        closestPtTriangle(W[0], W[1], W[2], bc);
        v = bc.closest;
        bary[0] = bc.bary[0], bary[1] = bc.bary[1], bary[2] = bc.bary[2];
        reduce(bc.used);
        return 0;
      case 4:
      {
        const bool sep = closestPtTetrahedron(W[0], W[1], W[2], W[3], bc);
        if (sep)
        {
          v = bc.closest;
          bary[0] = bc.bary[0], bary[1] = bc.bary[1], bary[2] = bc.bary[2], bary[3] = bc.bary[3];
          reduce(bc.used);
          return 0;
        }
        v = V3<T>(0, 0, 0);
        return bc.degenerate ? 2 : 1;
      }
      default:
        return 2;
    }
  }
  T4_HD V3<T> witnessA() const
  {
    V3<T> p(0, 0, 0);
    if (WITNESS)
      for (int i = 0; i < n; ++i)
        p += P[i] * bary[i];
    return p;
  }
};

enum
{
  GJK_FAR = 0,        // proven: core distance > cutoff
  GJK_SEPARATED = 1,  // converged, cores disjoint: v = closest vector (B -> A), |v| = core distance
  GJK_INTERSECT = 2,  // cores touch or overlap (simplex may enclose the origin)
  GJK_NEAR = 3        // (cull mode only) upper bound already <= accept: cores within `accept`
};

template <class T, bool WITNESS>
struct GjkOut
{
  V3<T> v;     // closest vector of the cores, from B to A
  V3<T> pA;    // witness on core A (WITNESS only)
  int iters;
  bool separated_by_1mm;  // some support plane proved the cores more than 1e-3 apart (a lower bound, not the distance)
  T lower2;               // best proven lower bound of the core distance, squared (0: none)
  Simplex<T, WITNESS> sx;
};

/*
 * cutoff : give up (GJK_FAR) as soon as the separating-axis lower bound exceeds it
 *          (tesseract_gjk_pair_detector.cpp:855-861, exit code 10; here with the true |v| instead of the running
 *          squared distance so that it can already fire on the first support)
 * accept : (cull mode, accept >= 0) stop with GJK_NEAR once |v| <= accept -- the pair is certainly within range
 * certify: (with accept) a pair that is within range keeps iterating, for at most `certify` more support pairs, until a
 *          support plane shows that the cores are more than 1 mm apart (out.separated_by_1mm), the simplex encloses
 *          the origin (GJK_INTERSECT) or the iteration converges: |v| <= accept is an UPPER bound of the distance, it
 *          does not tell a pair with separated cores (phases 1-2 of the detector) from a penetrating one (EPA)
 */
template <class T, bool WITNESS>
T4_HD int gjkDistance(const ShapeW<T>& A, const ShapeW<T>& B, V3<T> axis, bool with_margin, T cutoff, T accept,
                      GjkOut<T, WITNESS>& out, int max_iter, int certify = 0)
{
  Simplex<T, WITNESS>& sx = out.sx;
  bool near = false;
  sx.reset();
  V3<T> v = axis;
  if (v.len2() < Eps<T>::eps())
    v = V3<T>(0, 1, 0);
  T sqd = Eps<T>::large();
  bool have_v = false;  // v is a point of the Minkowski difference (after the first closest())
  int status = GJK_SEPARATED;
  int it = 0;
  out.separated_by_1mm = false;
  out.lower2 = T(0);
  for (; it < max_iter; ++it)
  {
    const V3<T> p = supportW(A, -v, with_margin);
    const V3<T> q = supportW(B, v, with_margin);
    const V3<T> w = p - q;
    const T delta = v.dot(w);
    const T vv = v.len2();
    // v.x >= delta for every point x of A - B: the cores are at least delta / |v| apart
    if (delta > T(0) && delta * delta > vv * T(1e-6))
      out.separated_by_1mm = true;
    if (delta > T(0) && vv > T(0))
      out.lower2 = t4max(out.lower2, delta * delta / vv);
    if (near && (out.separated_by_1mm || --certify < 0))
      break;
    if (delta > T(0) && delta * delta > vv * cutoff * cutoff)
    {
      status = GJK_FAR;
      break;
    }
    if (have_v)
    {
      if (sx.contains(w))
        break;  // support already in the simplex: converged exactly (polytopes)
      const T f0 = sqd - delta;
      if (f0 <= sqd * Eps<T>::rel2())
        break;  // duality gap closed
    }
    sx.add(w, p);
    V3<T> nv;
    const int c = sx.closest(nv);
    if (c == 1)
    {
      v = nv;
      status = GJK_INTERSECT;
      break;
    }
    if (c == 2)
    {
      // degenerate tetrahedron: keep the previous axis (code 3 of the reference); flat simplex through the origin
      // neighbourhood means touching -> let the penetration path decide
      status = have_v ? GJK_SEPARATED : GJK_INTERSECT;
      if (have_v)
        sx.removeAt(sx.n - 1), sx.closest(nv);
      break;
    }
    const T nsq = nv.len2();
    if (nsq <= Eps<T>::touch2())
    {
      v = nv;
      status = GJK_INTERSECT;
      break;
    }
    if (have_v && sqd - nsq <= Eps<T>::eps() * sqd)
    {
      // no progress (code 12): the simplex now holds the new point, its closest point is the answer
      v = nv;
      sqd = nsq;
      break;
    }
    v = nv;
    sqd = nsq;
    have_v = true;
    if (accept >= T(0) && sqd <= accept * accept)
    {
      near = true;
      if (certify <= 0 || out.separated_by_1mm)
      {
        status = GJK_NEAR;
        break;
      }
    }
  }
  if (near && status == GJK_SEPARATED)
    status = GJK_NEAR;  // within range whatever else the extra iterations found out
  out.v = v;
  out.iters = it + 1;
  if (WITNESS && (status == GJK_SEPARATED))
    out.pA = sx.witnessA();
  if (status == GJK_SEPARATED && !have_v)
    status = GJK_INTERSECT;  // never produced a closest point (max_iter == 0 or immediate degeneracy)
  return status;
}

}  // namespace t4
#endif
