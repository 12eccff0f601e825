/*
 * t4_epa.h -- penetration depth (minimum translation) of two intersecting convex cores by polytope expansion.
 *
 * Computes what btGjkEpaPenetrationDepthSolver::calcPenDepth / btGjkEpaSolver2::Penetration converge to
 * (called at tesseract_collision/src/bullet/tesseract_gjk_pair_detector.cpp:1024-1032; Bullet itself is an
 * un-vendored dependency): the face of the Minkowski difference A-B closest to the origin.
 * Differences from a literal port, all result-preserving at convergence:
 *   - margins (sphere / capsule radii, also of cast shapes) are NOT put into the support: the penetration depth of
 *     two ball-inflated cores is depth(cores) + rA + rB along the same normal, so the polytope stays finite and the
 *     expansion terminates exactly for sphere / capsule / box / hull / triangle pairs;
 *   - iterative visible-set marking with an explicit horizon instead of Bullet's recursive silhouette walk
 *     (no device recursion, bounded stack);
 *   - fixed-capacity vertex / face pools in thread-local memory with slot reuse.
 */
#ifndef T4_EPA_H_
#define T4_EPA_H_

#include "t4_gjk.h"

namespace t4
{
template <class T>
struct EpaResult
{
  V3<T> normal;  // outward normal of the closest face of A-B (unit)
  T depth;       // distance of that face from the origin (>= 0)
  V3<T> pA;      // witness on core A (world frame of the inputs)
  int iterations;
  bool converged;
};

// caps = btGjkEpaSolver2's EPA_MAX_VERTICES / EPA_MAX_FACES (128 / 256)
template <class T, int MAXV = 128, int MAXF = 256>
struct Epa
{
  struct Face
  {
    V3<T> n;
    int16_t v[3];
    int16_t adj[3];  // (neighbour slot << 2) | index of the twin edge in the neighbour; -1: none
    int16_t pad[2];
  };
  static constexpr int HZC = 32;  // horizon edges whose new-face plane is kept from the flood fill
  V3<T> W[MAXV], P[MAXV];
  Face F[MAXF];
  // plane distance of every face slot, +large for a free / removed slot: the closest-face search (the most frequent
  // memory access of an iteration) reads one dense array instead of two fields of 48-byte records
  T Fd[MAXF];
  // visit stamp of every face slot (dense for the same reason: three look-ups per face of the flood fill)
  int16_t Fm[MAXF];
  int nv, nf_hi;         // vertices used; high-water mark of face slots
  int16_t free_list[MAXF];
  int n_free;
  int16_t stack[MAXF], visible[MAXF];
  int16_t hz_a[MAXF], hz_b[MAXF], hz_g[MAXF];  // horizon edge (a -> b) and the packed adjacency entry across it
  V3<T> hz_n[HZC];
  T hz_d[HZC];

  const ShapeW<T>* A;
  const ShapeW<T>* B;

  T4_HDN void support(const V3<T>& d, V3<T>& w, V3<T>& p) const
  {
    p = supportW(*A, d, false);
    w = p - supportW(*B, -d, false);
  }
  T4_HD static T det3(const V3<T>& a, const V3<T>& b, const V3<T>& c) { return a.dot(b.cross(c)); }

  T4_HDN bool pushSupport(const V3<T>& d)
  {
    if (nv >= MAXV)
      return false;
    support(d, W[nv], P[nv]);
    ++nv;
    return true;
  }
  // btGjkEpaSolver2 GJK::EncloseOrigin: grow the simplex (W[0..nv)) to a non-degenerate tetrahedron
  T4_HD bool enclose4() const { return t4abs(det3(W[0] - W[3], W[1] - W[3], W[2] - W[3])) > T(0); }
  T4_HDN bool enclose3()
  {
    const V3<T> n = (W[1] - W[0]).cross(W[2] - W[0]);
    if (n.len2() > T(0))
    {
      pushSupport(n);
      if (enclose4())
        return true;
      --nv;
      pushSupport(-n);
      if (enclose4())
        return true;
      --nv;
    }
    return false;
  }
  T4_HDN bool enclose2()
  {
    const V3<T> d = W[1] - W[0];
    for (int i = 0; i < 3; ++i)
    {
      V3<T> axis(0, 0, 0);
      axis.set(i, 1);
      const V3<T> p = d.cross(axis);
      if (p.len2() > T(0))
      {
        pushSupport(p);
        if (enclose3())
          return true;
        --nv;
        pushSupport(-p);
        if (enclose3())
          return true;
        --nv;
      }
    }
    return false;
  }
  T4_HDN bool enclose1()
  {
    for (int i = 0; i < 3; ++i)
    {
      V3<T> axis(0, 0, 0);
      axis.set(i, 1);
      pushSupport(axis);
      if (enclose2())
        return true;
      --nv;
      pushSupport(-axis);
      if (enclose2())
        return true;
      --nv;
    }
    return false;
  }

  T4_HD bool makeFace(int a, int b, int c, Face& f, T& d) const
  {
    f.v[0] = (int16_t)a, f.v[1] = (int16_t)b, f.v[2] = (int16_t)c;
    f.adj[0] = f.adj[1] = f.adj[2] = -1;
    const V3<T> n = (W[b] - W[a]).cross(W[c] - W[a]);
    const T l = n.len();
    if (!(l > T(1e-20)))
      return false;
    f.n = n * (T(1) / l);
    d = f.n.dot(W[a]);
    return true;
  }
  T4_HD int allocFace()
  {
    if (n_free > 0)
      return free_list[--n_free];
    if (nf_hi < MAXF)
      return nf_hi++;
    return -1;
  }
  T4_HD static int edgeIndex(const Face& g, int from, int to)
  {
    for (int e = 0; e < 3; ++e)
      if (g.v[e] == from && g.v[(e + 1) % 3] == to)
        return e;
    return -1;
  }

  // ---- iteration state (members, so that the expansion can be resumed one iteration at a time: the lane-refill
  //      EPA kernel advances every lane by ONE iteration per trip and refills finished lanes at once)
  int mark_, it_, max_iter_;
  T tol_, saved_d;
  V3<T> saved_n;
  int16_t saved_v[3];  // vertices of the face that gave the tightest bound (W / P are append-only: indices stay valid)
  bool have_saved, deteriorated;

  // simplex: up to 4 points of A-B (and their A witnesses) around the origin, from the GJK that found the overlap.
  // false: no enclosing tetrahedron could be built (no penetration result)
  T4_HDN bool begin(const ShapeW<T>& a, const ShapeW<T>& b, const V3<T>* sw, const V3<T>* sp, int rank, int max_iter,
                    T tol)
  {
    A = &a;
    B = &b;
    nv = 0;
    for (int i = 0; i < rank && i < 4; ++i)
    {
      W[i] = sw[i];
      P[i] = sp[i];
      ++nv;
    }
    bool ok = false;
    if (nv == 0)
    {
      pushSupport(V3<T>(1, 0, 0));
    }
    if (nv == 1)
      ok = enclose1();
    else if (nv == 2)
      ok = enclose2();
    else if (nv == 3)
      ok = enclose3();
    else
      ok = enclose4();
    if (!ok || nv != 4)
      return false;
    if (det3(W[0] - W[3], W[1] - W[3], W[2] - W[3]) < T(0))
    {
      V3<T> t = W[0];
      W[0] = W[1];
      W[1] = t;
      t = P[0];
      P[0] = P[1];
      P[1] = t;
    }
    nf_hi = 0;
    n_free = 0;
    const int tet[4][3] = { { 0, 1, 2 }, { 1, 0, 3 }, { 2, 1, 3 }, { 0, 2, 3 } };
    for (int i = 0; i < 4; ++i)
    {
      if (!makeFace(tet[i][0], tet[i][1], tet[i][2], F[i], Fd[i]))
        return false;
      Fm[i] = 0;
      ++nf_hi;
    }
    for (int i = 0; i < 4; ++i)
      for (int e = 0; e < 3; ++e)
        for (int j = 0; j < 4; ++j)
          if (j != i)
          {
            const int ge = edgeIndex(F[j], F[i].v[(e + 1) % 3], F[i].v[e]);
            if (ge >= 0)
              F[i].adj[e] = (int16_t)((j << 2) | ge);
          }
    mark_ = 0;
    it_ = 0;
    max_iter_ = max_iter;
    tol_ = tol;
    saved_d = -Eps<T>::large();
    have_saved = false;
    deteriorated = false;
    return true;
  }

  // one expansion iteration: 0 = continue, 1 = finished (call finish), 2 = failed (no result)
  T4_HDN int step()
  {
    if (it_ >= max_iter_)
      return 1;
    int best = -1;
    {
      T bestd = Eps<T>::large();
      for (int i = 0; i < nf_hi; ++i)
        if (Fd[i] < bestd)
        {
          bestd = Fd[i];
          best = i;
        }
    }
    if (best < 0)
      return 2;
    // The closest-face distance is a LOWER bound of the depth and can only grow.  On curved supports (cylinder /
    // cone rims) the faces shrink to rounding-noise size near convergence and the hull eventually corrupts: the
    // bound then drops.  Stop there and report the tightest bound recorded.
    if (Fd[best] + T(1e-12) < saved_d)
    {
      deteriorated = true;
      return 1;
    }
    if (Fd[best] >= saved_d)
    {
      saved_d = Fd[best];
      saved_n = F[best].n;
      for (int k = 0; k < 3; ++k)
        saved_v[k] = F[best].v[k];
      have_saved = true;
    }
    if (nv >= MAXV)
      return 1;
    V3<T> w, p;
    support(F[best].n, w, p);
    const T wdist = F[best].n.dot(w) - Fd[best];
    if (wdist <= tol_)
      return 1;
    const int wi = nv;
    W[nv] = w;
    P[nv] = p;
    ++nv;
    // visible set (flood fill from the closest face) and, in the same pass, its horizon: an edge whose neighbour
    // fails the visibility test is a horizon edge whatever is found later, and the faces are popped in the order the
    // two-pass formulation walks them, so the horizon comes out in the same order.  The plane of the new face over
    // each edge is computed here, while the edge's vertices are at hand, and kept for the first HZC edges.
    ++mark_;
    int sp_ = 0, nvis = 0, nh = 0, first_bad = -1;
    bool degenerate = false;
    const T tiny = T(1e-14) * (T(1) + w.len2());
    stack[sp_++] = (int16_t)best;
    Fm[best] = (int16_t)mark_;
    while (sp_ > 0)
    {
      const int f = stack[--sp_];
      visible[nvis++] = (int16_t)f;
      for (int e = 0; e < 3; ++e)
      {
        const int a = F[f].adj[e];
        if (a >= 0)
        {
          const int g = a >> 2;
          if (Fm[g] == mark_)
            continue;
          if (F[g].n.dot(w) - Fd[g] > T(1e-14))
          {
            Fm[g] = (int16_t)mark_;
            stack[sp_++] = (int16_t)g;
            continue;
          }
        }
        const int va = F[f].v[e], vb = F[f].v[(e + 1) % 3];
        const V3<T> wa = W[va];
        const V3<T> n = (W[vb] - wa).cross(w - wa);
        const T l = n.len();
        if (a >= 0 && l < tiny)
          degenerate = true;
        if (nh < MAXF)
        {
          hz_a[nh] = (int16_t)va;
          hz_b[nh] = (int16_t)vb;
          hz_g[nh] = (int16_t)a;
          if (nh < HZC)
          {
            if (l > T(1e-20))
            {
              const V3<T> fn = n * (T(1) / l);
              hz_n[nh] = fn;
              hz_d[nh] = fn.dot(wa);
            }
            else if (first_bad < 0)
              first_bad = nh;
          }
          ++nh;
        }
      }
    }
    bool cached = true;
    if (degenerate)
    {
      // rare: absorb outside neighbours that would give a zero-area new face, and rebuild the horizon
      cached = false;
      for (int guard = 0; guard < 64; ++guard)
      {
        nh = 0;
        bool grew = false;
        for (int k = 0; k < nvis && !grew; ++k)
        {
          const int f = visible[k];
          for (int e = 0; e < 3 && !grew; ++e)
          {
            const int a = F[f].adj[e];
            const int g = a >= 0 ? (a >> 2) : -1;
            if (g >= 0 && Fm[g] == mark_)
              continue;
            const int va = F[f].v[e], vb = F[f].v[(e + 1) % 3];
            const V3<T> n = (W[vb] - W[va]).cross(w - W[va]);
            if (g >= 0 && n.len() < tiny && nvis < MAXF)
            {
              Fm[g] = (int16_t)mark_;
              visible[nvis++] = (int16_t)g;
              grew = true;
              break;
            }
            if (nh < MAXF)
            {
              hz_a[nh] = (int16_t)va;
              hz_b[nh] = (int16_t)vb;
              hz_g[nh] = (int16_t)a;
              ++nh;
            }
          }
        }
        if (!grew)
          break;
      }
    }
    if (nh < 3)
    {
      --nv;
      return 1;
    }
    // capacity check before touching the hull
    if (nh > (MAXF - nf_hi) + n_free + nvis)
    {
      --nv;
      return 1;
    }
    for (int k = 0; k < nvis; ++k)
    {
      Fd[visible[k]] = Eps<T>::large();
      free_list[n_free++] = visible[k];
    }
    // new faces (a, b, w): edge0 = horizon edge, edge1 = (b,w), edge2 = (w,a)
    bool bad = false;
    int first_new = -1;
    // remember the slots of the new faces in `stack` (free again at this point)
    int nnew = 0;
    for (int k = 0; k < nh; ++k)
    {
      const int slot = allocFace();
      Face nf;
      T nd = 0;
      if (cached && k < HZC)
      {
        if (slot < 0 || k == first_bad)
        {
          bad = true;
          break;
        }
        nf.v[0] = hz_a[k], nf.v[1] = hz_b[k], nf.v[2] = (int16_t)wi;
        nf.n = hz_n[k];
        nd = hz_d[k];
      }
      else if (slot < 0 || !makeFace(hz_a[k], hz_b[k], wi, nf, nd))
      {
        bad = true;
        break;
      }
      const int a = hz_g[k];
      nf.adj[0] = (int16_t)a;
      nf.adj[1] = nf.adj[2] = -1;
      F[slot] = nf;
      Fm[slot] = 0;
      Fd[slot] = nd;
      stack[nnew++] = (int16_t)slot;
      if (first_new < 0)
        first_new = slot;
      if (a >= 0)
        F[a >> 2].adj[a & 3] = (int16_t)(slot << 2);  // twin of the new face's edge 0
    }
    if (bad)
      return 1;
    // new face k = (hz_a[k], hz_b[k], w): its edge 1 continues in the face that starts at hz_b[k], its edge 2 comes
    // from the face that ends at hz_a[k]; matched on the dense index arrays instead of the face records
    for (int i = 0; i < nnew; ++i)
    {
      const int16_t ai = hz_a[i], bi = hz_b[i];
      int16_t a1 = -1, a2 = -1;
      for (int j = 0; j < nnew; ++j)
      {
        if (hz_a[j] == bi)
          a1 = (int16_t)((stack[j] << 2) | 2);  // (b, w) is the twin of that face's (w, a)
        if (hz_b[j] == ai)
          a2 = (int16_t)((stack[j] << 2) | 1);  // (w, a) is the twin of that face's (b, w)
      }
      Face& fi = F[stack[i]];
      fi.adj[1] = a1;
      fi.adj[2] = a2;
    }
    ++it_;
    return 0;
  }

  T4_HDN bool run(const ShapeW<T>& a, const ShapeW<T>& b, const V3<T>* sw, const V3<T>* sp, int rank, int max_iter,
                 T tol, EpaResult<T>& res)
  {
    res.converged = false;
    res.iterations = 0;
    if (!begin(a, b, sw, sp, rank, max_iter, tol))
      return false;
    int r;
    while ((r = step()) == 0)
    {
    }
    if (r == 2)
      return false;
    return finish(res);
  }

  // result after the expansion stopped
  T4_HDN bool finish(EpaResult<T>& res)
  {
    res.converged = false;
    int best = -1;
    res.iterations = it_;
    // closest ALIVE face (after a break the stored `best` may have been freed)
    best = -1;
    {
      T bestd = Eps<T>::large();
      for (int i = 0; i < nf_hi; ++i)
        if (Fd[i] < bestd)
        {
          bestd = Fd[i];
          best = i;
        }
    }
    if (best < 0)
      return false;
    if (have_saved && (deteriorated || Fd[best] < saved_d))
    {
      // report the recorded tightest bound; foot-point weights on its own triangle (clamped)
      const V3<T> pr = saved_n * saved_d;
      const V3<T> sw0 = W[saved_v[0]], sw1 = W[saved_v[1]], sw2 = W[saved_v[2]];
      T c0 = saved_n.dot((sw1 - pr).cross(sw2 - pr));
      T c1 = saved_n.dot((sw2 - pr).cross(sw0 - pr));
      T c2 = saved_n.dot((sw0 - pr).cross(sw1 - pr));
      c0 = t4max(c0, T(0)), c1 = t4max(c1, T(0)), c2 = t4max(c2, T(0));
      const T sum = c0 + c1 + c2;
      if (!(sum > T(0)))
        return false;
      res.normal = saved_n;
      res.depth = saved_d;
      res.pA = P[saved_v[0]] * (c0 / sum) + P[saved_v[1]] * (c1 / sum) + P[saved_v[2]] * (c2 / sum);
      return true;
    }
    // A flat face of A-B is usually split into several coplanar triangles with the same plane distance: take the one
    // that actually contains the foot point of the origin (signed barycentric coordinates), otherwise the witness
    // blend below would extrapolate.
    const T dband = Fd[best] + T(1e-9) * (T(1) + t4abs(Fd[best]));
    T b0 = 0, b1 = 0, b2 = 0, best_min = -Eps<T>::large();
    int pick = best;
    for (int i = 0; i < nf_hi; ++i)
    {
      if (Fd[i] > dband)  // also skips free slots (+large)
        continue;
      const Face& f = F[i];
      const V3<T> pr = f.n * Fd[i];
      const T c0 = f.n.dot((W[f.v[1]] - pr).cross(W[f.v[2]] - pr));
      const T c1 = f.n.dot((W[f.v[2]] - pr).cross(W[f.v[0]] - pr));
      const T c2 = f.n.dot((W[f.v[0]] - pr).cross(W[f.v[1]] - pr));
      const T sum = c0 + c1 + c2;
      if (!(sum > T(0)))
        continue;
      const T m = t4min(c0, t4min(c1, c2)) / sum;
      if (m > best_min)
      {
        best_min = m;
        pick = i;
        b0 = c0 / sum, b1 = c1 / sum, b2 = c2 / sum;
      }
    }
    if (best_min == -Eps<T>::large())
      return false;
    // clamp tiny negative weights (foot point on an edge) and renormalise
    b0 = t4max(b0, T(0)), b1 = t4max(b1, T(0)), b2 = t4max(b2, T(0));
    {
      const T sum = b0 + b1 + b2;
      b0 /= sum, b1 /= sum, b2 /= sum;
    }
    const Face& o = F[pick];
    res.normal = o.n;
    res.depth = Fd[pick];
    res.pA = P[o.v[0]] * b0 + P[o.v[1]] * b1 + P[o.v[2]] * b2;
    return true;
  }
};

}  // namespace t4
#endif
