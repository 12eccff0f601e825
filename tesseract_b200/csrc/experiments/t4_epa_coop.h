/*
 * EXPERIMENT, NOT COMPILED INTO libt4c.so.  Group-cooperative EPA with the polytope in shared memory: built in round 2,
 * parity-green on the B200 (GPU parity suite + golden scenarios), and measured 3.7x SLOWER than the thread-per-task
 * kernel it was meant to replace (configs[1] FIRST: EPA 22.0 ms vs 6.0 ms per 1 Mi states; CLOSEST: 305 ms vs 66 ms;
 * profiles/r3b_bench_coop_epa_*.json).  Why: see DESIGN.md section 4 "EPA: what did not work".  Kept as a record of the
 * design; the task body and the kernels that drove it follow the class below.
 *
 * t4_epa_coop.h -- penetration depth (minimum translation) of two intersecting convex cores by polytope expansion, written
 * for a GROUP of G lanes that work on one task in lock step with the polytope in SHARED memory.
 *
 * Computes what btGjkEpaPenetrationDepthSolver::calcPenDepth / btGjkEpaSolver2::Penetration converge to
 * (called at tesseract_collision/src/bullet/tesseract_gjk_pair_detector.cpp:1024-1032; Bullet itself is an
 * un-vendored dependency): the face of the Minkowski difference A-B closest to the origin.  Caps are
 * Bullet's (128 vertices, 256 faces, 255 iterations).  Differences from a literal port, all result-preserving at
 * convergence:
 *   - margins (sphere / capsule radii, also of cast shapes) are NOT put into the support: the penetration depth of
 *     two ball-inflated cores is depth(cores) + rA + rB along the same normal, so the polytope stays finite and the
 *     expansion terminates exactly for sphere / capsule / box / hull / triangle pairs;
 *   - visible-set flood fill with an explicit horizon instead of Bullet's recursive silhouette walk.
 *
 * Why a group and shared memory.  With one task per thread and the pools in thread-local memory (round 1) the kernel
 * ran at 5.7 % of the FP64 pipe: 17 KB of pools per thread, 10 GB of DRAM traffic per million states, 18 warp-cycles of
 * long-scoreboard stall per issued instruction.  Here the pools of a task live in shared memory and the lanes of its
 * group share the parts of an iteration that are data parallel:
 *     closest-face search      every lane scans a stride of the plane distances, arg-min by shuffle (ties -> lowest
 *                              slot, the answer of the serial scan)
 *     visibility               every lane tests a stride of the faces against the new vertex
 *     flood fill + horizon     ONE lane, integer work on the flags (order of the horizon = order of the new faces)
 *     new faces                one horizon edge per lane: plane, slot, twin links
 * while the support mapping -- the scalar part -- runs on the group's first lane.  A warp instruction costs the same
 * issue slot whatever its number of active lanes and the B200 issues 64 FP64 lane-operations per SM and clock, so the
 * group is kept SMALL where the polytope is small (flat-faced pairs, median 7 iterations, ~20 faces: several tasks
 * per warp) and a full warp only where it is large (curved cores: 15-75 iterations, 100+ faces).
 *
 * Every result is independent of G: reductions are exact (min / max with index tie-break, no floating-point sums
 * across lanes), and the serial sections run on one lane.  The host simulator runs the same code with G = 1.
 */
#ifndef T4_EPA_COOP_H_
#define T4_EPA_COOP_H_

#include <limits.h>

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
  bool capped;   // stopped on a pool capacity (vertices / faces): a larger pool could have gone on
};

// ---------------------------------------------------------------------------------------------------------------
// lock-step group of G lanes.  G = 1: serial (host simulator).
// ---------------------------------------------------------------------------------------------------------------
template <int G>
struct Coop;

template <>
struct Coop<1>
{
  static constexpr int lanes = 1;
  int lane = 0;
  T4_HD void sync() const {}
  T4_HD void argMin(double&, int&) const {}
  T4_HD void argMax(double&, int&) const {}
  T4_HD bool any(bool p) const { return p; }
  T4_HD uint32_t bcast(uint32_t v) const { return v; }
};

#ifdef T4_DEVICE_BUILD
template <int G>
struct Coop
{
  static constexpr int lanes = G;
  int lane;
  unsigned mask;
  __device__ __forceinline__ Coop()
  {
    const int wl = threadIdx.x & 31;
    lane = wl % G;
    mask = G == 32 ? 0xffffffffu : (((1u << G) - 1u) << (wl - lane));
  }
  __device__ __forceinline__ void sync() const { __syncwarp(mask); }
  // (key, index) of the smallest key; equal keys: the lowest index -- what a serial ascending scan with '<' finds
  __device__ __forceinline__ void argMin(double& k, int& i) const
  {
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1)
    {
      const double ok = __shfl_xor_sync(mask, k, off, G);
      const int oi = __shfl_xor_sync(mask, i, off, G);
      if (ok < k || (ok == k && oi < i))
        k = ok, i = oi;
    }
  }
  __device__ __forceinline__ void argMax(double& k, int& i) const
  {
#pragma unroll
    for (int off = G / 2; off > 0; off >>= 1)
    {
      const double ok = __shfl_xor_sync(mask, k, off, G);
      const int oi = __shfl_xor_sync(mask, i, off, G);
      if (ok > k || (ok == k && oi < i))
        k = ok, i = oi;
    }
  }
  __device__ __forceinline__ bool any(bool p) const { return __any_sync(mask, p) != 0; }
  __device__ __forceinline__ uint32_t bcast(uint32_t v) const { return __shfl_sync(mask, v, 0, G); }  // from lane 0
};
#endif

// ---------------------------------------------------------------------------------------------------------------
// storage of one task (shared memory on the device)
// ---------------------------------------------------------------------------------------------------------------
template <int MAXV, int MAXF>
struct EpaPool
{
  static_assert(MAXV <= 255 && MAXF <= 4096, "index widths");
  V3<double> W[MAXV];     // vertices of the polytope (points of A-B)
  V3<double> Fn[MAXF];    // unit face normals
  double Fd[MAXF];        // plane distances; +large marks a free slot (the closest-face search reads only this array)
  uint8_t Fv[MAXF][4];    // vertex indices
  int16_t Fadj[MAXF][4];  // (neighbour slot << 2) | index of the twin edge in the neighbour; -1: none
  uint8_t vis[MAXF];      // per iteration: 1 = sees the new vertex, 2 = visited by the flood fill
  int16_t free_list[MAXF];
  int16_t visible[MAXF], stack[MAXF];
  uint8_t hz_a[MAXF], hz_b[MAXF];     // horizon edge a -> b
  int16_t hz_g[MAXF], hz_slot[MAXF];  // packed adjacency entry across the edge; slot of the new face over it
  int32_t nvis, nh, ctl, pad;
  ShapeW<double> A, B;  // the pair, recentred (read by the support mapping of every iteration)
};

template <int G, int MAXV, int MAXF>
struct EpaCoop
{
  typedef double T;
  EpaPool<MAXV, MAXF>& S;
  V3<T>* Pw;  // witness on A of every vertex (global scratch: written once per vertex, three entries read at the end)
  Coop<G> g;
  // replicated in every lane
  int nv, nf_hi, n_free, it, max_iter;
  T tol;
  bool capped;
  // lane 0 only
  T saved_d;
  V3<T> saved_n;
  int saved_v[3];
  bool have_saved, deteriorated;

  T4_KM EpaCoop(EpaPool<MAXV, MAXF>& pool, V3<T>* witnesses) : S(pool), Pw(witnesses) {}

  T4_KMN void support(const V3<T>& d, V3<T>& w, V3<T>& p) const
  {
    p = supportW(S.A, d, false);
    w = p - supportW(S.B, -d, false);
  }
  T4_KM static T det3(const V3<T>& a, const V3<T>& b, const V3<T>& c) { return a.dot(b.cross(c)); }

  // ---- btGjkEpaSolver2 GJK::EncloseOrigin on a local simplex: grow it to a non-degenerate tetrahedron (lane 0)
  struct Seed
  {
    V3<T> w[4], p[4];
    int n;
  };
  T4_KMN void push(Seed& sd, const V3<T>& d) const
  {
    support(d, sd.w[sd.n], sd.p[sd.n]);
    ++sd.n;
  }
  T4_KM static bool enclose4(const Seed& sd)
  {
    return t4abs(det3(sd.w[0] - sd.w[3], sd.w[1] - sd.w[3], sd.w[2] - sd.w[3])) > T(0);
  }
  T4_KMN bool enclose3(Seed& sd) const
  {
    const V3<T> n = (sd.w[1] - sd.w[0]).cross(sd.w[2] - sd.w[0]);
    if (n.len2() > T(0))
    {
      push(sd, n);
      if (enclose4(sd))
        return true;
      --sd.n;
      push(sd, -n);
      if (enclose4(sd))
        return true;
      --sd.n;
    }
    return false;
  }
  T4_KMN bool enclose2(Seed& sd) const
  {
    const V3<T> d = sd.w[1] - sd.w[0];
    for (int i = 0; i < 3; ++i)
    {
      V3<T> axis(0, 0, 0);
      axis.set(i, 1);
      const V3<T> p = d.cross(axis);
      if (p.len2() > T(0))
      {
        push(sd, p);
        if (enclose3(sd))
          return true;
        --sd.n;
        push(sd, -p);
        if (enclose3(sd))
          return true;
        --sd.n;
      }
    }
    return false;
  }
  T4_KMN bool enclose1(Seed& sd) const
  {
    for (int i = 0; i < 3; ++i)
    {
      V3<T> axis(0, 0, 0);
      axis.set(i, 1);
      push(sd, axis);
      if (enclose2(sd))
        return true;
      --sd.n;
      push(sd, -axis);
      if (enclose2(sd))
        return true;
      --sd.n;
    }
    return false;
  }

  T4_KM bool planeOf(int a, int b, int c, V3<T>& n, T& d) const
  {
    const V3<T> wa = S.W[a];
    const V3<T> r = (S.W[b] - wa).cross(S.W[c] - wa);
    const T l = r.len();
    if (!(l > T(1e-20)))
      return false;
    n = r * (T(1) / l);
    d = n.dot(wa);
    return true;
  }

  // simplex: up to 4 points of A-B (and their A witnesses) around the origin, from the GJK that found the overlap
  // (lane 0 holds them).  false: no enclosing tetrahedron could be built (no penetration result).  Collective.
  T4_KMN bool begin(const V3<T>* sw, const V3<T>* sp, int rank, int max_iterations, T tolerance)
  {
    if (g.lane == 0)
    {
      Seed sd;
      sd.n = 0;
      for (int i = 0; i < rank && i < 4; ++i)
      {
        sd.w[i] = sw[i];
        sd.p[i] = sp[i];
        ++sd.n;
      }
      bool ok = false;
      if (sd.n == 0)
        push(sd, V3<T>(1, 0, 0));
      if (sd.n == 1)
        ok = enclose1(sd);
      else if (sd.n == 2)
        ok = enclose2(sd);
      else if (sd.n == 3)
        ok = enclose3(sd);
      else
        ok = enclose4(sd);
      ok = ok && sd.n == 4;
      if (ok)
      {
        if (det3(sd.w[0] - sd.w[3], sd.w[1] - sd.w[3], sd.w[2] - sd.w[3]) < T(0))
        {
          V3<T> t = sd.w[0];
          sd.w[0] = sd.w[1];
          sd.w[1] = t;
          t = sd.p[0];
          sd.p[0] = sd.p[1];
          sd.p[1] = t;
        }
        for (int i = 0; i < 4; ++i)
        {
          S.W[i] = sd.w[i];
          Pw[i] = sd.p[i];
        }
        const int tet[4][3] = { { 0, 1, 2 }, { 1, 0, 3 }, { 2, 1, 3 }, { 0, 2, 3 } };
        for (int i = 0; i < 4 && ok; ++i)
        {
          ok = planeOf(tet[i][0], tet[i][1], tet[i][2], S.Fn[i], S.Fd[i]);
          for (int e = 0; e < 3; ++e)
          {
            S.Fv[i][e] = (uint8_t)tet[i][e];
            S.Fadj[i][e] = -1;
          }
        }
        for (int i = 0; i < 4; ++i)
          for (int e = 0; e < 3; ++e)
            for (int j = 0; j < 4; ++j)
              if (j != i)
                for (int ge = 0; ge < 3; ++ge)  // twin edge: same vertices, opposite direction
                  if (tet[j][ge] == tet[i][(e + 1) % 3] && tet[j][(ge + 1) % 3] == tet[i][e])
                    S.Fadj[i][e] = (int16_t)((j << 2) | ge);
      }
      S.ctl = ok ? 1 : 0;
      S.pad = 0;
      saved_d = -Eps<T>::large();
      have_saved = false;
      deteriorated = false;
    }
    g.sync();
    const bool ok = S.ctl != 0;
    nv = 4;
    nf_hi = 4;
    n_free = 0;
    it = 0;
    max_iter = max_iterations;
    tol = tolerance;
    capped = false;
    g.sync();
    return ok;
  }

  // the closest alive face: plane distance and slot (INT_MAX: none).  Collective; the same answer in every lane.
  T4_KM void closestFace(T& bd, int& bi) const
  {
    bd = Eps<T>::large();
    bi = INT_MAX;
    for (int i = g.lane; i < nf_hi; i += G)
    {
      const T d = S.Fd[i];
      if (d < bd)
      {
        bd = d;
        bi = i;
      }
    }
    g.argMin(bd, bi);
  }

  // one expansion iteration: 0 = continue, 1 = finished (call finish), 2 = failed (no result).  Collective.
  T4_KMN int step()
  {
    if (it >= max_iter)
      return 1;
    T bd;
    int bi;
    closestFace(bd, bi);
    if (bi == INT_MAX)
      return 2;
    if (g.lane == 0)
    {
      int ctl = 0;
      // The closest-face distance is a LOWER bound of the depth and can only grow.  On curved supports (cylinder /
      // cone rims) the faces shrink to rounding-noise size near convergence and the hull eventually corrupts: the
      // bound then drops.  Stop there and report the tightest bound recorded.
      if (bd + T(1e-12) < saved_d)
      {
        deteriorated = true;
        ctl = 1;
      }
      else
      {
        if (bd >= saved_d)
        {
          saved_d = bd;
          saved_n = S.Fn[bi];
          for (int k = 0; k < 3; ++k)
            saved_v[k] = S.Fv[bi][k];
          have_saved = true;
        }
        if (nv >= MAXV)
          ctl = 3;  // vertex pool full
        else
        {
          const V3<T> n = S.Fn[bi];
          V3<T> w, p;
          support(n, w, p);
          if (n.dot(w) - bd <= tol)
            ctl = 1;
          else
          {
            S.W[nv] = w;
            Pw[nv] = p;
          }
        }
      }
      S.ctl = ctl;
    }
    g.sync();
    const int ctl = S.ctl;
    if (ctl)
    {
      capped = capped || ctl == 3;
      return 1;
    }
    const int wi = nv;
    const V3<T> w = S.W[wi];
    ++nv;
    const T tiny = T(1e-14) * (T(1) + w.len2());
    // ---- which faces see the new vertex
    for (int i = g.lane; i < nf_hi; i += G)
    {
      const T d = S.Fd[i];
      S.vis[i] = (d < Eps<T>::large() && S.Fn[i].dot(w) - d > T(1e-14)) ? 1 : 0;
    }
    g.sync();
    // ---- flood fill from the closest face over the visible faces; an edge whose neighbour does not see the vertex is
    //      a horizon edge.  The order in which the horizon comes out fixes the slots of the new faces: one lane.
    if (g.lane == 0)
    {
      int sp = 0, nvis = 0, nh = 0;
      S.stack[sp++] = (int16_t)bi;
      S.vis[bi] = 2;
      while (sp > 0)
      {
        const int f = S.stack[--sp];
        S.visible[nvis++] = (int16_t)f;
        for (int e = 0; e < 3; ++e)
        {
          const int a = S.Fadj[f][e];
          if (a >= 0)
          {
            const int q = a >> 2;
            if (S.vis[q] == 2)
              continue;
            if (S.vis[q] == 1)
            {
              S.vis[q] = 2;
              S.stack[sp++] = (int16_t)q;
              continue;
            }
          }
          if (nh < MAXF)
          {
            S.hz_a[nh] = S.Fv[f][e];
            S.hz_b[nh] = S.Fv[f][(e + 1) % 3];
            S.hz_g[nh] = (int16_t)a;
            ++nh;
          }
          else
            nh = MAXF + 1;  // horizon longer than the pool: reported as a capacity stop below
        }
      }
      S.nvis = nvis;
      S.nh = nh;
    }
    g.sync();
    int nvis = S.nvis, nh = S.nh;
    if (nh > MAXF)
    {
      --nv;
      capped = true;
      return 1;
    }
    // ---- geometry of the new faces (one horizon edge per lane): zero-area faces over an edge that has a neighbour
    //      mean the neighbour is coplanar with the new vertex and must be absorbed (rare, serial)
    bool degenerate = false;
    for (int k = g.lane; k < nh; k += G)
    {
      const V3<T> wa = S.W[S.hz_a[k]];
      const T l = (S.W[S.hz_b[k]] - wa).cross(w - wa).len();
      if (S.hz_g[k] >= 0 && l < tiny)
        degenerate = true;
    }
    if (g.any(degenerate))
    {
      if (g.lane == 0)
      {
        for (int guard = 0; guard < 64; ++guard)
        {
          nh = 0;
          bool grew = false;
          for (int k = 0; k < nvis && !grew; ++k)
          {
            const int f = S.visible[k];
            for (int e = 0; e < 3 && !grew; ++e)
            {
              const int a = S.Fadj[f][e];
              const int q = a >= 0 ? (a >> 2) : -1;
              if (q >= 0 && S.vis[q] == 2)
                continue;
              const int va = S.Fv[f][e], vb = S.Fv[f][(e + 1) % 3];
              const V3<T> n = (S.W[vb] - S.W[va]).cross(w - S.W[va]);
              if (q >= 0 && n.len() < tiny && nvis >= MAXF)
                S.pad = 1;  // cannot absorb: pool full
              if (q >= 0 && n.len() < tiny && nvis < MAXF)
              {
                S.vis[q] = 2;
                S.visible[nvis++] = (int16_t)q;
                grew = true;
                break;
              }
              if (nh < MAXF)
              {
                S.hz_a[nh] = (uint8_t)va;
                S.hz_b[nh] = (uint8_t)vb;
                S.hz_g[nh] = (int16_t)a;
                ++nh;
              }
            }
          }
          if (!grew)
            break;
        }
        S.nvis = nvis;
        S.nh = nh;
      }
      g.sync();
      nvis = S.nvis;
      nh = S.nh;
      capped = capped || S.pad != 0;
    }
    bool bad = false;
    for (int k = g.lane; k < nh; k += G)
    {
      const V3<T> wa = S.W[S.hz_a[k]];
      const T l = (S.W[S.hz_b[k]] - wa).cross(w - wa).len();
      if (!(l > T(1e-20)))
        bad = true;
    }
    bad = g.any(bad);
    // nothing has been modified so far: a horizon that cannot be closed, or does not fit, ends the expansion with the
    // hull of the previous iteration intact
    const int nfree_tot = n_free + nvis;
    const bool no_room = nh > (MAXF - nf_hi) + nfree_tot;
    if (nh < 3 || bad || no_room)
    {
      --nv;
      capped = capped || no_room;
      return 1;
    }
    // ---- commit: the visible faces become free slots ...
    for (int k = g.lane; k < nvis; k += G)
    {
      const int f = S.visible[k];
      S.Fd[f] = Eps<T>::large();
      S.free_list[n_free + k] = (int16_t)f;
    }
    g.sync();
    // ... and every horizon edge (a, b) gets its face (a, b, w): edge 0 = the horizon edge, edge 1 = (b, w), edge 2 =
    // (w, a).  Slots are handed out in horizon order: the free list last-in first-out, then the high-water mark.
    for (int k = g.lane; k < nh; k += G)
    {
      const int slot = k < nfree_tot ? S.free_list[nfree_tot - 1 - k] : nf_hi + (k - nfree_tot);
      S.hz_slot[k] = (int16_t)slot;
      const int va = S.hz_a[k], vb = S.hz_b[k];
      const V3<T> wa = S.W[va];
      const V3<T> r = (S.W[vb] - wa).cross(w - wa);
      const V3<T> fn = r * (T(1) / r.len());
      S.Fn[slot] = fn;
      S.Fd[slot] = fn.dot(wa);
      S.Fv[slot][0] = (uint8_t)va, S.Fv[slot][1] = (uint8_t)vb, S.Fv[slot][2] = (uint8_t)wi;
      const int a = S.hz_g[k];
      S.Fadj[slot][0] = (int16_t)a;
      if (a >= 0)
        S.Fadj[a >> 2][a & 3] = (int16_t)(slot << 2);  // twin of the new face's edge 0
    }
    if (nh <= nfree_tot)
      n_free = nfree_tot - nh;
    else
    {
      n_free = 0;
      nf_hi += nh - nfree_tot;
    }
    g.sync();
    // new face k = (a_k, b_k, w): its edge 1 continues in the face that starts at b_k, its edge 2 comes from the face
    // that ends at a_k
    for (int k = g.lane; k < nh; k += G)
    {
      const int ai = S.hz_a[k], bk = S.hz_b[k];
      int16_t a1 = -1, a2 = -1;
      for (int j = 0; j < nh; ++j)
      {
        if (S.hz_a[j] == bk)
          a1 = (int16_t)((S.hz_slot[j] << 2) | 2);  // (b, w) is the twin of that face's (w, a)
        if (S.hz_b[j] == ai)
          a2 = (int16_t)((S.hz_slot[j] << 2) | 1);  // (w, a) is the twin of that face's (b, w)
      }
      const int slot = S.hz_slot[k];
      S.Fadj[slot][1] = a1;
      S.Fadj[slot][2] = a2;
    }
    g.sync();
    ++it;
    return 0;
  }

  // result after the expansion stopped.  Collective; `res` is meaningful in lane 0.
  T4_KMN bool finish(EpaResult<T>& res)
  {
    res.converged = false;
    res.iterations = it;
    res.capped = capped;
    T bd;
    int best;
    closestFace(bd, best);
    if (best == INT_MAX)
      return false;
    if (g.lane == 0)
      S.ctl = (have_saved && (deteriorated || bd < saved_d)) ? 1 : 0;
    g.sync();
    const bool use_saved = S.ctl != 0;
    g.sync();
    if (use_saved)
    {
      if (g.lane == 0)
      {
        // report the recorded tightest bound; foot-point weights on its own triangle (clamped)
        bool ok = false;
        const V3<T> pr = saved_n * saved_d;
        const V3<T> sw0 = S.W[saved_v[0]], sw1 = S.W[saved_v[1]], sw2 = S.W[saved_v[2]];
        T c0 = saved_n.dot((sw1 - pr).cross(sw2 - pr));
        T c1 = saved_n.dot((sw2 - pr).cross(sw0 - pr));
        T c2 = saved_n.dot((sw0 - pr).cross(sw1 - pr));
        c0 = t4max(c0, T(0)), c1 = t4max(c1, T(0)), c2 = t4max(c2, T(0));
        const T sum = c0 + c1 + c2;
        if (sum > T(0))
        {
          res.normal = saved_n;
          res.depth = saved_d;
          res.pA = Pw[saved_v[0]] * (c0 / sum) + Pw[saved_v[1]] * (c1 / sum) + Pw[saved_v[2]] * (c2 / sum);
          ok = true;
        }
        S.ctl = ok ? 1 : 0;
      }
      g.sync();
      const bool ok = S.ctl != 0;
      g.sync();
      return ok;
    }
    // A flat face of A-B is usually split into several coplanar triangles with the same plane distance: take the one
    // that actually contains the foot point of the origin (largest minimum barycentric coordinate; ties -> lowest
    // slot), otherwise the witness blend below would extrapolate.
    const T dband = bd + T(1e-9) * (T(1) + t4abs(bd));
    T best_min = -Eps<T>::large();
    int pick = INT_MAX;
    for (int i = g.lane; i < nf_hi; i += G)
    {
      const T d = S.Fd[i];
      if (d > dband)  // also skips free slots (+large)
        continue;
      const V3<T> n = S.Fn[i];
      const V3<T> pr = n * d;
      const V3<T> w0 = S.W[S.Fv[i][0]], w1 = S.W[S.Fv[i][1]], w2 = S.W[S.Fv[i][2]];
      const T c0 = n.dot((w1 - pr).cross(w2 - pr));
      const T c1 = n.dot((w2 - pr).cross(w0 - pr));
      const T c2 = n.dot((w0 - pr).cross(w1 - pr));
      const T sum = c0 + c1 + c2;
      if (!(sum > T(0)))
        continue;
      const T m = t4min(c0, t4min(c1, c2)) / sum;
      if (m > best_min)
      {
        best_min = m;
        pick = i;
      }
    }
    g.argMax(best_min, pick);
    if (pick == INT_MAX)
      return false;
    if (g.lane == 0)
    {
      const T d = S.Fd[pick];
      const V3<T> n = S.Fn[pick];
      const V3<T> pr = n * d;
      const int v0 = S.Fv[pick][0], v1 = S.Fv[pick][1], v2 = S.Fv[pick][2];
      const V3<T> w0 = S.W[v0], w1 = S.W[v1], w2 = S.W[v2];
      const T c0 = n.dot((w1 - pr).cross(w2 - pr));
      const T c1 = n.dot((w2 - pr).cross(w0 - pr));
      const T c2 = n.dot((w0 - pr).cross(w1 - pr));
      const T sum = c0 + c1 + c2;
      // clamp tiny negative weights (foot point on an edge) and renormalise
      T b0 = t4max(c0 / sum, T(0)), b1 = t4max(c1 / sum, T(0)), b2 = t4max(c2 / sum, T(0));
      const T bs = b0 + b1 + b2;
      b0 /= bs, b1 /= bs, b2 /= bs;
      res.normal = n;
      res.depth = d;
      res.pA = Pw[v0] * b0 + Pw[v1] * b1 + Pw[v2] * b2;
    }
    return true;
  }
};

// ---- the task body (was in t4_kernels2.h) ----------------------------------------------------------------------
#if 0
// One EPA task -- phase 3 (penetration) and phase 4 of the exact query -- worked on by a group of G lanes with the
// polytope in the shared-memory pool S (t4_epa.h).  The group's first lane does the scalar parts (seed GJK, support
// mapping, phase 4, result bridging); all lanes share the data-parallel parts of the expansion.
// FORWARD: this is the small-polytope tier -- tasks with a curved core (cylinder / cone: 15-75 iterations, 100+ faces)
// and tasks that run into the pool capacity are forwarded to the full-size tier through the retry queue.  A forwarded
// task is run again from the start there, so the result does not depend on the tiering.
template <int G, int MAXV, int MAXF, bool FORWARD>
T4_KD void epaTaskBody(uint32_t t, const KParams& P, EpaPool<MAXV, MAXF>& S, V3<double>* Pw)
{
  EpaCoop<G, MAXV, MAXF> epa(S, Pw);
  const bool leader = epa.g.lane == 0;
  // leader state
  uint32_t c = 0;
  int64_t s = 0;
  int a = 0, b = 0, st = GJK_SEPARATED;
  double maxDist2 = 0;
  V3<double> offset;
  bt::DetectorState<double> St;
  bt::DetectorOut<double> d;
  GjkOut<double, true> gj;
  gj.sx.n = 0;
  if (leader)
  {
    int run = 0;  // 0: nothing to do, 1: expand the polytope, 2: no expansion (the seed GJK decided), 3: forward
    const EpaTask& T = P.tasks[P.task_order ? P.task_order[t] : t];
    c = T.cand;
    s = P.cand_state[c];
    if (!((P.test_type == T4C_TEST_FIRST || P.hit_only) && P.hit[s]))
    {
      a = (int)P.cand_a[c], b = (int)P.cand_b[c];
      loadPairShapesD(P, s, a, b, S.A, S.B);
      if (FORWARD && (S.A.curvedCore() || S.B.curvedCore()))
        run = 3;
      else
      {
        maxDist2 = detectorMaxDist2(S.A, S.B, P.max_margin_d);
        offset = bt::recentre(S.A, S.B);
        St.isValid = T.is_valid;
        St.status = 0;
        St.distance = T.distance;
        St.normalInB = V3<double>(T.normalInB[0], T.normalInB[1], T.normalInB[2]);
        St.pointOnA = V3<double>(T.pointOnA[0], T.pointOnA[1], T.pointOnA[2]);
        St.pointOnB = V3<double>(T.pointOnB[0], T.pointOnB[1], T.pointOnB[2]);
        St.orgNormalInB = V3<double>(T.orgNormalInB[0], T.orgNormalInB[1], T.orgNormalInB[2]);
        d.has = false;
        d.gjk_iters = 0;
        d.epa_iters = 0;
        d.p1_iters = 0;
        d.used_epa = true;
        st = bt::phase3Seed(S.A, S.B, gj);
        d.gjk_iters += gj.iters;
        run = st == GJK_INTERSECT ? 1 : 2;
      }
    }
    S.ctl = run;
  }
  epa.g.sync();
  int run = S.ctl;
  epa.g.sync();
  EpaResult<double> er;
  er.iterations = 0;
  er.capped = false;
  bool ok = false;
  if (run == 1)
  {
    if (epa.begin(gj.sx.W, gj.sx.P, gj.sx.n, 255, 1e-12))
    {
      int r;
      while ((r = epa.step()) == 0)
      {
      }
      ok = r != 2 && epa.finish(er);
    }
    if (FORWARD && epa.capped)
      run = 3;
  }
  if (leader && run == 3)
  {
    const uint32_t k = qAppend(&P.qcount[Q_RETRY]);
    P.epa_retry[k] = t;  // sized for every task of the launch
  }
  else if (leader && run != 0)
  {
    bt::phase3Apply(S.A, S.B, St, d, st, gj.v, gj.pA, ok, er);
    bt::detectPhase4(S.A, S.B, offset, maxDist2, St, d);
    statAdd(P.stats, ST_EPA_GJK_ITERS, (uint32_t)d.gjk_iters);
    statAdd(P.stats, ST_EPA_PAIRS, 1u);
    statAdd(P.stats, ST_EPA_ITERS, (uint32_t)d.epa_iters);
    statAdd(P.stats, ST_SUPV64, S.A.evals + S.B.evals);
    statAdd(P.stats, ST_SUPV_EPA, S.A.evals + S.B.evals);
    PairResult<double> pr;
    copyOut(d, pr);
    if (pr.has)
    {
      if (P.cast)
      {
        // continuous data needs the shapes at their world poses (not recentred)
        ShapeW<double> A, B;
        loadPairShapesD(P, s, a, b, A, B);
        emitContact(P, s, a, b, A, B, pr);
      }
      else
        emitContact(P, s, a, b, S.A, S.B, pr);
    }
  }
  epa.g.sync();  // the pool is reused by the group's next task
}

#endif
// ---- the kernels and their launcher (were in t4c_exact.cu) ------------------------------------------------------
#if 0
// ---- EPA: groups of G lanes per task, polytope pools in shared memory (t4_epa.h).  Persistent grid: a group takes the
//      next task from a device counter as soon as it is done with the last one.
//      tier 1 (FORWARD): small pools, small groups -- flat-faced pairs (median 7 iterations, ~20 faces);
//      tier 2: Bullet's full caps (128 vertices / 256 faces), one warp per task -- curved cores and whatever overflowed
//      tier 1, read from the retry queue (or every task directly when there is no retry queue).
#ifndef T4_EPA_SMALL_G
#define T4_EPA_SMALL_G 4
#endif
#define T4_EPA_SMALL_V 20
#define T4_EPA_SMALL_F 36
#define T4_EPA_SMALL_THREADS 64
#define T4_EPA_BIG_THREADS 128

template <int G, int MAXV, int MAXF, bool FORWARD, int THREADS>
__global__ void __launch_bounds__(THREADS) epaCoopKernel(KParams P, int64_t n)
{
  extern __shared__ __align__(16) unsigned char epa_smem[];
  typedef EpaPool<MAXV, MAXF> Pool;
  constexpr int GROUPS = THREADS / G;
  const int grp = threadIdx.x / G;
  Pool& S = reinterpret_cast<Pool*>(epa_smem)[grp];
  V3<double>* Pw = reinterpret_cast<V3<double>*>(P.epa_scratch) + ((size_t)blockIdx.x * GROUPS + grp) * MAXV;
  Coop<G> g;
  const bool from_retry = !FORWARD && P.epa_retry != nullptr;
  const uint32_t count = from_retry ? P.qcount[Q_RETRY] : (uint32_t)n;
  for (;;)
  {
    uint32_t i = 0;
    if (g.lane == 0)
      i = atomicAdd(&P.qcount[FORWARD ? Q_FETCH : Q_FETCH_EPA], 1u);
    i = g.bcast(i);
    if (i >= count)
      break;
    epaTaskBody<G, MAXV, MAXF, FORWARD>(from_retry ? P.epa_retry[i] : i, P, S, Pw);
  }
}

typedef void (*EpaKernelFn)(KParams, int64_t);
struct EpaLaunchCfg
{
  EpaKernelFn fn;
  int threads, groups, maxv, blocks_per_sm;
  size_t smem;
};
static EpaLaunchCfg epaCfg(bool small)
{
  static EpaLaunchCfg cfg[2];
  static bool ready = false;
  if (!ready)
  {
    cfg[0].fn = epaCoopKernel<T4_EPA_SMALL_G, T4_EPA_SMALL_V, T4_EPA_SMALL_F, true, T4_EPA_SMALL_THREADS>;
    cfg[0].threads = T4_EPA_SMALL_THREADS, cfg[0].groups = T4_EPA_SMALL_THREADS / T4_EPA_SMALL_G;
    cfg[0].maxv = T4_EPA_SMALL_V;
    cfg[0].smem = cfg[0].groups * sizeof(EpaPool<T4_EPA_SMALL_V, T4_EPA_SMALL_F>);
    cfg[1].fn = epaCoopKernel<32, 128, 256, false, T4_EPA_BIG_THREADS>;
    cfg[1].threads = T4_EPA_BIG_THREADS, cfg[1].groups = T4_EPA_BIG_THREADS / 32, cfg[1].maxv = 128;
    cfg[1].smem = cfg[1].groups * sizeof(EpaPool<128, 256>);
    for (auto& c : cfg)
    {
      cudaFuncSetAttribute(c.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c.smem);
      int nb = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, c.fn, c.threads, c.smem) != cudaSuccess || nb < 1)
        nb = 1;
      c.blocks_per_sm = nb;
    }
    ready = true;
  }
  return cfg[small ? 0 : 1];
}

// bytes of witness scratch the two tiers need (tier 1 first, tier 2 behind it)
size_t t4EpaScratchBytes(int sm_count)
{
  const EpaLaunchCfg a = epaCfg(true), b = epaCfg(false);
  return ((size_t)sm_count * a.blocks_per_sm * a.groups * a.maxv + (size_t)sm_count * b.blocks_per_sm * b.groups * b.maxv) *
         sizeof(V3<double>);
}

// P.qcount[Q_FETCH], [Q_FETCH_EPA] and [Q_RETRY] must be zero; P.epa_retry holds n entries (null: tier 2 only);
// P.epa_scratch holds t4EpaScratchBytes().  Returns the number of launches.
int t4LaunchEpa(const KParams& P, int64_t n, int sm_count, cudaStream_t st)
{
  const EpaLaunchCfg a = epaCfg(true), b = epaCfg(false);
  KParams Q = P;
  int launches = 0;
  if (P.epa_retry)
  {
    const int64_t want = (n + a.groups - 1) / a.groups;
    const int grid = (int)t4max<int64_t>(1, t4min<int64_t>(want, (int64_t)sm_count * a.blocks_per_sm));
    a.fn<<<grid, a.threads, a.smem, st>>>(Q, n);
    ++launches;
  }
  Q.epa_scratch = P.epa_scratch + (size_t)sm_count * a.blocks_per_sm * a.groups * a.maxv * 3;
  // tier 2 does not know its task count on the host (it is the retry counter): a full persistent grid unless the
  // whole launch is smaller
  const int64_t want = (n + b.groups - 1) / b.groups;
  const int grid = (int)t4max<int64_t>(1, t4min<int64_t>(want, (int64_t)sm_count * b.blocks_per_sm));
  b.fn<<<grid, b.threads, b.smem, st>>>(Q, n);
  return launches + 1;
}

#endif
}  // namespace t4
#endif
