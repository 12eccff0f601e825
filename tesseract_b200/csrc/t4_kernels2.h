/*
 * t4_kernels2.h -- narrowphase kernel bodies: FP32 cull, FP64 exact query + result bridging, CLOSEST reduction.
 * See t4_kernels.h for the pipeline overview and reference citations.
 */
#ifndef T4_KERNELS2_H_
#define T4_KERNELS2_H_

#include "t4_kernels.h"

namespace t4
{
template <class T>
T4_KD void loadPairShapes(const KParams& P, int64_t s, int a, int b, ShapeW<T>& A, ShapeW<T>& B)
{
  const DShape& Sa = P.shapes[a];
  const DShape& Sb = P.shapes[b];
  Xf<T> a0, a1, b0, b1;
  shapePose<T>(P, Sa, a, s, a0, a1);
  shapePose<T>(P, Sb, b, s, b0, b1);
  makeShapeW<T>(A, Sa, P.verts, P.sup_cells, P.sup_cand, a0, a1, P.cast && Sa.alink >= 0);
  makeShapeW<T>(B, Sb, P.verts, P.sup_cells, P.sup_cand, b0, b1, P.cast && Sb.alink >= 0);
}

T4_KD void loadPairShapesD(const KParams& P, int64_t s, int a, int b, ShapeW<double>& A, ShapeW<double>& B)
{
  const DShape& Sa = P.shapes[a];
  const DShape& Sb = P.shapes[b];
  Xf<double> a0, a1, b0, b1;
  shapePoseD(P, Sa, a, s, a0, a1);
  shapePoseD(P, Sb, b, s, b0, b1);
  makeShapeW<double>(A, Sa, P.verts, P.sup_cells, P.sup_cand, a0, a1, P.cast && Sa.alink >= 0);
  makeShapeW<double>(B, Sb, P.verts, P.sup_cells, P.sup_cand, b0, b1, P.cast && Sb.alink >= 0);
  // FP64 shape parameters (the FP32 record is only good for culling)
  const double* pa = P.shape_d + (size_t)a * 16;
  const double* pb = P.shape_d + (size_t)b * 16;
  if (A.type == SH_BOX)
    A.he = V3<double>(pa[0], pa[1], pa[2]);
  else if (A.type != SH_HULL && A.type != SH_TRIANGLE)
    A.r = pa[0], A.h = pa[1];
  if (B.type == SH_BOX)
    B.he = V3<double>(pb[0], pb[1], pb[2]);
  else if (B.type != SH_HULL && B.type != SH_TRIANGLE)
    B.r = pb[0], B.h = pb[1];
}

T4_KD bool analyticPair(int ta, int tb)
{
  if (ta == SH_SPHERE && (tb == SH_SPHERE || tb == SH_TRIANGLE || tb == SH_CAPSULE))
    return true;
  if (ta == SH_TRIANGLE && tb == SH_SPHERE)
    return true;
  if (ta == SH_CAPSULE && (tb == SH_CAPSULE || tb == SH_SPHERE))
    return true;
  return false;
}

T4_KD uint32_t orderedBits(float f);

T4_KD void pushSurvivor(const KParams& P, uint32_t c, int tier)
{
  uint32_t* q = tier == 0 ? P.surv_w : (tier == 1 ? P.surv2_w : (tier == 2 ? P.surv3_w : P.surv4_w));
  const uint32_t k = qAppend(&P.qcount[Q_SURV + tier]);
  if (k < P.surv_cap)
    q[k] = c;
}

// The reference's GJK may stop early with an upper bound up to ~1e-4 m above the true distance (t4_bullet.h), so a pair
// is accepted on the FP32 bound alone only when it is this far inside the margin; the rest goes to the exact pass.
#define T4_ACCEPT_SLACK 5e-4f

// item = candidate index.  FP32 GJK used ONLY to discard pairs that are provably farther apart than the contact
// cut-off, and to recognise pairs that are certainly inside it:
//   - flags-only output: a certain contact sets the state's flag, nothing else is queued for it;
//   - FIRST with records: ONE certain, non-penetrating pair per state goes to the primary queue (its record is the
//     state's contact), ONE certainly penetrating pair per state to the second queue, every other survivor to the
//     third; the later queues are only looked at for states that are still without a contact afterwards
//     (ContactTestData::done semantics, core/common.h:129-133).
// Everything that is reported goes through the exact FP64 pass.
T4_KD void cullBody(int64_t item, const KParams& P)
{
  const int64_t c = item + (int64_t)P.cull_base;
  const int64_t s = P.cand_state[c];
  const bool first_rec = P.test_type == T4C_TEST_FIRST && !P.hit_only;
  if (P.hit_only && P.hit[s])
    return;
  const int a = (int)P.cand_a[c], b = (int)P.cand_b[c];
  ShapeW<float> A, B;
  loadPairShapes<float>(P, s, a, b, A, B);
  if (!A.cast && !B.cast && analyticPair(A.type, B.type))
  {
    if (P.cand_class)
      P.cand_class[c] = 0;
    if (P.key_min_ub)
      P.cand_lb[c] = -1e30f;  // analytic pairs carry no bounds: never pruned
    pushSurvivor(P, (uint32_t)c, first_rec ? 3 : 0);
    return;
  }
  const V3<float> axis0 = A.x0.p - B.x0.p;
  const V3<float> offset = (A.x0.p + B.x0.p) * 0.5f;
  A.x0.p -= offset;
  B.x0.p -= offset;
  if (A.cast)
    A.x1.p -= offset;
  if (B.cast)
    B.x1.p -= offset;
  const float msum = A.margin() + B.margin();
  const float cutoff = msum + P.max_margin + T4_CULL_SLACK;
  float accept = cutoff;  // anything within the cut-off is decided by the exact pass
  bool bool_mode = false;
  if ((P.hit_only || first_rec) && P.calc_dist)
  {
    const float pm = t4min((float)pairMargin(P, P.shapes[a].alink, P.shapes[b].link), P.max_margin);
    accept = msum + pm - T4_ACCEPT_SLACK;
    bool_mode = accept > 0.f;
    if (!bool_mode)
      accept = cutoff;
  }
  if (!P.calc_dist)
    accept = -1.f;  // separated cores are never reported (tesseract_gjk_pair_detector.cpp:830-834)
  GjkOut<float, false> g;
  // FIRST with records: a pair that is certainly within range iterates on (a few support pairs at most) until it is
  // also known whether its cores are apart -- the difference between one cheap GJK and an EPA run in the exact pass
  const int st = gjkDistance<float, false>(A, B, axis0, false, cutoff, accept, g, 48, (first_rec && bool_mode) ? 8 : 0);
  statAdd(P.stats, ST_GJK_PAIRS, 1);
  statAdd(P.stats, ST_CULL_ITERS, (uint32_t)g.iters);
  statAdd(P.stats, ST_SUPV, A.evals + B.evals);
  if (st == GJK_FAR)
    return;
  // low 6 bits: the exact pass of this pair will need about as many iterations (sort key: keeps a warp's lanes in
  // step); bit 7: the cores are certainly separated (see gjkBodyInner)
  if (P.cand_class)
    P.cand_class[c] = (uint8_t)((g.iters < 63 ? g.iters : 63) | (g.separated_by_1mm ? 0x80 : 0));
  if (st == GJK_SEPARATED && !P.calc_dist)
    return;
  if (P.key_min_ub)
  {
    // distance bounds of this pair from the FP32 iteration: the closest point of the last simplex is a point of A - B
    // (upper bound of the core distance), the best support plane a lower bound; intersecting cores: both zero
    const bool sep = st == GJK_SEPARATED || st == GJK_NEAR;
    const float ub = (sep ? t4sqrt(g.v.len2()) : 0.f) - msum;
    const float lb = (sep ? t4sqrt(g.lower2) : 0.f) - msum;
    P.cand_lb[c] = lb;
    const int grp = P.link_group[P.shapes[b].link];
    if (grp >= 0)
    {
      const size_t slot = ((size_t)s * P.n_alinks + P.shapes[a].alink) * P.n_groups + grp;
#ifdef T4_DEVICE_BUILD
      atomicMin(&P.key_min_ub[slot], orderedBits(ub));
#else
      if (orderedBits(ub) < P.key_min_ub[slot])
        P.key_min_ub[slot] = orderedBits(ub);
#endif
    }
  }
  if (P.hit_only)
  {
    // simplex encloses the origin (cores overlap), or |v| <= margins + pair margin - slack: certainly a contact
    if ((st == GJK_INTERSECT && (g.v.len2() == 0.f || bool_mode)) || (st == GJK_NEAR && bool_mode))
    {
      P.hit[s] = 1;
      return;
    }
  }
  if (first_rec)
  {
    // tier 0: one pair per state that is certainly in contact AND whose cores are certainly apart (phase 2 of the
    // detector only: no EPA, phase 1 skipped under the same certificate); tiers 1 / 2: one certainly penetrating
    // pair per state (needs EPA; only looked at when the earlier tiers left the state without a contact) -- against a
    // flat-faced or point-core shape (median 7 EPA iterations) before one against a cylinder / cone (15-75 iterations:
    // the polytope has to approximate the curved surface); tier 3: the rest.
    const bool certain = st == GJK_NEAR && bool_mode && g.separated_by_1mm;
    const bool deep = st == GJK_INTERSECT && g.v.len2() == 0.f && P.calc_pen;
    const bool curved = A.curvedCore() || B.curvedCore();
    int tier = 3;
    if (certain && (orU32(&P.cull_claim[s], 1u) & 1u) == 0u)
      tier = 0;
    else if (deep && !curved && (orU32(&P.cull_claim[s], 2u) & 2u) == 0u)
      tier = 1;
    else if (deep && curved && (orU32(&P.cull_claim[s], 4u) & 4u) == 0u)
      tier = 2;
    pushSurvivor(P, (uint32_t)c, tier);
    return;
  }
  pushSurvivor(P, (uint32_t)c, 0);
}

// ContactTestType::CLOSEST keeps ONE record per (state, link pair): the minimum distance (core/common.h:151-157).  When
// link B is a compound of many shapes (an octomap of a million voxels, a mesh), hundreds of its shapes are inside the
// margin of a robot link, and all but one of their exact results are thrown away.  item = survivor of the cull: it stays
// in the exact queue only if its distance LOWER bound does not exceed the smallest UPPER bound seen for its key by more
// than the slack -- what the exact pass (which may stop up to ~1e-4 above the true distance, t4_bullet.h) and the FP32
// arithmetic of the bounds can differ by.  The reported minimum is unchanged: the pair the reference would report has
// lower bound <= its reference distance <= every other pair's reference distance <= that pair's upper bound + 1e-4.
#define T4_PRUNE_SLACK 5e-4f
T4_KD float orderedFloat(uint32_t u);
T4_KD void pruneBody(int64_t i, const KParams& P)
{
  const uint32_t c = P.surv[i];
  const int b = (int)P.cand_b[c];
  const int grp = P.link_group[P.shapes[b].link];
  bool keep = true;
  if (grp >= 0)
  {
    const int64_t s = P.cand_state[c];
    const int a = (int)P.cand_a[c];
    const size_t slot = ((size_t)s * P.n_alinks + P.shapes[a].alink) * P.n_groups + grp;
    const uint32_t m = P.key_min_ub[slot];  // 0xffffffff: no pair of this key has published a bound (analytic pairs)
    keep = m == 0xffffffffu || P.cand_lb[c] <= orderedFloat(m) + T4_PRUNE_SLACK;
  }
  if (keep)
  {
    const uint32_t k = qAppend(&P.qcount[Q_SURV2]);
    P.surv2_w[k] = c;  // at most as many as came in
  }
}

T4_KD uint32_t orderedBits(float f)
{
  uint32_t u;
#ifdef T4_DEVICE_BUILD
  u = __float_as_uint(f);
#else
  union
  {
    float f;
    uint32_t u;
  } cv;
  cv.f = f;
  u = cv.u;
#endif
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
T4_KD float orderedFloat(uint32_t u)
{
  const uint32_t v = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef T4_DEVICE_BUILD
  return __uint_as_float(v);
#else
  union
  {
    float f;
    uint32_t u;
  } cv;
  cv.u = v;
  return cv.f;
#endif
}
T4_KD unsigned long long orderedBits64(double d)
{
  unsigned long long u;
#ifdef T4_DEVICE_BUILD
  u = (unsigned long long)__double_as_longlong(d);
#else
  union
  {
    double d;
    unsigned long long u;
  } cv;
  cv.d = d;
  u = cv.u;
#endif
  return (u >> 63) ? ~u : (u | 0x8000000000000000ULL);
}

// TesseractBroadphaseBridgedManifoldResult::addContactPoint -> add{Discrete,Cast}SingleResult -> processResult
// (bullet_utils.h:388-441,527-611,757-802 ; core/common.h:114-167)
T4_KDN void emitContact(const KParams& P, int64_t s, int a, int b, const ShapeW<double>& A, const ShapeW<double>& B,
                       const PairResult<double>& pr)
{
  if (!pr.has)
    return;
  if (pr.distance > P.max_margin_d)  // reject beyond the max margin (:717, :759)
    return;
  const DShape& Sa = P.shapes[a];
  const DShape& Sb = P.shapes[b];
  // per-pair margin unless both calculate_* flags are off (core/common.h:122-124)
  if ((P.calc_dist || P.calc_pen) && pr.distance > pairMargin(P, Sa.alink, Sb.link))
    return;
  if (P.hit)
    P.hit[s] = 1;
  if (P.hit_only || !P.raw)
    return;
  if (P.test_type == T4C_TEST_FIRST && casU32(&P.first_claim[s], 0u, 1u) != 0u)
    return;
  const uint32_t slot = qAppend(&P.qcount[Q_RAW]);
  if (slot >= P.raw_cap)
    return;
  const V3<double> pointA = pr.pointOnB + pr.normalOnB * pr.distance;
  t4c_contact ct;
  ct.state = (int32_t)(s + P.state_base);
  ct.flags = pr.used_epa ? 1u : 0u;
  ct.distance = (float)pr.distance;
  ct.cc_time[0] = ct.cc_time[1] = -1.f;
  ct.cc_type[0] = ct.cc_type[1] = T4C_CC_NONE;
  ct.reserved[0] = ct.reserved[1] = 0;
  double local[2][3] = { { 0, 0, 0 }, { 0, 0, 0 } };
  bool swap = false;
  if (P.cast)
  {
    // addCastSingleResult (bullet_utils.h:527-611): continuous data per cast side; a single cast side is reported as
    // index 1 (:595-607)
    const size_t offA = poseOffset(P, s, Sa.alink);
    const Xf<double> linkA0 = Xf<double>::from12(P.poses0d + offA);
    ShapeW<double> Au = A;
    Au.cast = false;
    double t;
    int ty;
    V3<double> pl = linkA0.invApply(pointA);
    if (Sb.alink >= 0)
    {
      continuousData(Au, linkA0, pointA, -pr.normalOnB, t, ty, pl);
      ct.cc_time[0] = (float)t, ct.cc_type[0] = (uint8_t)ty;
      local[0][0] = pl.x, local[0][1] = pl.y, local[0][2] = pl.z;
      const size_t offB = poseOffset(P, s, Sb.alink);
      const Xf<double> linkB0 = Xf<double>::from12(P.poses0d + offB);
      ShapeW<double> Bu = B;
      Bu.cast = false;
      V3<double> plb = linkB0.invApply(pr.pointOnB);
      continuousData(Bu, linkB0, pr.pointOnB, pr.normalOnB, t, ty, plb);
      ct.cc_time[1] = (float)t, ct.cc_type[1] = (uint8_t)ty;
      local[1][0] = plb.x, local[1][1] = plb.y, local[1][2] = plb.z;
    }
    else
    {
      swap = true;
      continuousData(Au, linkA0, pointA, -pr.normalOnB, t, ty, pl);
      ct.cc_time[1] = (float)t, ct.cc_type[1] = (uint8_t)ty;
      local[1][0] = pl.x, local[1][1] = pl.y, local[1][2] = pl.z;
    }
  }
  const int i0 = swap ? 1 : 0, i1 = swap ? 0 : 1;
  const double sg = swap ? 1.0 : -1.0;
  ct.link[i0] = Sa.link, ct.link[i1] = Sb.link;
  ct.shape_id[i0] = (int16_t)Sa.shape_id, ct.shape_id[i1] = (int16_t)Sb.shape_id;
  ct.subshape_id[i0] = Sa.subshape_id, ct.subshape_id[i1] = Sb.subshape_id;
  ct.point[i0][0] = (float)pointA.x, ct.point[i0][1] = (float)pointA.y, ct.point[i0][2] = (float)pointA.z;
  ct.point[i1][0] = (float)pr.pointOnB.x, ct.point[i1][1] = (float)pr.pointOnB.y, ct.point[i1][2] = (float)pr.pointOnB.z;
  ct.normal[0] = (float)(sg * pr.normalOnB.x), ct.normal[1] = (float)(sg * pr.normalOnB.y);
  ct.normal[2] = (float)(sg * pr.normalOnB.z);
  P.raw[slot] = ct;
  if (P.raw_d64)
    P.raw_d64[slot] = pr.distance;
  if (P.raw_local)
  {
    float* L = P.raw_local + (size_t)slot * 6;
    for (int j = 0; j < 2; ++j)
      for (int q = 0; q < 3; ++q)
        L[3 * j + q] = (float)local[j][q];
  }
}

// work counters of one exact-pass item, flushed once per item (every exit of the body)
struct ExactCounters
{
  uint32_t pairs = 0, p1 = 0, p2 = 0, supv = 0, skipped = 0, violations = 0;
};

// item = survivor index: phases 1-2 of the exact (FP64) query.  Pairs that need the penetration solver are handed to
// the EPA kernel through the task queue, so that a warp never mixes the short GJK path with the long EPA path.
T4_KD void gjkBodyInner(int64_t k, const KParams& P, ExactCounters& cnt)
{
  const uint32_t c = P.surv[k];
  const int64_t s = P.cand_state[c];
  if ((P.test_type == T4C_TEST_FIRST || P.hit_only) && P.hit[s])
    return;
  const int a = (int)P.cand_a[c], b = (int)P.cand_b[c];
  ShapeW<double> A, B;
  loadPairShapesD(P, s, a, b, A, B);
  PairResult<double> pr;
  cnt.pairs = 1;
  if (analyticPairQuery(A, B, P.max_margin_d, pr))
  {
    emitContact(P, s, a, b, A, B, pr);
    return;
  }
  // recentre in place (the world-pose copies are only needed for the continuous data of cast contacts: reloaded there)
  const double maxDist2 = detectorMaxDist2(A, B, P.max_margin_d);
  const V3<double> offset = bt::recentre(A, B);
  bt::DetectorState<double> S;
  bt::DetectorOut<double> d;
  const bool budgeted = P.gjk_budget > 0 && P.long_w;
  // Phase 1 (the boolean GJK) has ONE output: "the cores intersect" or not.  When the FP32 cull found a support plane
  // that puts the cores more than 1 mm apart -- a thousand times its own rounding error and far beyond every tolerance
  // of the boolean GJK (its "touching" exits need the origin within ~1e-8 of a simplex of A - B, or, for a segment,
  // within 1.5e-8 / |AB| with |AB| >= the separation) -- that output is known and the 3-4 support pairs are not spent.
  // Only with calculate_distance on: without it the reference tells "separated" (-1) from "undecided after 1000
  // iterations" (-2).  P.verify_skip (test hook) runs phase 1 anyway and counts the pairs whose outcome contradicts the
  // certificate.
  const bool certified = P.calc_dist && P.cand_class && (P.cand_class[c] & 0x80);
  const int r = bt::detectPhase12(A, B, offset, maxDist2, P.calc_pen != 0, P.calc_dist != 0, S, d,
                                  budgeted ? P.gjk_budget : (1 << 30), certified && !P.verify_skip);
  if (certified)
  {
    cnt.skipped = 1;
    if (P.verify_skip && S.status == 0 && r != bt::DET_BUDGET)
      cnt.violations = 1;
  }
  if (r == bt::DET_PHASE4)
    bt::detectPhase4(A, B, offset, maxDist2, S, d);
  cnt.p1 = (uint32_t)d.p1_iters;
  cnt.p2 = (uint32_t)(d.gjk_iters - d.p1_iters);
  cnt.supv = A.evals + B.evals;
  if (r == bt::DET_BUDGET)
  {
    // a pair that cycles towards the detector's 1000-iteration caps would hold its whole launch for milliseconds
    // with one active lane: such pairs are collected and run together at the end of the chunk
    const uint32_t t = qAppend(&P.qcount[Q_LONG]);
    if (t < P.long_cap)
      P.long_w[t] = c;
    return;
  }
  if (r == bt::DET_PHASE3)
  {
    const uint32_t t = qAppend(&P.qcount[P.task_q]);
    if (t < P.task_cap)
    {
      EpaTask& T = P.tasks[t];
      T.cand = c;
      T.is_valid = S.isValid;
      T.distance = S.distance;
      T.normalInB[0] = S.normalInB.x, T.normalInB[1] = S.normalInB.y, T.normalInB[2] = S.normalInB.z;
      T.pointOnA[0] = S.pointOnA.x, T.pointOnA[1] = S.pointOnA.y, T.pointOnA[2] = S.pointOnA.z;
      T.pointOnB[0] = S.pointOnB.x, T.pointOnB[1] = S.pointOnB.y, T.pointOnB[2] = S.pointOnB.z;
      T.orgNormalInB[0] = S.orgNormalInB.x, T.orgNormalInB[1] = S.orgNormalInB.y, T.orgNormalInB[2] = S.orgNormalInB.z;
    }
    return;
  }
  copyOut(d, pr);
  if (!pr.has)
    return;
  if (P.cast)
    loadPairShapesD(P, s, a, b, A, B);
  emitContact(P, s, a, b, A, B, pr);
}

T4_KD void gjkBody(int64_t k, const KParams& P)
{
  ExactCounters cnt;
  gjkBodyInner(k, P, cnt);
  statAdd(P.stats, ST_EXACT_PAIRS, cnt.pairs);
  statAdd(P.stats, ST_P1_ITERS, cnt.p1);
  statAdd(P.stats, ST_P2_ITERS, cnt.p2);
  statAdd(P.stats, ST_SUPV64, cnt.supv);
  statAdd(P.stats, ST_P1_SKIPPED, cnt.skipped);
  if (cnt.violations)
    statAdd(P.stats, ST_P1_VIOLATIONS, cnt.violations);
}

// item = EPA task: phase 3 (penetration) and phase 4 of the exact query
template <class EPA_T>
T4_KD void epaBody(int64_t t, const KParams& P, EPA_T& epa)
{
  const EpaTask& T = P.tasks[P.task_order ? P.task_order[t] : (uint32_t)t];
  const uint32_t c = T.cand;
  const int64_t s = P.cand_state[c];
  if ((P.test_type == T4C_TEST_FIRST || P.hit_only) && P.hit[s])
    return;
  const int a = (int)P.cand_a[c], b = (int)P.cand_b[c];
  ShapeW<double> A, B;
  loadPairShapesD(P, s, a, b, A, B);
  const double maxDist2 = detectorMaxDist2(A, B, P.max_margin_d);
  const V3<double> offset = bt::recentre(A, B);
  bt::DetectorState<double> S;
  S.isValid = T.is_valid;
  S.status = 0;
  S.distance = T.distance;
  S.normalInB = V3<double>(T.normalInB[0], T.normalInB[1], T.normalInB[2]);
  S.pointOnA = V3<double>(T.pointOnA[0], T.pointOnA[1], T.pointOnA[2]);
  S.pointOnB = V3<double>(T.pointOnB[0], T.pointOnB[1], T.pointOnB[2]);
  S.orgNormalInB = V3<double>(T.orgNormalInB[0], T.orgNormalInB[1], T.orgNormalInB[2]);
  bt::DetectorOut<double> d;
  d.has = false;
  d.gjk_iters = 0;
  d.epa_iters = 0;
  d.p1_iters = 0;
  d.used_epa = false;
  bt::detectPhase3(A, B, S, d, epa);
  bt::detectPhase4(A, B, offset, maxDist2, S, d);
  statAdd(P.stats, ST_EPA_GJK_ITERS, (uint32_t)d.gjk_iters);
  statAdd(P.stats, ST_EPA_PAIRS, 1u);
  statAdd(P.stats, ST_EPA_ITERS, (uint32_t)d.epa_iters);
  statAdd(P.stats, ST_SUPV64, A.evals + B.evals);
  statAdd(P.stats, ST_SUPV_EPA, A.evals + B.evals);
  PairResult<double> pr;
  copyOut(d, pr);
  if (!pr.has)
    return;
  if (P.cast)
    loadPairShapesD(P, s, a, b, A, B);
  emitContact(P, s, a, b, A, B, pr);
}

T4_KD unsigned long long mix64(unsigned long long x)
{
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}

T4_KD bool sameKey(const t4c_contact& x, const t4c_contact& y)
{
  if (x.state != y.state)
    return false;
  return (x.link[0] == y.link[0] && x.link[1] == y.link[1]) || (x.link[0] == y.link[1] && x.link[1] == y.link[0]);
}

// ContactTestType::CLOSEST (core/common.h:151-157): keep the minimum-distance contact per (state, link pair).
// item = raw contact index.  Pass 1 finds / claims the hash slot of the key and atomically minimises the ordered bits
// of the FP64 distance (the reference compares doubles: two triangles of one mesh often tie in FP32); pass 2 lets the
// first record that still finds its own distance in the slot take it (CAS to a sentinel) and copy itself out.
T4_KD uint32_t closestSlot(const KParams& P, uint32_t i, bool insert)
{
  const t4c_contact& c = P.raw[i];
  const unsigned lo = (unsigned)(c.link[0] < c.link[1] ? c.link[0] : c.link[1]);
  const unsigned hi = (unsigned)(c.link[0] < c.link[1] ? c.link[1] : c.link[0]);
  uint32_t h = (uint32_t)mix64(((unsigned long long)(unsigned)c.state << 32) ^ ((unsigned long long)lo << 21) ^ hi) &
               (P.ht_size - 1);
  for (uint32_t probe = 0; probe < P.ht_size; ++probe)
  {
    int32_t rep = P.ht_rep[h];
    if (rep < 0 && insert)
    {
      rep = casI32(&P.ht_rep[h], -1, (int32_t)i);
      if (rep < 0)
        return h;
    }
    if (rep >= 0 && sameKey(P.raw[rep], c))
      return h;
    if (rep < 0 && !insert)
      return 0xffffffffu;
    h = (h + 1) & (P.ht_size - 1);
  }
  return 0xffffffffu;
}

T4_KD void closestInsertBody(int64_t i, const KParams& P)
{
  const uint32_t h = closestSlot(P, (uint32_t)i, true);
  if (h == 0xffffffffu)
    return;
  minU64(&P.ht_best[h], orderedBits64(P.raw_d64[i]));
}

T4_KD void closestEmitBody(int64_t i, const KParams& P)
{
  const uint32_t h = closestSlot(P, (uint32_t)i, false);
  if (h == 0xffffffffu)
    return;
  const unsigned long long key = orderedBits64(P.raw_d64[i]);
  if (P.ht_best[h] != key || casU64(&P.ht_best[h], key, 0xfffffffffffffffeULL) != key)
    return;
  const uint32_t slot = qAppend(&P.qcount[Q_OUT]);
  if (slot >= P.out_cap)
    return;
  P.out[slot] = P.raw[i];
  if (P.out_local && P.raw_local)
    for (int q = 0; q < 6; ++q)
      P.out_local[(size_t)slot * 6 + q] = P.raw_local[(size_t)i * 6 + q];
}

}  // namespace t4
#endif
