"""The parity rule between the CUDA path and the CPU oracle (north_star tolerance) -- ONE strict
comparator used by every randomized parity test, CPU tier (host simulation of the kernel bodies) and GPU tier alike.

  1. key set per state identical, except keys whose distance lies within BAND (1e-5 m) of the pair margin; a state's
     hit flag may differ only when such a key is the reason;
  2. per common contact (canonicalised so that link[0] < link[1]): |d distance| <= 1e-5;
  3. normals within 1e-5 (skipped when the cores touch, |d| < 1e-9), nearest points within 1e-5;
  4. calculate_penetration = false sentinel records (distance ~ -DBL_MIN): sign only (Appendix B.6).

Two COUNTED exception classes replace rule 3; each needs a checked geometric reason, and the report carries their
counts and the measured deviations (no silent allowance):

  NONUNIQUE   the closest pair is not unique (parallel faces / edges).  Required: same distance and normal (1e-5),
              pB - pA = d n on both sides (1e-5), and -- when link poses are supplied -- both device witnesses lie ON
              their own shapes (point-in-shape test in pure numpy / scipy, independent of oracle and device).
  ORACLE_OFF_EXACT  the oracle -- the restated Bullet path -- is itself off the exact geometry: for barely penetrating
              polytopes (depth < 1e-4) the restated btGjkEpaSolver2 can leave its GJK on a repeated support vertex and
              report a small POSITIVE distance along an unrelated axis, and the answer then depends on the order of the
              pair (observed: hull x box -6.4e-5 exact, oracle +2.06e-4 as (hull, box) and -6.4e-5 as (box, hull)).
              Whether a real Bullet build does the same cannot be checked here (Bullet is not installed).  Accepted only
              when the DEVICE agrees to 1e-5 with the exact Minkowski-hull value computed by scipy (tests/
              second_opinion.py: no iteration, no tolerance); needs link poses and polytope / sphere shapes.
  CURVED_PEN  penetrating pair with a cylinder / cone core.  The reference's own penetration solver stops unconverged
              there (Bullet's 128-vertex polytope cap, tesseract_gjk_pair_detector.cpp:1024-1032 -> btGjkEpaSolver2):
              the depth bound agrees to 1e-5 but the face normal of a polytope that approximates a curved surface is
              only defined to ~sqrt(gap).  Required: distance rule 2 and self-consistency.  The normal / point
              deviations and the distance of the witnesses from their shapes are measured and returned (histogram).
"""
import numpy as np

TOL = 1e-5
BAND = 1e-5
F32 = 4e-7  # the device record stores distance / points / normal as FP32: half an ulp at 3 m

SPHERE, BOX, CYLINDER, CONE, CAPSULE, HULL, TRIANGLE = range(7)
CURVED = (CYLINDER, CONE)
HIST_EDGES = (1e-5, 1e-4, 1e-3, 1e-2, 3e-2, 1e-1, 1.0)


def canon(c):
    """(key, shape ids, distance, points, normal, cc_time, cc_type) with link[0] < link[1]"""
    l0, l1 = int(c["link"][0]), int(c["link"][1])
    p = np.asarray(c["point"], dtype=np.float64)
    n = np.asarray(c["normal"], dtype=np.float64)
    t = np.asarray(c["cc_time"], dtype=np.float64)
    ty = np.asarray(c["cc_type"]).astype(int)
    sid = (int(c["shape_id"][0]), int(c["subshape_id"][0]), int(c["shape_id"][1]), int(c["subshape_id"][1]))
    if l0 > l1:
        l0, l1 = l1, l0
        p = p[::-1]
        n = -n
        t = t[::-1]
        ty = ty[::-1]
        sid = (sid[2], sid[3], sid[0], sid[1])
    return (int(c["state"]), l0, l1), sid, float(c["distance"]), p, n, t, ty


def index(contacts, per_shape):
    out = {}
    for c in contacts:
        key, sid, d, p, n, t, ty = canon(c)
        k = key + sid if per_shape else key
        out[k] = (d, p, n, t, ty, sid)
    return out


# ---------------------------------------------------------------------------------------------------------------------
# independent point-on-shape test
# ---------------------------------------------------------------------------------------------------------------------
class ShapeTester:
    def __init__(self, builder):
        self.b = builder
        self.by_id = {(s["link"], s["shape_id"], s["subshape_id"]): s for s in builder.shapes}
        self.verts = (np.concatenate(builder.vertices) if builder.vertices else np.zeros((0, 3))).reshape(-1, 3)
        self._hull = {}

    def types(self, link, shape_id, sub):
        return self.by_id[(link, shape_id, sub)]["type"]

    def _hull_planes(self, off, cnt):
        key = (off, cnt)
        if key not in self._hull:
            from scipy.spatial import ConvexHull
            self._hull[key] = ConvexHull(self.verts[off:off + cnt]).equations
        return self._hull[key]

    def outside_by(self, link, shape_id, sub, link_pose, p_world):
        """signed slack of p against the shape: <= 0 inside or on the surface, else (an upper bound of) the distance"""
        s = self.by_id[(link, shape_id, sub)]
        L = np.asarray(s["pose"], dtype=np.float64).reshape(3, 4)
        T = np.asarray(link_pose, dtype=np.float64).reshape(3, 4)
        R = T[:, :3] @ L[:, :3]
        o = T[:, :3] @ L[:, 3] + T[:, 3]
        p = R.T @ (np.asarray(p_world, dtype=np.float64) - o)
        t, par = s["type"], s["params"]
        if t == SPHERE:
            return float(np.linalg.norm(p) - par[0])
        if t == BOX:
            return float(np.max(np.abs(p) - np.asarray(par[:3])))
        if t == CYLINDER:
            return float(max(np.hypot(p[0], p[1]) - par[0], abs(p[2]) - par[1]))
        if t == CONE:
            r, h = par[0], par[1]
            rad = r * (0.5 - p[2] / h)  # radius of the cone at height z (origin at mid height)
            side = (np.hypot(p[0], p[1]) - rad) * h / np.hypot(r, h)
            return float(max(side, abs(p[2]) - 0.5 * h))
        if t == CAPSULE:
            z = min(max(p[2], -par[1]), par[1])
            return float(np.linalg.norm(p - np.array([0, 0, z])) - par[0])
        if t == HULL:
            eq = self._hull_planes(s["vertex_offset"], s["vertex_count"])
            return float(np.max(eq[:, :3] @ p + eq[:, 3]))
        v = self.verts[s["vertex_offset"]:s["vertex_offset"] + 3]  # TRIANGLE: distance to the triangle
        return _point_triangle_distance(p, v[0], v[1], v[2])

    def world_core(self, link, shape_id, sub, link_pose):
        """(vertices of the core in the world frame, margin radius) for polytopes and spheres, else None"""
        s = self.by_id[(link, shape_id, sub)]
        t, par = s["type"], s["params"]
        if t == SPHERE:
            loc, r = np.zeros((1, 3)), par[0]
        elif t == BOX:
            h = np.asarray(par[:3])
            loc, r = np.array([[sx * h[0], sy * h[1], sz * h[2]] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)]), 0.0
        elif t == HULL:
            loc, r = self.verts[s["vertex_offset"]:s["vertex_offset"] + s["vertex_count"]], 0.0
        elif t == TRIANGLE:
            loc, r = self.verts[s["vertex_offset"]:s["vertex_offset"] + 3], 0.0
        else:
            return None
        L = np.asarray(s["pose"], dtype=np.float64).reshape(3, 4)
        T = np.asarray(link_pose, dtype=np.float64).reshape(3, 4)
        R = T[:, :3] @ L[:, :3]
        o = T[:, :3] @ L[:, 3] + T[:, 3]
        return loc @ R.T + o, r


def _point_triangle_distance(p, a, b, c):
    ab, ac, ap = b - a, c - a, p - a
    d1, d2 = ab @ ap, ac @ ap
    if d1 <= 0 and d2 <= 0:
        return float(np.linalg.norm(ap))
    bp = p - b
    d3, d4 = ab @ bp, ac @ bp
    if d3 >= 0 and d4 <= d3:
        return float(np.linalg.norm(bp))
    vc = d1 * d4 - d3 * d2
    if vc <= 0 and d1 >= 0 and d3 <= 0:
        return float(np.linalg.norm(ap - ab * (d1 / (d1 - d3))))
    cp = p - c
    d5, d6 = ab @ cp, ac @ cp
    if d6 >= 0 and d5 <= d6:
        return float(np.linalg.norm(cp))
    vb = d5 * d2 - d1 * d6
    if vb <= 0 and d2 >= 0 and d6 <= 0:
        return float(np.linalg.norm(ap - ac * (d2 / (d2 - d6))))
    va = d3 * d6 - d5 * d4
    if va <= 0 and (d4 - d3) >= 0 and (d5 - d6) >= 0:
        w = (d4 - d3) / ((d4 - d3) + (d5 - d6))
        return float(np.linalg.norm(p - (b + (c - b) * w)))
    n = np.cross(ab, ac)
    return float(abs(ap @ n) / np.linalg.norm(n))


# ---------------------------------------------------------------------------------------------------------------------
# the comparator
# ---------------------------------------------------------------------------------------------------------------------
def compare(builder, con_g, con_o, hit_g, hit_o, margin_of, per_shape=False, check_cc=False, link_poses=None,
            state_base=0):
    """con_g / con_o: device / oracle contact records; hit_g / hit_o: per-state flags (or None);
    margin_of(l0, l1) -> pair margin; link_poses: (n_states, n_links, 3, 4) FP64 poses for the witness tests (discrete
    queries only) or None.  Returns the report dict; raises AssertionError on any violation."""
    G, O = index(con_g, per_shape), index(con_o, per_shape)
    tester = ShapeTester(builder)
    rep = dict(n_gpu=len(G), n_oracle=len(O), band_exempt=0, max_dd=0.0, max_dp=0.0, max_dn=0.0, nonunique=0,
               curved_pen=0, curved_dn_hist=[0] * (len(HIST_EDGES) + 1), curved_max_dn=0.0, curved_max_dp=0.0,
               curved_max_witness_slack=0.0, max_dt=0.0, oracle_off_exact=0, oracle_off_exact_max=0.0,
               witness_checks=0, flag_mismatch=0, sentinel=0)
    exempt_states = set()
    for k in set(G) ^ set(O):
        src = O if k in O else G
        d = src[k][0]
        m = margin_of(k[1], k[2])
        assert abs(d - m) <= BAND + F32, f"key {k} only in {'oracle' if k in O else 'device'}: d={d}, margin={m}"
        rep["band_exempt"] += 1
        exempt_states.add(k[0])
    if hit_g is not None and hit_o is not None:
        diff = np.nonzero(np.asarray(hit_g) != np.asarray(hit_o))[0]
        rep["flag_mismatch"] = int(len(diff))
        for s in diff:
            assert int(s) + state_base in exempt_states, f"state {int(s) + state_base}: hit flag differs without a margin-band key"

    def on_shapes(k, sid, pts, enforce=True):
        """both witnesses lie on their own shapes (independent geometric test); returns the larger slack"""
        if link_poses is None:
            return 0.0
        st = k[0] - state_base
        worst = 0.0
        for side, link in ((0, k[1]), (1, k[2])):
            slack = tester.outside_by(link, sid[2 * side], sid[2 * side + 1], link_poses[st, link], pts[side])
            if enforce:
                assert slack <= TOL + F32, f"{k}: device witness {side} is {slack:.3g} m off its shape"
            worst = max(worst, slack)
        rep["witness_checks"] += 1
        return worst

    for k in set(G) & set(O):
        dg, pg, ng, tg, tyg, sidg = G[k]
        do, po, no, to, tyo, sido = O[k]
        if do > -1e-300 and do < 0 and abs(do) < 1e-30:  # calculate_penetration = false sentinel: sign only
            assert dg < 0 or abs(dg) < 1e-30, f"{k}: sentinel sign {dg}"
            rep["sentinel"] += 1
            continue
        dd = abs(dg - do)
        dn_pre = float(np.linalg.norm(ng - no)) if abs(do) > 1e-9 else 0.0

        def device_matches_exact():
            """ORACLE_OFF_EXACT (module comment): True when exact geometry confirms the DEVICE's distance and normal"""
            if link_poses is None:
                return False
            import second_opinion as so
            st = k[0] - state_base
            try:
                ca = tester.world_core(k[1], sidg[0], sidg[1], link_poses[st, k[1]])
                cb = tester.world_core(k[2], sidg[2], sidg[3], link_poses[st, k[2]])
            except KeyError:
                return False
            if ca is None or cb is None:
                return False
            d_exact, n_exact = so.exact_pair(ca[0], cb[0], ca[1], cb[1])
            if abs(dg - d_exact) > TOL + F32:
                return False
            # exact_pair's normal points from the second shape to the first; the record's from link 0 to link 1
            if n_exact is not None and abs(d_exact) > 1e-9 and np.linalg.norm(ng + n_exact) > 1e-4:
                return False
            rep["oracle_off_exact"] += 1
            rep["oracle_off_exact_max"] = max(rep["oracle_off_exact_max"], abs(do - d_exact))
            return True

        if (dd > TOL + F32 or dn_pre > 1e-3) and device_matches_exact():
            # how far the ORACLE is off is recorded, not bounded: it is not a device tolerance (worst seen: 9 mm for two
            # interpenetrating flat triangles of a mesh-mesh pair, where GJK / EPA on zero-volume shapes is ill-posed;
            # sign and normal of a 3 um triangle-triangle contact)
            continue
        assert dd <= TOL + F32, f"{k}: distance {dg} vs {do}"
        rep["max_dd"] = max(rep["max_dd"], dd)
        dp = max(np.linalg.norm(pg[0] - po[0]), np.linalg.norm(pg[1] - po[1]))
        dn = float(np.linalg.norm(ng - no)) if abs(do) > 1e-9 else 0.0
        res_g = float(np.linalg.norm((pg[1] - pg[0]) - dg * ng))
        curved = False
        if do < 0:
            try:
                ta, tb = tester.types(k[1], sidg[0], sidg[1]), tester.types(k[2], sidg[2], sidg[3])
                curved = ta in CURVED or tb in CURVED
            except KeyError:
                curved = False
        if curved and (dp > TOL + F32 or dn > TOL + F32):
            assert res_g <= TOL + 4 * F32, f"{k}: inconsistent device contact ({res_g})"
            # witnesses of an unconverged polytope face: reported, not enforced (the face need not contain the foot
            # point of the origin; Bullet's own blend has the same property)
            rep["curved_max_witness_slack"] = max(rep["curved_max_witness_slack"], on_shapes(k, sidg, pg, False))
            rep["curved_pen"] += 1
            rep["curved_max_dn"] = max(rep["curved_max_dn"], dn)
            rep["curved_max_dp"] = max(rep["curved_max_dp"], dp)
            rep["curved_dn_hist"][int(np.searchsorted(HIST_EDGES, dn))] += 1
            continue
        if dp > TOL + F32:
            res_o = float(np.linalg.norm((po[1] - po[0]) - do * no))
            assert dn <= TOL + F32, f"{k}: normal {ng} vs {no} (d={do})"
            assert res_g <= TOL + 4 * F32 and res_o <= TOL + 4 * F32, f"{k}: inconsistent contact {res_g} {res_o}"
            on_shapes(k, sidg, pg)
            rep["nonunique"] += 1
        else:
            assert dn <= TOL + F32, f"{k}: normal {ng} vs {no} (d={do})"
            rep["max_dp"] = max(rep["max_dp"], dp)
            rep["max_dn"] = max(rep["max_dn"], dn)
        if check_cc:
            assert list(tyg) == list(tyo), f"{k}: cc_type {tyg} vs {tyo}"
            if dp <= TOL + F32:
                # cc_time is a RATIO of two lengths of the order of the sweep (bullet_utils.h:506-523): a witness that
                # moves by dp changes it by ~dp / sweep, so metres at 1e-5 are fractions at 1e-4 for centimetre sweeps
                dt = float(np.max(np.abs(tg - to)))
                assert dt <= 1e-4, f"{k}: cc_time {tg} vs {to}"
                rep["max_dt"] = max(rep["max_dt"], dt)
    return rep
