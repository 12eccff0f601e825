"""Parity rules between the CUDA path and the CPU oracle (north_star tolerance).

  - key set per state: identical, except keys whose ORACLE distance lies within `band` (1e-5 m) of the pair margin
  - per contact (canonicalised so that link[0] < link[1]): |d distance| <= tol, nearest points <= tol, normals <= tol
    (normals are skipped when the cores touch, |d| < 1e-9); a contact whose witness pair is not unique (parallel
    features) may instead satisfy: same distance and normal, and pB - pA = d * n on both sides.
"""
import numpy as np

TOL = 1e-5


def canon(c):
    """(key, distance, pA, pB, normal, cc_time, cc_type) with link[0] < link[1]"""
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
        out[k] = (d, p, n, t, ty)
    return out


def compare(gpu, ora, margin_of, per_shape=False, tol=TOL, band=TOL, check_cc=False):
    """margin_of(l0, l1) -> pair margin.  Returns a dict of statistics; raises AssertionError on a violation."""
    G, O = index(gpu, per_shape), index(ora, per_shape)
    stats = dict(n_gpu=len(G), n_oracle=len(O), band_exempt=0, max_dd=0.0, max_dp=0.0, max_dn=0.0, nonunique=0)
    for k in set(G) ^ set(O):
        src = O if k in O else G
        d = src[k][0]
        m = margin_of(k[1], k[2])
        assert abs(d - m) <= band + 1e-7, f"key {k} only in {'oracle' if k in O else 'gpu'}: d={d}, margin={m}"
        stats["band_exempt"] += 1
    for k in set(G) & set(O):
        dg, pg, ng, tg, tyg = G[k]
        do, po, no, to, tyo = O[k]
        dd = abs(dg - do)
        assert dd <= tol + 2e-7 * max(1.0, abs(do)), f"{k}: distance {dg} vs {do}"
        stats["max_dd"] = max(stats["max_dd"], dd)
        if do <= -1e30 or abs(do) < 1e-30:
            continue  # calculate_penetration=false sentinel: sign only
        dp = max(np.linalg.norm(pg[0] - po[0]), np.linalg.norm(pg[1] - po[1]))
        dn = np.linalg.norm(ng - no) if abs(do) > 1e-9 else 0.0
        if dp > tol + 1e-6 or dn > tol + 1e-6:
            # non-unique witnesses / normals (parallel features, or equal-depth faces): both answers must be
            # self-consistent: pB - pA = d * n
            res_g = np.linalg.norm((pg[1] - pg[0]) - dg * ng)
            res_o = np.linalg.norm((po[1] - po[0]) - do * no)
            assert res_g <= 10 * tol and res_o <= 10 * tol, f"{k}: inconsistent contact {res_g} {res_o}"
            assert dn <= tol + 1e-6 or dd <= tol, f"{k}: normal {ng} vs {no}"
            stats["nonunique"] += 1
            stats["nonunique_example"] = (k, dp, dn)
        else:
            stats["max_dp"] = max(stats["max_dp"], dp)
            stats["max_dn"] = max(stats["max_dn"], dn)
        if check_cc:
            assert list(tyg) == list(tyo), f"{k}: cc_type {tyg} vs {tyo}"
            if dp <= tol + 1e-6:
                assert np.allclose(tg, to, atol=1e-4), f"{k}: cc_time {tg} vs {to}"
    return stats
