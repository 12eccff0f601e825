"""Shared bodies of the randomized parity tests: CUDA path (real library on the GPU tier, host simulation of the same
kernel bodies on the CPU tier) vs the CPU oracle, on BASELINE.json config-2 style scenes.

Tolerances (north_star): colliding link-pair set exact except within 1e-5 m of the margin; distance / nearest points /
normals within 1e-5.  One documented exception class: PENETRATING pairs with a curved core (cylinder, cone).  There the
reference's own EPA does not converge (Bullet stops at its 128-vertex cap, tesseract_gjk_pair_detector.cpp:1024-1032 ->
btGjkEpaSolver2); both sides report the tightest depth bound, which agrees to 1e-5, but the face normal / witness of a
polytope approximating a curved surface is only defined to ~sqrt(gap): checked at 3e-2 plus self-consistency (depth itself agrees to 1e-5).
"""
import numpy as np

from tesseract_b200 import capi
from tesseract_b200 import scene as S
from oracle import pyoracle

import parity

CURVED = (capi.CYLINDER, capi.CONE)


def link_types(b):
    t = {}
    for s in b.shapes:
        t.setdefault(s["link"], set()).add(s["type"])
    return t


def compare_batch(b, con_g, con_o, hit_g, hit_o, margin, per_shape=False, check_cc=False):
    assert np.array_equal(hit_g, hit_o), f"{int((hit_g != hit_o).sum())} hit flags differ"
    types = link_types(b)
    G, O = parity.index(con_g, per_shape), parity.index(con_o, per_shape)
    for k in set(G) ^ set(O):
        d = (O.get(k) or G.get(k))[0]
        assert abs(d - margin) <= 1e-5 + 1e-7, f"key {k} on one side only, d={d}"
    stats = dict(n=len(G), max_dd=0.0, max_dp=0.0, max_dn=0.0, curved_pen=0, curved_max_dn=0.0, nonunique=0)
    for k in set(G) & set(O):
        dg, pg, ng, tg, tyg = G[k]
        do, po, no, to, tyo = O[k]
        dd = abs(dg - do)
        assert dd <= 1e-5, f"{k}: distance {dg} vs {do}"
        stats["max_dd"] = max(stats["max_dd"], dd)
        dp = max(np.linalg.norm(pg[0] - po[0]), np.linalg.norm(pg[1] - po[1]))
        dn = float(np.linalg.norm(ng - no))
        curved = bool((types.get(k[1], set()) | types.get(k[2], set())) & set(CURVED))
        if curved and do < 0:
            stats["curved_pen"] += 1
            stats["curved_max_dn"] = max(stats["curved_max_dn"], dn)
            assert dn <= 3e-2, f"{k}: curved penetration normal {ng} vs {no}"
            assert np.linalg.norm((pg[1] - pg[0]) - dg * ng) <= 1e-5, f"{k}: inconsistent contact"
            continue
        if dp > 1e-5 or dn > 1e-5:
            # non-unique witnesses (parallel features): same distance and normal, self-consistent points
            assert dn <= 1e-5, f"{k}: normal {ng} vs {no} (d={do})"
            assert np.linalg.norm((pg[1] - pg[0]) - dg * ng) <= 2e-5, f"{k}: inconsistent contact"
            stats["nonunique"] += 1
            continue
        stats["max_dp"] = max(stats["max_dp"], dp)
        stats["max_dn"] = max(stats["max_dn"], dn)
        if check_cc:
            assert list(tyg) == list(tyo), f"{k}: cc_type {tyg} vs {tyo}"
            assert np.allclose(tg, to, atol=1e-5), f"{k}: cc_time {tg} vs {to}"
    return stats


def flip_statistics(b, con_g, con_o, hit_g, hit_o, margin, per_shape=False):
    """Full device path (FK on the GPU).  The reference detector is DISCONTINUOUS in its inputs: its GJK stops on
    branch tests between quantities that differ by rounding noise (t4_bullet.h), so a 1-ulp change of a link pose
    (device sincos vs glibc) can move a contact to the other side of such a branch.  Both answers are then valid GJK
    bounds of the same pair.  Everything else must agree to 1e-5; the flipped fraction is bounded and reported."""
    G, O = parity.index(con_g, per_shape), parity.index(con_o, per_shape)
    one_sided = [k for k in set(G) ^ set(O) if abs((O.get(k) or G.get(k))[0] - margin) > 1e-4]
    assert not one_sided, one_sided[:5]
    both = sorted(set(G) & set(O))
    dd = np.array([abs(G[k][0] - O[k][0]) for k in both])
    dn = np.array([np.linalg.norm(G[k][2] - O[k][2]) for k in both])
    flipped = int(((dd > 1e-5) | (dn > 1e-4)).sum())
    assert dd.max() <= 2e-4, dd.max()
    assert flipped <= max(3, 2e-3 * len(both)), (flipped, len(both))
    assert int((hit_g != hit_o).sum()) <= max(1, 1e-4 * len(hit_o))
    return dict(n=len(both), flipped=flipped, frac=flipped / max(1, len(both)), max_dd=float(dd.max()))


def discrete_config2(make_checker, n_states, n_prims, test_type, seed=1234, threads=4, device_fk=False):
    b = S.build_config2(n_prims=n_prims, margin=0.05)
    q = S.random_states(b, n_states, seed=seed, dtype=np.float64)
    bc = make_checker(b)
    hit_o, con_o, _ = pyoracle.check_discrete(b, q, test_type, n_threads=threads)
    if device_fk:
        hit_g, con_g, stats = bc.check_discrete(q, test_type)
        if test_type == capi.TEST_FIRST:
            assert int((hit_g != hit_o).sum()) <= max(1, 1e-4 * len(hit_o))
            return dict(n=len(con_g))
        return flip_statistics(b, con_g, con_o, hit_g, hit_o, 0.05, per_shape=(test_type == capi.TEST_ALL))
    # identical FP64 link poses on both sides (the oracle's FK): the exact pass must then track the oracle
    hit_g, con_g, _, stats = bc.check_link_poses(pyoracle.fk(b, q), None, test_type)
    if test_type == capi.TEST_FIRST:
        # FIRST is order dependent in the reference (core/common.h:129-133): only emptiness is comparable
        assert np.array_equal(hit_g, hit_o)
        assert len(con_g) == int(hit_g.sum())
        return dict(n=len(con_g))
    return compare_batch(b, con_g, con_o, hit_g, hit_o, 0.05, per_shape=(test_type == capi.TEST_ALL))


def flags_only(make_checker, n_states, n_prims, seed=77, threads=4, margin=0.05, pen=True, dist=True):
    b = S.build_config2(n_prims=n_prims, margin=margin)
    q = S.random_states(b, n_states, seed=seed, dtype=np.float64)
    bc = make_checker(b)
    hit_g, _, _ = bc.check_discrete(q, capi.TEST_FIRST, calculate_penetration=pen, calculate_distance=dist,
                                    want_contacts=False)
    hit_o, _, _ = pyoracle.check_discrete(b, q, capi.TEST_FIRST, calculate_penetration=pen, calculate_distance=dist)
    assert np.array_equal(hit_g, hit_o), f"{int((hit_g != hit_o).sum())} flags differ"
    return int(hit_g.sum())


def cast_config3(make_checker, n_traj, n_wp, n_prims, lvs=0.05, threads=4, device_fk=False):
    b = S.build_config2(n_prims=n_prims, margin=0.05)
    traj = S.random_trajectories(b, n_traj, n_wp, seed=7)
    q0, q1, it, st = pyoracle.lvs_expand(traj, lvs, cast=True)
    bc = make_checker(b)
    hit_o, con_o, _ = pyoracle.check_cast(b, q0, q1, capi.TEST_CLOSEST, n_threads=threads)
    if device_fk:
        # LVS expansion + FK + cast check through the trajectory entry point
        hit_g, con_g, loc, stats = bc.check_trajectories(traj, lvs, True, capi.TEST_CLOSEST)
        assert len(hit_g) == len(q0)
        return flip_statistics(b, con_g, con_o, hit_g, hit_o, 0.05)
    hit_g, con_g, loc, stats = bc.check_link_poses(pyoracle.fk(b, q0), pyoracle.fk(b, q1), capi.TEST_CLOSEST)
    return compare_batch(b, con_g, con_o, hit_g, hit_o, 0.05, check_cc=True)


def discrete_lvs(make_checker, n_traj, n_wp, n_prims, lvs=0.05, threads=4):
    """checkTrajectory(LVS_DISCRETE) (core/utils.h:309-428): LVS sub-states through the trajectory entry point"""
    b = S.build_config2(n_prims=n_prims, margin=0.05)
    traj = S.random_trajectories(b, n_traj, n_wp, seed=11, sigma=0.03)
    q0, _, it, st = pyoracle.lvs_expand(traj, lvs, cast=False)
    bc = make_checker(b)
    hit_o, con_o, _ = pyoracle.check_discrete(b, q0, capi.TEST_CLOSEST, n_threads=threads)
    hit_g, con_g, loc, stats = bc.check_trajectories(traj, lvs, False, capi.TEST_CLOSEST)
    assert len(hit_g) == len(q0)
    return flip_statistics(b, con_g, con_o, hit_g, hit_o, 0.05)


def _lbracket(seed):
    """small concave fixture: an L-bracket as a closed triangle mesh (two boxes sharing a face)"""
    rng = np.random.default_rng(seed)
    a, t, w = rng.uniform(0.15, 0.3), rng.uniform(0.02, 0.04), rng.uniform(0.1, 0.2)
    boxes = [((0, 0, 0), (a, w, t)), ((0, 0, 0), (t, w, a))]
    V, F = [], []
    for (x0, y0, z0), (dx, dy, dz) in boxes:
        o = len(V)
        V += [(x0 + i * dx, y0 + j * dy, z0 + k * dz) for i in (0, 1) for j in (0, 1) for k in (0, 1)]
        quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
        for q in quads:
            F += [(o + q[0], o + q[1], o + q[2]), (o + q[0], o + q[2], o + q[3])]
    return np.array(V, dtype=np.float64), np.array(F, dtype=np.int64)


def config4_voxels(make_checker, leaves, plane, n_states, threads=4, test_type=capi.TEST_CLOSEST):
    """configs[3]: ABB hull links vs voxels + concave mesh, distance queries at margin 0.10 (FK bypass: exact poses)"""
    b = S.build_config4(leaves, plane, margin=0.10)
    q = S.random_states(b, n_states, seed=99, dtype=np.float64)
    bc = make_checker(b)
    T = pyoracle.fk(b, q)
    hit_g, con_g, _, stats = bc.check_link_poses(T, None, test_type)
    hit_o, con_o, _ = pyoracle.check_discrete(b, q, test_type, n_threads=threads)
    out = compare_batch(b, con_g, con_o, hit_g, hit_o, 0.10)
    out["colliding"] = int(hit_g.sum())
    return out


def config5_dual_arm(make_checker, n_states, n_prims, threads=4, test_type=capi.TEST_CLOSEST):
    """configs[4]: two iiwa with cross-arm pairs + concave fixtures + primitives (FK bypass: exact poses)"""
    fixtures = []
    for i in range(4):
        v, f = _lbracket(5 + i)
        fixtures.append((f"fixture_{i}", v, f, S.pose((0.45 * (1 if i % 2 else -1), 0.3 * (1 if i < 2 else -1), 0.35))))
    b = S.build_config5(n_prims=n_prims, margin=0.05, fixtures=fixtures)
    q = S.random_states(b, n_states, seed=2024, dtype=np.float64)
    bc = make_checker(b)
    T = pyoracle.fk(b, q)
    hit_g, con_g, _, stats = bc.check_link_poses(T, None, test_type)
    hit_o, con_o, _ = pyoracle.check_discrete(b, q, test_type, n_threads=threads)
    out = compare_batch(b, con_g, con_o, hit_g, hit_o, 0.05)
    out["colliding"] = int(hit_g.sum())
    out["self_or_cross"] = int(sum(1 for c in con_g if c["link"][0] < 16 and c["link"][1] < 16))
    return out


def fk_parity(make_checker, robot="iiwa14", n=200):
    b = S.SceneBuilder()
    S.add_robot(b, S.load_robot(robot))
    q = S.random_states(b, n, seed=5, dtype=np.float64)
    bc = make_checker(b)
    Tg = bc.fk(q)
    To = pyoracle.fk(b, q)
    err = float(np.abs(Tg - To).max())
    assert err <= 1e-9, err
    return err
