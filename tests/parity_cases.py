"""Shared bodies of the randomized parity tests: CUDA path (real library on the GPU tier, host simulation of the same
kernel bodies on the CPU tier) vs the CPU oracle, on BASELINE.json config-2 style scenes.

Every comparison goes through tests/parity.py compare(): the north_star rule (colliding link-pair set exact except
within 1e-5 m of the margin; distance / nearest points / normals within 1e-5) with two counted, geometrically verified
exception classes (non-unique witnesses, unconverged curved penetration).  This holds for the FK-bypass entries AND for
the full device path: the device FK is bit-identical to the host's (t4_sincos.h), so no looser rule is needed there.
"""
import json
import os

import numpy as np

from tesseract_b200 import capi
from tesseract_b200 import scene as S
from oracle import pyoracle

import parity

REPORTS = []  # every comparison appends its report here; the GPU tier writes them to gpurun_out/parity_reports.json


def compare_batch(b, con_g, con_o, hit_g, hit_o, margin, per_shape=False, check_cc=False, link_poses=None, label=""):
    rep = parity.compare(b, con_g, con_o, hit_g, hit_o, lambda l0, l1: margin, per_shape=per_shape, check_cc=check_cc,
                         link_poses=link_poses)
    rep["label"] = label
    REPORTS.append(rep)
    out = os.environ.get("T4_PARITY_REPORT")
    if out:
        with open(out, "w") as f:
            json.dump(REPORTS, f, indent=1)
    return rep


def discrete_config2(make_checker, n_states, n_prims, test_type, seed=1234, threads=4, device_fk=False):
    b = S.build_config2(n_prims=n_prims, margin=0.05)
    q = S.random_states(b, n_states, seed=seed, dtype=np.float64)
    bc = make_checker(b)
    hit_o, con_o, _ = pyoracle.check_discrete(b, q, test_type, n_threads=threads)
    T = pyoracle.fk(b, q)
    if device_fk:
        hit_g, con_g, stats = bc.check_discrete(q, test_type)  # joint vectors in: FK on the device, the timed path
    else:
        hit_g, con_g, _, stats = bc.check_link_poses(T, None, test_type)  # FK bypass: the oracle's FP64 poses
    if test_type == capi.TEST_FIRST:
        # FIRST is order dependent in the reference (core/common.h:129-133): only emptiness is comparable
        assert np.array_equal(hit_g, hit_o), f"{int((hit_g != hit_o).sum())} hit flags differ"
        assert len(con_g) == int(hit_g.sum())
        return dict(n=len(con_g))
    return compare_batch(b, con_g, con_o, hit_g, hit_o, 0.05, per_shape=(test_type == capi.TEST_ALL), link_poses=T,
                         label=f"config2 {'device-FK' if device_fk else 'FK-bypass'} type {test_type} n {n_states}")


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
    else:
        hit_g, con_g, loc, stats = bc.check_link_poses(pyoracle.fk(b, q0), pyoracle.fk(b, q1), capi.TEST_CLOSEST)
    return compare_batch(b, con_g, con_o, hit_g, hit_o, 0.05, check_cc=True,
                         label=f"config3 cast {'device-FK + device LVS' if device_fk else 'FK-bypass'} {n_traj}x{n_wp}")


def discrete_lvs(make_checker, n_traj, n_wp, n_prims, lvs=0.05, threads=4):
    """checkTrajectory(LVS_DISCRETE) (core/utils.h:309-428): LVS sub-states through the trajectory entry point"""
    b = S.build_config2(n_prims=n_prims, margin=0.05)
    traj = S.random_trajectories(b, n_traj, n_wp, seed=11, sigma=0.03)
    q0, _, it, st = pyoracle.lvs_expand(traj, lvs, cast=False)
    bc = make_checker(b)
    hit_o, con_o, _ = pyoracle.check_discrete(b, q0, capi.TEST_CLOSEST, n_threads=threads)
    hit_g, con_g, loc, stats = bc.check_trajectories(traj, lvs, False, capi.TEST_CLOSEST)
    assert len(hit_g) == len(q0)
    return compare_batch(b, con_g, con_o, hit_g, hit_o, 0.05, link_poses=pyoracle.fk(b, q0),
                         label=f"discrete LVS device-FK {n_traj}x{n_wp}")


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
    out = compare_batch(b, con_g, con_o, hit_g, hit_o, 0.10, link_poses=T, label=f"config4 voxels n {n_states}")
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
    out = compare_batch(b, con_g, con_o, hit_g, hit_o, 0.05, link_poses=T, label=f"config5 dual arm n {n_states}")
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
    # bit-identical: same sine / cosine as the host libm (t4_sincos.h), same operation order, no FMA contraction
    bad = int((Tg.view(np.uint64) != To.view(np.uint64)).sum())
    assert bad == 0, f"{bad} of {Tg.size} pose elements differ (max {float(np.abs(Tg - To).max())})"
    return 0.0


def abb_mesh_config1(make_checker, n_states, threads=4, device_fk=True):
    """SURVEY 8(d) config 1 / BASELINE configs[0]: ABB IRB2400 with its links loaded AS GIVEN -- triangle Mesh shapes, one
    triangle shape per face (bullet_utils.cpp:111-132) -- self + environment check against one static 1 m box at
    (1.0, 0, 0.5), ContactTestType::ALL, margin 0.05; the all-zeros state plus uniform-in-limits states.  The self check
    of mesh links is compound x compound: every triangle of the smaller link walks the child tree of the other link
    (t4_kernels.h broadActiveBody)."""
    b = S.SceneBuilder()
    root, moving = S.add_robot(b, S.load_robot("abb_irb2400"), convex=False)
    b.add_collision_object("box", 0, [S.Box(1.0, 1.0, 1.0)], [S.identity()], pose_=S.pose((1.0, 0.0, 0.5)))
    b.set_active(moving)
    b.set_margin(0.05)
    q = np.concatenate([np.zeros((1, 6)), S.random_states(b, n_states - 1, seed=1, dtype=np.float64)])
    bc = make_checker(b)
    hit_o, con_o, _ = pyoracle.check_discrete(b, q, capi.TEST_ALL, n_threads=threads)
    T = pyoracle.fk(b, q)
    if device_fk:
        hit_g, con_g, stats = bc.check_discrete(q, capi.TEST_ALL)
    else:
        hit_g, con_g, _, stats = bc.check_link_poses(T, None, capi.TEST_ALL)
    rep = compare_batch(b, con_g, con_o, hit_g, hit_o, 0.05, per_shape=True, link_poses=T,
                        label=f"config1 ABB triangle-mesh links, self + box, ALL, n {n_states}")
    rep["self_contacts"] = int(sum(1 for c in con_g if b.link_active[c["link"][0]] and b.link_active[c["link"][1]]))
    rep["candidates"] = stats["candidates"]
    return rep
