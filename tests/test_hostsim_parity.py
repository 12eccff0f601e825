"""CPU tier: the CUDA kernel bodies (host simulation, tests/hostsim) vs the CPU oracle on randomized robot scenes."""
import pytest

from tesseract_b200 import capi
from tesseract_b200.managers import BatchChecker

import parity_cases as pc
from hostsim import build as hostsim


def mk(b):
    return BatchChecker(b, lib=hostsim.lib())


@pytest.mark.parametrize("tt", [capi.TEST_FIRST, capi.TEST_CLOSEST, capi.TEST_ALL])
def test_discrete_config2(tt):
    print(pc.discrete_config2(mk, 400, 10000, tt))


@pytest.mark.parametrize("tt", [capi.TEST_FIRST, capi.TEST_CLOSEST])
def test_discrete_config2_device_fk(tt):
    """full path: FK on the device (see parity_cases.flip_statistics)"""
    print(pc.discrete_config2(mk, 300, 10000, tt, device_fk=True))


def test_cast_trajectories_device_fk():
    print(pc.cast_config3(mk, 4, 10, 10000, device_fk=True))


def test_flags_only():
    assert pc.flags_only(mk, 600, 10000) > 0


def test_flags_only_boolean_request():
    """calculate_penetration = calculate_distance = false at margin 0 (the OMPL-validator use, SURVEY 8(d) mode A)"""
    assert pc.flags_only(mk, 600, 10000, margin=0.0, pen=False, dist=False) > 0


def test_discrete_lvs():
    print(pc.discrete_lvs(mk, 4, 10, 10000))


def test_cast_lvs():
    print(pc.cast_config3(mk, 6, 12, 10000))


@pytest.mark.parametrize("robot", ["iiwa14", "abb_irb2400"])
def test_fk(robot):
    pc.fk_parity(mk, robot)


def test_config1_abb_mesh_links():
    """configs[0]: ABB with triangle-mesh links, self + environment, ALL (compound x compound through the child trees)"""
    out = pc.abb_mesh_config1(mk, 12)
    print(out)
    assert out["n_gpu"] > 0


def test_first_in_rounds():
    """FIRST queries in rounds over the robot's shapes (t4c.cu runBatch; on by default for >= 16384 states): same
    flags, one record per colliding state, each record a contact the oracle also reports"""
    import subprocess, sys, os
    code = ("import sys; sys.path.insert(0, 'tests'); import parity_cases as pc; from hostsim import build as h; "
            "from tesseract_b200.managers import BatchChecker; from tesseract_b200 import capi; "
            "mk = lambda b: BatchChecker(b, lib=h.lib()); "
            "print(pc.discrete_config2(mk, 400, 10000, capi.TEST_FIRST)); "
            "print(pc.flags_only(mk, 600, 10000)); "
            "print(pc.flags_only(mk, 600, 10000, margin=0.0, pen=False, dist=False))")
    env = dict(os.environ, T4C_ROUND_MIN_STATES="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout + r.stderr


@pytest.mark.parametrize("hooks", [dict(T4C_LIVE_ORDER="0"), dict(T4C_ROUND_MERGE="0"), dict(T4C_ROUND_MERGE="300")])
def test_first_in_rounds_ab_switches(hooks):
    """the A/B switches of the round scheduler (ordered live-state lists off, slice growth off / extreme): same flags
    as the oracle, one record per colliding state"""
    import subprocess, sys, os
    code = ("import sys; sys.path.insert(0, 'tests'); import parity_cases as pc; from hostsim import build as h; "
            "from tesseract_b200.managers import BatchChecker; from tesseract_b200 import capi; "
            "mk = lambda b: BatchChecker(b, lib=h.lib()); "
            "print(pc.discrete_config2(mk, 300, 10000, capi.TEST_FIRST)); "
            "print(pc.flags_only(mk, 400, 10000))")
    env = dict(os.environ, T4C_ROUND_MIN_STATES="1", **hooks)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout + r.stderr


def test_config4_voxels_and_mesh():
    """configs[3] style: ABB hull links vs octomap voxels (fixture box_2m.bt) + the plane_4m.stl triangle mesh"""
    import os
    import numpy as np
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    leaves = np.load(os.path.join(gold, "box_2m_octree.npz"))["leaves"].astype(np.float64)
    leaves[:, :3] += (0.0, -1.0, 0.6)
    pl = np.load(os.path.join(gold, "plane_4m.npz"))
    out = pc.config4_voxels(mk, leaves, (pl["vertices"], pl["faces"]), 60)
    print(out)
    assert out["colliding"] > 0


def test_config5_dual_arm():
    """configs[4] style: two iiwa (cross-arm pairs) + concave L-bracket fixtures + primitives"""
    out = pc.config5_dual_arm(mk, 600, 6000, threads=8)
    print(out)
    assert out["self_or_cross"] > 0


def test_chunked_batches_with_rounds():
    """batches larger than one chunk (default 1 Mi states; forced to 128 here) with FIRST rounds on: state indices,
    contact offsets and per-chunk queues stay consistent; T4C_DRAIN_MIN=1 makes the host-buffer entries download the
    finished records after every round and chunk (the progressive copy of hostRun) and T4C_UPLOAD_PIECE_MIN=16 the
    piecewise q upload / FK; T4C_GJK_BUDGET=4 sends most pairs through the deferred long-pair pass (results must not
    change: a deferred pair is run again from the start)"""
    import subprocess, sys, os
    code = ("import sys; sys.path.insert(0, 'tests'); import parity_cases as pc; from hostsim import build as h; "
            "from tesseract_b200.managers import BatchChecker; from tesseract_b200 import capi; "
            "mk = lambda b: BatchChecker(b, lib=h.lib()); "
            "print(pc.discrete_config2(mk, 500, 10000, capi.TEST_FIRST)); "
            "print(pc.discrete_config2(mk, 500, 10000, capi.TEST_CLOSEST)); "
            "print(pc.flags_only(mk, 700, 10000)); "
            "print(pc.cast_config3(mk, 5, 12, 10000, device_fk=True))")
    env = dict(os.environ, T4C_ROUND_MIN_STATES="1", T4C_CHUNK_STATES="128", T4C_DRAIN_MIN="1",
               T4C_UPLOAD_PIECE_MIN="16", T4C_GJK_BUDGET="4")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout + r.stderr


def test_phase1_certificate_never_contradicted():
    """T4C_VERIFY_SKIP: phase 1 runs for certified pairs too and contradictions are counted (see test_gpu_fullsize)"""
    import subprocess, sys, os
    code = ("import sys; sys.path.insert(0, 'tests'); import numpy as np; from hostsim import build as h; "
            "from tesseract_b200 import capi, scene as S; from tesseract_b200.managers import BatchChecker; "
            "b = S.build_config2(n_prims=10000, margin=0.05); bc = BatchChecker(b, lib=h.lib()); "
            "q = S.random_states(b, 1500, seed=5, dtype=np.float64); "
            "hh, c, st = bc.check_discrete(q, capi.TEST_CLOSEST); print(st['phase1_skipped'], st['phase1_skip_violations']); "
            "assert st['phase1_skipped'] > 1000 and st['phase1_skip_violations'] == 0, st")
    env = dict(os.environ, T4C_VERIFY_SKIP="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout + r.stderr


def test_round_scheduler_edge_cases():
    """FIRST in rounds on the corner inputs: a scene without static shapes (no environment tree: self-collision only),
    batches of 0 / 1 / 2 states, cast segments, flags without distance and penetration"""
    import subprocess, sys, os
    code = r'''
import sys; sys.path.insert(0, "tests")
import numpy as np
from hostsim import build as h
from tesseract_b200.managers import BatchChecker
from tesseract_b200 import capi, scene as S
from oracle import pyoracle
lib = h.lib()
b = S.SceneBuilder(); S.add_robot(b, S.load_robot("abb_irb2400")); b.set_margin(0.05)
q = S.random_states(b, 300, seed=3, dtype=np.float64)
hit, con, st = BatchChecker(b, lib=lib).check_discrete(q, capi.TEST_FIRST)
assert np.array_equal(hit, pyoracle.check_discrete(b, q, capi.TEST_FIRST)[0]) and len(con) == int(hit.sum())
b2 = S.build_config2(n_prims=500, margin=0.05)
bc2 = BatchChecker(b2, lib=lib)
for n in (1, 2, 0):
    q2 = S.random_states(b2, max(n, 1), seed=5, dtype=np.float64)[:n]
    hit, con, st = bc2.check_discrete(q2, capi.TEST_FIRST)
    want = pyoracle.check_discrete(b2, q2, capi.TEST_FIRST)[0] if n else np.zeros(0, np.uint8)
    assert np.array_equal(hit, want) and len(con) == int(hit.sum())
q0 = S.random_states(b2, 200, seed=6, dtype=np.float64)
q1 = q0 + 0.05 * np.random.default_rng(1).standard_normal(q0.shape)
hit, con, loc, st = bc2.check_cast(q0, q1, capi.TEST_FIRST)
assert np.array_equal(hit, pyoracle.check_cast(b2, q0, q1, capi.TEST_FIRST)[0]) and len(con) == int(hit.sum())
hit, con, st = bc2.check_discrete(q0, capi.TEST_FIRST, calculate_penetration=False, calculate_distance=False,
                                  want_contacts=False)
assert np.array_equal(hit, pyoracle.check_discrete(b2, q0, capi.TEST_FIRST, calculate_penetration=False,
                                                   calculate_distance=False)[0])
print("edge cases ok")
'''
    env = dict(os.environ, T4C_ROUND_MIN_STATES="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "edge cases ok" in r.stdout, r.stdout + r.stderr
