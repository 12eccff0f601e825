"""GPU tier: libt4c.so on a B200 vs the CPU oracle on randomized robot scenes (BASELINE.json configs[1], [2])."""
import os

import pytest

from tesseract_b200 import capi
from tesseract_b200.managers import BatchChecker

import parity_cases as pc

pytestmark = pytest.mark.gpu
THREADS = os.cpu_count() or 4


def mk(b):
    return BatchChecker(b)


@pytest.mark.parametrize("tt", [capi.TEST_FIRST, capi.TEST_CLOSEST, capi.TEST_ALL])
def test_discrete_config2(tt):
    print(pc.discrete_config2(mk, 20000, 10000, tt, threads=THREADS))


@pytest.mark.parametrize("tt", [capi.TEST_FIRST, capi.TEST_CLOSEST, capi.TEST_ALL])
def test_discrete_config2_device_fk(tt):
    """the timed path -- joint vectors in, FK on the device -- under the same strict rule as the FK-bypass entries"""
    print(pc.discrete_config2(mk, 20000, 10000, tt, device_fk=True, threads=THREADS))


def test_sincos_device():
    """t4_sincos.h evaluated by a kernel on the device: the host libm's bits on 2e6 arguments"""
    import test_sincos as ts
    print(ts.check_bits(ts.run_lib(capi.load_library()), 2000000, table_range_only=True), "arguments, 0 mismatches")


def test_budgeted_gjk_async_pass():
    """FIRST rounds + a GJK budget of 4 support pairs on 32 768 states: most pairs are deferred to the long-pair pass,
    which runs on its own stream beside the deferred tiers and the EPA launch; results must not change (a deferred
    pair is run again from the start)"""
    import subprocess, sys
    code = ("import sys, os; sys.path.insert(0, 'tests'); import parity_cases as pc; "
            "from tesseract_b200.managers import BatchChecker; from tesseract_b200 import capi; "
            "mk = lambda b: BatchChecker(b); t = os.cpu_count(); "
            "print(pc.discrete_config2(mk, 32768, 10000, capi.TEST_FIRST, device_fk=True, threads=t)); "
            "print(pc.discrete_config2(mk, 32768, 10000, capi.TEST_CLOSEST, device_fk=True, threads=t)); "
            "print(pc.flags_only(mk, 32768, 10000, threads=t))")
    env = dict(os.environ, T4C_GJK_BUDGET="4", T4C_DRAIN_MIN="1", T4C_UPLOAD_PIECE_MIN="4096")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=1800)
    print(r.stdout[-2000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


def test_cast_trajectories_device_fk():
    print(pc.cast_config3(mk, 64, 30, 10000, device_fk=True, threads=THREADS))


def test_flags_only():
    assert pc.flags_only(mk, 30000, 10000, threads=THREADS) > 0


def test_flags_only_boolean_request():
    assert pc.flags_only(mk, 30000, 10000, margin=0.0, pen=False, dist=False, threads=THREADS) > 0


def test_discrete_lvs():
    print(pc.discrete_lvs(mk, 64, 30, 10000, threads=THREADS))


def test_cast_lvs():
    print(pc.cast_config3(mk, 64, 30, 10000, threads=THREADS))


@pytest.mark.parametrize("robot", ["iiwa14", "abb_irb2400"])
def test_fk(robot):
    pc.fk_parity(mk, robot, n=5000)


def test_config4_monkey_octomap():
    """configs[3]: ABB hull links vs the 414 194 occupied leaves of blender_monkey.bt + plane_4m.stl, margin 0.10"""
    import numpy as np
    gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    leaves = np.load(os.path.join(gold, "blender_monkey_octree.npz"))["leaves"].astype(np.float64)
    pl = np.load(os.path.join(gold, "plane_4m.npz"))
    b_kw = dict(robot_xyz=(0.0, -1.6, 0.0))
    from tesseract_b200 import scene as S
    orig = S.build_config4
    try:
        S.build_config4 = lambda l, p, margin=0.10: orig(l, p, margin=margin, **b_kw)
        out = pc.config4_voxels(mk, leaves, (pl["vertices"], pl["faces"]), 256, threads=THREADS)
    finally:
        S.build_config4 = orig
    print(out)
    assert out["colliding"] > 0


def test_config5_dual_arm():
    out = pc.config5_dual_arm(mk, 4000, 10000, threads=THREADS)
    print(out)
    assert out["self_or_cross"] > 0


def test_config1_abb_mesh_links():
    """configs[0]: ABB with triangle-mesh links as given, self + environment, ALL, zero state + 1000 random states"""
    out = pc.abb_mesh_config1(mk, 1001, threads=THREADS)
    print(out)
    assert out["n_gpu"] > 0
