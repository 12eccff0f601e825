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


@pytest.mark.parametrize("tt", [capi.TEST_FIRST, capi.TEST_CLOSEST])
def test_discrete_config2_device_fk(tt):
    """full path: FK on the device (see parity_cases.flip_statistics)"""
    print(pc.discrete_config2(mk, 20000, 10000, tt, device_fk=True, threads=THREADS))


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
