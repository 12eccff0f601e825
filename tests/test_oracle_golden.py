"""Pins the CPU oracle against every known-answer test the reference holds for the hot path
(SURVEY.md section 4.1 / 8(c)).  CPU only."""
import numpy as np
import pytest

from oracle import pyoracle
from tesseract_b200 import capi
from tesseract_b200 import scene as S

import scenarios as sc


def make(builder, cast):
    return pyoracle.OracleManager(builder, cast)


@pytest.mark.parametrize("convex", [False, True])
@pytest.mark.parametrize("tt", [capi.TEST_FIRST, capi.TEST_CLOSEST, capi.TEST_ALL])
def test_box_box(convex, tt):
    sc.box_box(make, convex, tt)


def test_sphere_sphere_primitive():
    sc.sphere_sphere_primitive(make)


def test_sphere_sphere_convex():
    sc.sphere_sphere_convex(make)


@pytest.mark.parametrize("convex", [False, True])
def test_box_sphere(convex):
    sc.box_sphere(make, convex)


@pytest.mark.parametrize("kind", ["cylinder", "cone", "capsule"])
def test_box_cyl_cone_capsule(kind):
    sc.box_cylinder_cone_capsule(make, kind)


def test_mesh_mesh():
    sc.mesh_mesh(make)


@pytest.mark.parametrize("convex,tol", [(False, 0.001), (True, 0.02)])
def test_octomap_sphere(convex, tol):
    sc.octomap_sphere(make, convex, tol)


def test_octomap_mesh_2712():
    sc.octomap_mesh(make)


@pytest.mark.parametrize("cast", [False, True])
def test_octomap_octomap(cast):
    sc.octomap_octomap(make, cast)


@pytest.mark.parametrize("cast", [False, True])
def test_compound_compound(cast):
    sc.compound_compound(make, cast)


@pytest.mark.parametrize("convex", [False, True])
def test_large_dataset_300(convex):
    sc.large_dataset(make, convex)


def test_multi_clone_2700():
    sc.multi_clone(make)


def test_box_box_cast():
    sc.box_box_cast(make)


@pytest.mark.parametrize("convex", [False, True])
def test_sphere_sphere_cast(convex):
    sc.sphere_sphere_cast(make, convex)


def test_fk_iiwa_zero_pose():
    """tesseract_kinematics/test/kinematics_test_utils.h:441-540: iiwa zero-pose link frames."""
    b = S.SceneBuilder()
    S.add_robot(b, S.load_robot("iiwa14"))
    T = pyoracle.fk(b, np.zeros((1, 7)))[0]
    li = b.link_index
    np.testing.assert_allclose(T[li["link_2"]][:, 3], [-0.00043624, 0, 0.36], atol=1e-9)
    assert abs(T[li["link_4"]][2, 3] - 0.78) < 1e-9
    assert abs(T[li["link_6"]][2, 3] - 1.18) < 1e-9
    assert abs(T[li["tool0"]][2, 3] - 1.306) < 1e-9
    for l in range(T.shape[0]):
        np.testing.assert_allclose(T[l][:, :3], np.eye(3), atol=1e-12)


def test_fk_matches_independent_chain_product():
    """KDL-vs-OFKT style cross-check (tesseract_environment_unit.cpp:153-195): the oracle's tree FK vs an
    independent numpy chain product T_origin * Rot(axis, q), 1e-9."""
    robot = S.load_robot("abb_irb2400")
    b = S.SceneBuilder()
    S.add_robot(b, robot)
    rng = np.random.default_rng(0)
    q = S.random_states(b, 20, seed=3, dtype=np.float64)
    T = pyoracle.fk(b, q)
    for n in range(len(q)):
        cur = np.eye(4)
        k = 0
        for j in robot["joints"]:
            if j["parent"] == "base_link" and j["child"] == "base":
                continue
            O = np.eye(4)
            O[:3, :] = np.array(j["origin"])
            R = np.eye(4)
            if j["type"] == "revolute":
                a = np.array(j["axis"]) / np.linalg.norm(j["axis"])
                K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
                R[:3, :3] = np.eye(3) + np.sin(q[n, k]) * K + (1 - np.cos(q[n, k])) * K @ K
                k += 1
            cur = cur @ O @ R
            np.testing.assert_allclose(T[n, b.link_index[j["child"]]], cur[:3, :], atol=1e-9)
    del rng


def test_lvs_expand_rule():
    """core/utils.h:189-205,328-347: cnt = ceil(dist/lvs)+1, LinSpaced; discrete checks sub-states 0..cnt-2 and
    the last waypoint itself."""
    traj = np.zeros((1, 3, 2))
    traj[0, 1] = [0.3, 0.4]   # dist 0.5
    traj[0, 2] = [0.3, 0.42]  # dist 0.02 (< lvs)
    q0, q1, it, st = pyoracle.lvs_expand(traj, 0.1, cast=True)
    assert len(q0) == 5 + 1  # ceil(0.5/0.1)+1 = 6 points -> 5 sub-segments, then one plain segment
    np.testing.assert_allclose(q0[1], [0.06, 0.08], atol=1e-15)
    np.testing.assert_allclose(q1[4], [0.3, 0.4], atol=0)
    assert list(st) == [0, 0, 0, 0, 0, 1]
    q0, _, it, st = pyoracle.lvs_expand(traj, 0.1, cast=False)
    assert len(q0) == 5 + 1 + 1
    np.testing.assert_allclose(q0[-1], [0.3, 0.42], atol=0)
