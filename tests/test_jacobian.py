"""Device Jacobians and contact gradients against finite differences of the CPU oracle.

The reference validates calcJacobian the same way: numericalJacobian (tesseract_kinematics/core/utils.h:88-125, central
idea: perturb one joint, difference the link pose) at 1e-3 tolerances (kinematics_test_utils.h runJacobianTest).  Here the
oracle FK plays the kinematics and the step is 1e-6."""
import ctypes as C

import numpy as np
import pytest

from oracle import pyoracle
from tesseract_b200 import capi, scene as S
from tesseract_b200.managers import BatchChecker


def numerical_jacobian(b, q, link, point_local, h=1e-6):
    n, nq = q.shape
    T0 = pyoracle.fk(b, q)[:, link]
    p0 = T0[:, :, :3] @ point_local + T0[:, :, 3]
    J = np.zeros((n, 6, nq))
    for j in range(nq):
        qp = q.copy()
        qp[:, j] += h
        T1 = pyoracle.fk(b, qp)[:, link]
        p1 = T1[:, :, :3] @ point_local + T1[:, :, 3]
        J[:, 0:3, j] = (p1 - p0) / h
        dR = T1[:, :, :3] @ np.transpose(T0[:, :, :3], (0, 2, 1))  # world-frame incremental rotation
        J[:, 3, j] = (dR[:, 2, 1] - dR[:, 1, 2]) / (2 * h)
        J[:, 4, j] = (dR[:, 0, 2] - dR[:, 2, 0]) / (2 * h)
        J[:, 5, j] = (dR[:, 1, 0] - dR[:, 0, 1]) / (2 * h)
    return J


def device_jacobian(bc, q, link, point_local):
    lib = bc.lib
    n, nq = q.shape
    J = np.zeros((n, 6, nq))
    dp = C.POINTER(C.c_double)
    pt = np.ascontiguousarray(point_local, dtype=np.float64)
    capi.check(lib.t4c_jacobian(bc.model, bc.scene.h, q.ctypes.data_as(dp), n, link, pt.ctypes.data_as(dp),
                                J.ctypes.data_as(dp)), lib)
    return J


def jacobian_case(mk, robot, n=200):
    b = S.SceneBuilder()
    S.add_robot(b, S.load_robot(robot))
    q = S.random_states(b, n, seed=17, dtype=np.float64)
    bc = mk(b)
    rng = np.random.default_rng(3)
    worst = 0.0
    for link in range(len(b.link_names)):
        pt = rng.uniform(-0.2, 0.2, size=3)
        Jd = device_jacobian(bc, q, link, pt)
        Jn = numerical_jacobian(b, q, link, pt)
        worst = max(worst, float(np.abs(Jd - Jn).max()))
    assert worst <= 5e-6, worst
    return worst


def gradient_case(mk, n=60, n_prims=3000, h=1e-6):
    b = S.build_config2(n_prims=n_prims, margin=0.05)
    q = S.random_states(b, n, seed=23, dtype=np.float64)
    bc = mk(b)
    hit, con, _ = bc.check_discrete(q, capi.TEST_CLOSEST)
    assert len(con) > 20
    con = np.ascontiguousarray(con)
    nq = q.shape[1]
    grad = np.zeros((len(con), nq))
    dp = C.POINTER(C.c_double)
    capi.check(bc.lib.t4c_contact_gradients(bc.scene.h, bc.model, q.ctypes.data_as(dp), n, con.ctypes.data, len(con),
                                            grad.ctypes.data_as(dp)), bc.lib)
    # finite differences of the ORACLE's closest distance of the same (state, link pair)
    def distances(qq):
        _, co, _ = pyoracle.check_discrete(b, qq, capi.TEST_CLOSEST, n_threads=4)
        return {(int(c["state"]), min(c["link"]), max(c["link"])): float(c["distance"]) for c in co}
    d0 = distances(q)
    keys = [(int(c["state"]), int(min(c["link"])), int(max(c["link"]))) for c in con]
    num = np.full((len(con), nq), np.nan)
    for j in range(nq):
        qp, qm = q.copy(), q.copy()
        qp[:, j] += h
        qm[:, j] -= h
        dp_, dm_ = distances(qp), distances(qm)
        for i, k in enumerate(keys):
            if k in dp_ and k in dm_:
                num[i, j] = (dp_[k] - dm_[k]) / (2 * h)
    ok = ~np.isnan(num).any(axis=1)
    err = np.abs(grad[ok] - num[ok]).max(axis=1)
    # the closest FEATURE pair can switch inside the step for a few contacts (kinks of the distance function): the
    # bulk must agree to 1e-4, and every contact to the size of a kink
    assert np.quantile(err, 0.9) <= 1e-4, np.quantile(err, 0.9)
    assert ok.sum() >= 0.8 * len(con)
    return dict(contacts=int(ok.sum()), median=float(np.median(err)), p90=float(np.quantile(err, 0.9)), max=float(err.max()))


def hostsim_checker(b):
    from hostsim import build as hostsim
    return BatchChecker(b, lib=hostsim.lib())


@pytest.mark.parametrize("robot", ["iiwa14", "abb_irb2400"])
def test_jacobian_hostsim(robot):
    print(jacobian_case(hostsim_checker, robot))


def test_contact_gradients_hostsim():
    print(gradient_case(hostsim_checker))


@pytest.mark.gpu
@pytest.mark.parametrize("robot", ["iiwa14", "abb_irb2400"])
def test_jacobian_device(robot):
    print(jacobian_case(lambda b: BatchChecker(b), robot, n=5000))


@pytest.mark.gpu
def test_contact_gradients_device():
    print(gradient_case(lambda b: BatchChecker(b), n=400, n_prims=10000))
