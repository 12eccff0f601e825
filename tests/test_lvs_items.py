"""t4c_lvs_items / t4c_lvs_item_states (host-only entries of the C ABI) against the oracle's LVS expansion
(oracle: tesseract_environment/core/utils.h:189-205, 328-347 restated).  CPU tier: no device work involved."""
import numpy as np
import pytest

from oracle import pyoracle
from tesseract_b200 import capi, scene as S
from tesseract_b200.managers import BatchChecker
from hostsim import build as hostsim


@pytest.fixture(scope="module")
def checker():
    b = S.build_config2(n_prims=50, margin=0.05)
    return BatchChecker(b, lib=hostsim.lib())


def trajectories(n_traj, n_wp, n_q, seed):
    rng = np.random.default_rng(seed)
    t = rng.uniform(-1.5, 1.5, size=(n_traj, n_wp, n_q))
    t[0, 1] = t[0, 0]                     # zero-length step: visited once, never subdivided
    t[1, 2] = t[1, 1] + 1e-4              # shorter than the longest valid segment
    t[2, 1:] = -t[2, :1]                  # |high| < |low| and sign flips: LinSpaced's flipped evaluation order
    return t


@pytest.mark.parametrize("cast", [False, True])
@pytest.mark.parametrize("lvs", [0.05, 0.4, np.inf])
def test_items_match_oracle(checker, cast, lvs):
    traj = trajectories(5, 4, 7, 7)
    q0_o, q1_o, it_o, st_o = pyoracle.lvs_expand(traj, lvs if np.isfinite(lvs) else 1e300, cast)
    it, st, sub = checker.lvs_items(traj, lvs, cast)
    assert len(it) == len(it_o)
    assert np.array_equal(it, it_o) and np.array_equal(st, st_o)
    # sub-step index restarts at 0 on every (trajectory, step)
    key = it.astype(np.int64) * 1000 + st
    first = np.r_[True, key[1:] != key[:-1]]
    assert np.all(sub[first] == 0) and np.all(np.diff(sub)[~first[1:]] == 1)
    if not np.isfinite(lvs):
        assert len(it) == 5 * (3 if cast else 4) and np.all(sub == 0)
    # joint vectors of a strided selection and of everything: bit-identical to the expansion
    for sel in (np.arange(len(it)), np.arange(0, len(it), 7), np.array([len(it) - 1])):
        q0, q1 = checker.lvs_item_states(traj, lvs, cast, sel)
        assert np.array_equal(q0, q0_o[sel])
        if cast:
            assert np.array_equal(q1, q1_o[sel])


def test_item_states_argument_errors(checker):
    traj = trajectories(3, 4, 7, 3)
    n = len(checker.lvs_items(traj, 0.1, False)[0])
    with pytest.raises(RuntimeError):
        checker.lvs_item_states(traj, 0.1, False, np.array([3, 3]))
    with pytest.raises(RuntimeError):
        checker.lvs_item_states(traj, 0.1, False, np.array([n]))
    with pytest.raises(RuntimeError):
        checker.lvs_items(traj, 0.0, False)
