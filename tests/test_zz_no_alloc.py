"""t4c_fk_noalloc: the allocation-free variant of t4c_fk (fixed device scratch array inside the library image).

No counterpart in the reference.  Collected last on purpose.
The GPU test calls the entry through the C ABI twice in a row -- the second call finds the first
call's bytes in the shared scratch -- and compares with t4c_fk bit for bit (the same kernel, the rule of test_fk).
"""
import ctypes as C

import numpy as np
import pytest

from tesseract_b200 import capi, scene as S
from tesseract_b200.managers import BatchChecker


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _robot(lib=None):
    b = S.SceneBuilder()
    S.add_robot(b, S.load_robot("iiwa14"))
    return b, BatchChecker(b, lib=lib)


def _noalloc(bc, q):
    q = np.ascontiguousarray(q, dtype=np.float64).reshape(-1, bc.fk_desc.n_q)
    out = np.full((len(q), len(bc.b.link_names), 3, 4), np.nan)
    st = bc.lib.t4c_fk_noalloc(bc.model, bc.scene.h, _dp(q), len(q), _dp(out))
    return st, out


def _exercise(bc, b):
    lib = bc.lib
    nq, nl = bc.fk_desc.n_q, len(b.link_names)
    per_state = nq * 8 + nl * 96
    n_max = (capi.FK_NOALLOC_SCRATCH_BYTES - 256) // per_state
    # nothing resident yet: the entry refuses instead of uploading
    q0 = S.random_states(b, 16, seed=1, dtype=np.float64)
    st, _ = _noalloc(bc, q0)
    assert st == capi.T4C_ERR_INVALID
    assert b"never uploads" in lib.t4c_last_error()
    ref0 = bc.fk(q0)  # the existing entry makes scene + model resident
    # two calls in a row with different inputs and sizes: the second runs over the first one's leftovers
    qa = S.random_states(b, min(3000, n_max), seed=2, dtype=np.float64)
    qb = S.random_states(b, 777, seed=3, dtype=np.float64)
    for q in (qa, qb, q0):
        st, out = _noalloc(bc, q)
        capi.check(st, lib)
        ref = bc.fk(q)
        assert not np.isnan(out).any()
        bad = int((out.view(np.uint64) != ref.view(np.uint64)).sum())
        assert bad == 0, f"{bad} of {out.size} pose elements differ from t4c_fk (max {float(np.abs(out - ref).max())})"
    assert np.array_equal(_noalloc(bc, q0)[1], ref0)
    # the largest batch that fits, and one state more
    qm = S.random_states(b, n_max, seed=4, dtype=np.float64)
    st, out = _noalloc(bc, qm)
    capi.check(st, lib)
    assert np.array_equal(out, bc.fk(qm))
    st, _ = _noalloc(bc, S.random_states(b, n_max + 1, seed=5, dtype=np.float64))
    assert st == capi.T4C_ERR_CAPACITY
    assert b"fixed array" in lib.t4c_last_error()
    # empty batch
    st, _ = _noalloc(bc, np.zeros((0, nq)))
    assert st == capi.T4C_OK


def test_fk_noalloc_host_logic():
    """CPU tier: the entry's checks and staging layout on the host simulator (same code, host array)"""
    from hostsim import build as hostsim
    b, bc = _robot(hostsim.lib())
    _exercise(bc, b)


@pytest.mark.gpu
def test_fk_noalloc_matches_fk_on_the_device():
    b, bc = _robot()
    _exercise(bc, b)
