"""GPU tier, BASELINE.json configs[1] at FULL size (1 Mi joint states, 10k primitives): size-independent properties
of the batched check plus an oracle comparison on a random sub-sample."""
import os

import numpy as np
import pytest

from oracle import pyoracle
from tesseract_b200 import capi, scene as S
from tesseract_b200.managers import BatchChecker

pytestmark = pytest.mark.gpu
N = 1 << 20
MARGIN = 0.05


@pytest.fixture(scope="module")
def setup():
    b = S.build_config2(n_prims=10000, margin=MARGIN)
    q = S.random_states(b, N, seed=1234, dtype=np.float64)
    return b, q, BatchChecker(b)


def test_flag_agreement_and_first_records(setup):
    b, q, bc = setup
    flags, _, st0 = bc.check_discrete(q, capi.TEST_FIRST, want_contacts=False)
    hit1, con1, st1 = bc.check_discrete(q, capi.TEST_FIRST, capacity=N)
    # FIRST with records: exactly one record per colliding state, flags identical to the flags-only pass
    assert np.array_equal(flags, hit1)
    assert len(con1) == int(hit1.sum()) == st1["colliding_states"]
    assert len(np.unique(con1["state"])) == len(con1)
    assert np.array_equal(np.sort(con1["state"]), np.flatnonzero(hit1))
    assert (con1["distance"] <= MARGIN + 1e-6).all()
    nrm = np.linalg.norm(con1["normal"].astype(np.float64), axis=1)
    assert np.abs(nrm[con1["distance"] > -1e30] - 1.0).max() < 1e-4
    # idempotence: a second pass over the same batch gives the same flags
    flags2, _, _ = bc.check_discrete(q, capi.TEST_FIRST, want_contacts=False)
    assert np.array_equal(flags, flags2)
    # oracle on a random sub-sample of the full batch
    idx = np.sort(np.random.default_rng(3).choice(N, 4000, replace=False))
    hit_o, _, _ = pyoracle.check_discrete(b, q[idx], capi.TEST_FIRST, want_contacts=False,
                                          n_threads=os.cpu_count() or 4)
    assert np.array_equal(flags[idx], hit_o)


def test_closest_vs_first_and_all(setup):
    b, q, bc = setup
    n = 1 << 17
    hit_c, con_c, _ = bc.check_discrete(q[:n], capi.TEST_CLOSEST)
    hit_a, con_a, _ = bc.check_discrete(q[:n], capi.TEST_ALL)
    hit_f, _, _ = bc.check_discrete(q[:n], capi.TEST_FIRST, want_contacts=False)
    assert np.array_equal(hit_c, hit_f) and np.array_equal(hit_a, hit_f)
    assert (con_c["distance"] <= MARGIN + 1e-6).all()
    # CLOSEST = per (state, link pair) minimum of ALL (core/common.h:151-157)
    def keyed_min(con):
        lo = np.minimum(con["link"][:, 0], con["link"][:, 1]).astype(np.int64)
        hi = np.maximum(con["link"][:, 0], con["link"][:, 1]).astype(np.int64)
        key = (con["state"].astype(np.int64) << 40) | (lo << 20) | hi
        order = np.lexsort((con["distance"], key))
        k, d = key[order], con["distance"][order]
        first = np.r_[True, k[1:] != k[:-1]]
        return k[first], d[first]
    kc, dc = keyed_min(con_c)
    ka, da = keyed_min(con_a)
    assert len(con_c) == len(kc), "CLOSEST must hold one record per key"
    assert np.array_equal(kc, ka)
    assert np.array_equal(dc, da)


def test_phase1_certificate_never_contradicted():
    """The exact pass skips its boolean GJK (phase 1) for pairs the FP32 cull certified as more than 1 mm apart.  Under
    the T4C_VERIFY_SKIP hook the phase runs anyway and every certified pair whose boolean GJK says "intersect" is
    counted: 1 Mi states of configs[1] in FIRST and 256 Ki in CLOSEST mode, zero contradictions allowed."""
    import subprocess, sys
    code = ("import numpy as np; from tesseract_b200 import capi, scene as S; "
            "from tesseract_b200.managers import BatchChecker; "
            "b = S.build_config2(n_prims=10000, margin=0.05); bc = BatchChecker(b); "
            "q = S.random_states(b, 1 << 20, seed=1234, dtype=np.float64); "
            "h, c, st = bc.check_discrete(q, capi.TEST_FIRST, capacity=1 << 20); "
            "print('FIRST', st['phase1_skipped'], st['phase1_skip_violations']); "
            "assert st['phase1_skipped'] > 100000 and st['phase1_skip_violations'] == 0, st; "
            "h, c, st = bc.check_discrete(q[:1 << 18], capi.TEST_CLOSEST); "
            "print('CLOSEST', st['phase1_skipped'], st['phase1_skip_violations']); "
            "assert st['phase1_skipped'] > 100000 and st['phase1_skip_violations'] == 0, st")
    env = dict(os.environ, T4C_VERIFY_SKIP="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=1200)
    print(r.stdout[-1000:])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
