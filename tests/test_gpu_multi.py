"""GPU tier, two or more devices: the multi-GPU entries of the C ABI (t4_multi.h) against the single-device path and the
CPU oracle.  One process, one host thread per device (t4c_check_discrete_multi): contiguous shards, scene replicated,
hit flags + contact records + counters gathered on device 0 by one grouped ncclSend / ncclRecv exchange."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import pyoracle
from tesseract_b200 import capi, scene as S
from tesseract_b200.managers import BatchChecker, create_scene

import parity

pytestmark = pytest.mark.gpu


def n_devices():
    try:
        return capi.load_library().t4c_device_count()
    except Exception:
        return 0


@pytest.fixture(scope="module")
def rig():
    nd = min(n_devices(), 4)
    if nd < 2:
        pytest.skip("needs at least two CUDA devices")
    lib = capi.load_library()
    b = S.build_config2(n_prims=10000, margin=0.05)
    scenes = [create_scene(lib, b, d) for d in range(nd)]
    fkd = b.fk_desc()
    model = C.c_void_p()
    capi.check(lib.t4c_fk_model_create(C.byref(fkd), scenes[0].h, C.byref(model)), lib)
    devs = (C.c_int32 * nd)(*range(nd))
    comms = (C.c_void_p * nd)()
    capi.check(lib.t4c_comm_create_local(devs, nd, comms), lib)
    yield lib, b, scenes, model, comms, nd
    for c in comms:
        lib.t4c_comm_destroy(c)
    lib.t4c_fk_model_destroy(model)


def run_multi(lib, scenes, model, comms, nd, q, tt, cap):
    n = len(q)
    hit = np.zeros(n, dtype=np.uint8)
    con = np.zeros(cap, dtype=capi.CONTACT_DTYPE)
    res = capi.Result()
    res.hit = hit.ctypes.data_as(C.POINTER(C.c_uint8))
    res.contacts, res.capacity = con.ctypes.data, cap
    hs = (C.c_void_p * nd)(*[s.h for s in scenes])
    req = capi.Request(tt, 1, 1)
    capi.check(lib.t4c_check_discrete_multi(hs, model, comms, nd, q.ctypes.data_as(C.POINTER(C.c_double)), n,
                                            C.byref(req), C.byref(res)), lib)
    return hit, con[:int(res.n_contacts)], res.stats.as_dict()


@pytest.mark.parametrize("tt", [capi.TEST_FIRST, capi.TEST_CLOSEST])
def test_multi_matches_single_device_and_oracle(rig, tt):
    lib, b, scenes, model, comms, nd = rig
    n = 60000 + 7  # not a multiple of the device count: ragged last shard
    q = S.random_states(b, n, seed=4321, dtype=np.float64)
    hit_m, con_m, st_m = run_multi(lib, scenes, model, comms, nd, q, tt, 16 * n)
    bc = BatchChecker(b)
    hit_s, con_s, st_s = bc.check_discrete(q, tt)
    assert np.array_equal(hit_m, hit_s)
    assert st_m["states"] == n and st_m["colliding_states"] == int(hit_s.sum())
    assert len(con_m) == len(con_s)
    if tt == capi.TEST_CLOSEST:
        # same records whatever the sharding: compare as sets keyed by (state, link pair)
        Gm, Gs = parity.index(con_m, False), parity.index(con_s, False)
        assert set(Gm) == set(Gs)
        assert max(abs(Gm[k][0] - Gs[k][0]) for k in Gm) == 0.0
        # and against the oracle on the states around a shard boundary
        per = (n + nd - 1) // nd
        lo = per - 1500
        sl = slice(lo, lo + 3000)
        hit_o, con_o, _ = pyoracle.check_discrete(b, q[sl], tt, n_threads=os.cpu_count() or 4)
        sub = con_m[(con_m["state"] >= lo) & (con_m["state"] < lo + 3000)].copy()
        sub["state"] -= lo
        rep = parity.compare(b, sub, con_o, hit_m[sl], hit_o, lambda a, c: 0.05, link_poses=pyoracle.fk(b, q[sl]))
        print(rep)
    else:
        assert np.array_equal(np.sort(con_m["state"]), np.flatnonzero(hit_m))
        # every gathered record (most of them reach the root through the progressive peer-memory push) is, field for
        # field, the record the single-device ALL query holds for the same (state, link pair, shapes)
        hit_a, con_a, _ = bc.check_discrete(q, capi.TEST_ALL)
        Ga = parity.index(con_a, True)
        Gm = parity.index(con_m, True)
        missing = [k for k in Gm if k not in Ga]
        assert not missing, f"{len(missing)} gathered FIRST records have no counterpart in the ALL result"
        assert max(abs(Gm[k][0] - Ga[k][0]) for k in Gm) == 0.0
        assert max(float(np.abs(np.asarray(Gm[k][1]) - np.asarray(Ga[k][1])).max()) for k in Gm) == 0.0
        idx = np.sort(np.random.default_rng(5).choice(n, 3000, replace=False))
        hit_o, _, _ = pyoracle.check_discrete(b, q[idx], tt, want_contacts=False, n_threads=os.cpu_count() or 4)
        assert np.array_equal(hit_m[idx], hit_o)


def test_sharded_processes_push_records_over_peer_memory(tmp_path):
    """One PROCESS per device (the torchrun layout of bench.py): t4c_check_discrete_sharded with the root's gather slab
    mapped into the peers through CUDA IPC; records are pushed while the rounds compute (test hooks make the small
    shards run in rounds and push in small pieces).  tests/sharded_worker.py holds the comparison."""
    import subprocess
    import sys
    nd = min(n_devices(), 2)
    if nd < 2:
        pytest.skip("needs at least two CUDA devices")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, T4C_ROUND_MIN_STATES="1", T4C_DRAIN_MIN="256", T4C_TRACE_PASSES="1")
    idfile = str(tmp_path / "comm_id")
    procs = [subprocess.Popen([sys.executable, os.path.join(root, "tests", "sharded_worker.py"), str(r), str(nd), idfile,
                               "20000"], cwd=root, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(nd)]
    outs = []
    try:
        for p in procs:
            outs.append(p.communicate(timeout=900)[0])
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, f"rank {r}:\n{o[-3000:]}"
    assert "SHARDED OK" in outs[0], outs[0][-3000:]
    # the records did travel through the slab (a rank that could not map it would still pass: it sends everything)
    import re
    pushed = [int(m.group(1)) for m in re.finditer(r"pushed (\d+) of \d+ records", outs[0])]
    assert pushed and max(pushed) > 0, "no record reached the root through the peer-memory push"
    print(outs[0][-400:])
