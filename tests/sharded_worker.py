"""One rank of the process-per-GPU test of t4c_check_discrete_sharded (tests/test_gpu_multi.py starts `world` of these).

    python tests/sharded_worker.py <rank> <world> <id file> <n states per rank>

Every rank checks its own shard of one seeded batch on device `rank`; the records reach the root through the gather
slab (CUDA IPC mapping between the processes, pushed while the rounds compute) and the final NCCL exchange.  The root
then runs the WHOLE batch on its own device and compares: same flags, one FIRST record per colliding state, every
gathered record equal field for field to the single-device ALL record of the same (state, link pair, shapes); CLOSEST
records equal as a set.  Prints "SHARDED OK" on the root.
"""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank, world, idfile, per = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], int(sys.argv[4])
    import torch
    from tesseract_b200 import capi, scene as S
    from tesseract_b200.managers import BatchChecker
    import parity

    lib = capi.load_library()
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    b = S.build_config2(n_prims=10000, margin=0.05)
    desc = b.scene_desc()
    scene = C.c_void_p()
    capi.check(lib.t4c_scene_create(C.byref(desc), rank, C.byref(scene)), lib)
    fkd = b.fk_desc()
    model = C.c_void_p()
    capi.check(lib.t4c_fk_model_create(C.byref(fkd), scene, C.byref(model)), lib)
    # the 128-byte communicator id travels through a file
    buf = (C.c_uint8 * 128)()
    if rank == 0:
        capi.check(lib.t4c_comm_unique_id(buf), lib)
        with open(idfile + ".tmp", "wb") as f:
            f.write(bytes(buf))
        os.replace(idfile + ".tmp", idfile)
    else:
        t0 = time.time()
        while not os.path.exists(idfile):
            if time.time() - t0 > 120:
                raise RuntimeError("no communicator id")
            time.sleep(0.05)
        buf = (C.c_uint8 * 128)(*open(idfile, "rb").read())
    comm = C.c_void_p()
    capi.check(lib.t4c_comm_create(buf, rank, world, rank, C.byref(comm)), lib)

    n = per * world - 3  # ragged last shard
    q = S.random_states(b, n, seed=99, dtype=np.float64)
    a, e = rank * per, min(n, (rank + 1) * per)
    m = e - a
    q_dev = torch.from_numpy(q[a:e].copy()).to(dev)
    for tt, cps in ((capi.TEST_FIRST, 1), (capi.TEST_FIRST, 1), (capi.TEST_CLOSEST, 16)):
        cap = cps * per
        hit_dev = torch.zeros(max(m, 1), dtype=torch.uint8, device=dev)
        con_dev = torch.zeros(cap * 80, dtype=torch.uint8, device=dev)
        io = capi.DeviceIO()
        io.q, io.q_layout = q_dev.data_ptr(), capi.Q_F64_ROWMAJOR
        io.hit, io.contacts, io.capacity = hit_dev.data_ptr(), con_dev.data_ptr(), cap
        g = capi.Gathered()
        if rank == 0:
            g_hit = torch.zeros(n, dtype=torch.uint8, device=dev)
            g_con = torch.zeros(world * cap * 80, dtype=torch.uint8, device=dev)
            g.hit, g.contacts, g.capacity = g_hit.data_ptr(), g_con.data_ptr(), world * cap
        req = capi.Request(tt, 1, 1)
        capi.check(lib.t4c_check_discrete_sharded(scene, model, comm, 0, m, a, C.byref(req), C.byref(io),
                                                  C.byref(g) if rank == 0 else None), lib)
        torch.cuda.synchronize()
        if rank != 0:
            continue
        hit_m = g_hit.cpu().numpy()
        con_m = g_con.cpu().numpy()[:int(g.n_contacts) * 80].view(capi.CONTACT_DTYPE)
        assert int(g.n_states) == n, (g.n_states, n)
        bc = BatchChecker(b)
        if tt == capi.TEST_FIRST:
            hit_s, con_a, _ = bc.check_discrete(q, capi.TEST_ALL)
            assert np.array_equal(hit_m, hit_s), "flags differ from the single-device run"
            assert np.array_equal(np.sort(con_m["state"]), np.flatnonzero(hit_m)), "not one record per colliding state"
            Ga, Gm = parity.index(con_a, True), parity.index(con_m, True)
            missing = [k for k in Gm if k not in Ga]
            assert not missing, f"{len(missing)} gathered records have no counterpart in the ALL result"
            assert max(abs(Gm[k][0] - Ga[k][0]) for k in Gm) == 0.0
            assert max(float(np.abs(np.asarray(Gm[k][1]) - np.asarray(Ga[k][1])).max()) for k in Gm) == 0.0
        else:
            hit_s, con_s, _ = bc.check_discrete(q, tt)
            assert np.array_equal(hit_m, hit_s)
            Gs, Gm = parity.index(con_s, False), parity.index(con_m, False)
            assert set(Gs) == set(Gm)
            assert max(abs(Gm[k][0] - Gs[k][0]) for k in Gm) == 0.0
        print("mode", tt, "records", len(con_m), "colliding", int(hit_m.sum()), flush=True)
    lib.t4c_comm_destroy(comm)
    lib.t4c_fk_model_destroy(model)
    lib.t4c_scene_destroy(scene)
    if rank == 0:
        print("SHARDED OK", flush=True)


if __name__ == "__main__":
    main()
