"""CPU tier: the N > 1 path (state sharding + the single gather of hit flags) with torch.distributed gloo,
world_size 2 and 3; kernels run in the host simulator; rank 0 compares the gathered flags with the CPU oracle."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_states, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from hostsim import build as hostsim
    from tesseract_b200 import capi, scene as S, sharding
    from tesseract_b200.managers import BatchChecker
    b = S.build_config2(n_prims=2000, margin=0.05)
    q = S.random_states(b, n_states, seed=4321, dtype=np.float64)
    bc = BatchChecker(b, lib=hostsim.lib())
    full = sharding.check_discrete_sharded(bc, q, dist, capi.TEST_FIRST)
    if rank == 0:
        np.save(out_path, full)
    else:
        assert full is None
    dist.barrier()
    dist.destroy_process_group()


def test_shard_range_covers_everything():
    from tesseract_b200.sharding import shard_range
    for n in (0, 1, 7, 64, 1000003):
        for w in (1, 2, 3, 4, 8):
            blocks = [shard_range(n, r, w) for r in range(w)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            for (a0, a1), (b0, b1) in zip(blocks, blocks[1:]):
                assert a1 == b0 and a0 <= a1


@pytest.mark.parametrize("world,n_states", [(2, 96), (3, 50)])
def test_sharded_check_matches_oracle(tmp_path, world, n_states):
    from hostsim import build as hostsim
    hostsim.build()
    out = str(tmp_path / "hits.npy")
    mp.spawn(_worker, args=(world, _free_port(), n_states, out), nprocs=world, join=True)
    got = np.load(out)
    from oracle import pyoracle
    from tesseract_b200 import capi, scene as S
    b = S.build_config2(n_prims=2000, margin=0.05)
    q = S.random_states(b, n_states, seed=4321, dtype=np.float64)
    hit_o, _, _ = pyoracle.check_discrete(b, q, capi.TEST_FIRST, want_contacts=False, n_threads=4)
    assert got.shape == (n_states,)
    assert np.array_equal(got, hit_o)
