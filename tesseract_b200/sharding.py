"""Multi-GPU plumbing of the batched check: one process per GPU, contiguous block partition of the state range,
scene replicated on every rank, ONE exchange at the end -- a gather of the per-state hit flags to rank 0
(NCCL over NVLink on the GPU box, gloo in the CPU tests).  The reference relies on the same independence
with one manager clone per thread (tesseract_collision/include/tesseract_collision/test_suite/
collision_multi_threaded_unit.hpp:79-127)."""
import numpy as np


def shard_range(n, rank, world):
    """Contiguous block partition: n_g = ceil(n / world); the last ranks may get a short (or empty) block."""
    n_g = (n + world - 1) // world
    lo = min(n, rank * n_g)
    return lo, min(n, lo + n_g)


def gather_hits(hit_local, n_total, dist, device=None, dst=0):
    """Gather the per-rank hit flags (uint8) to rank `dst`; returns the full (n_total,) array there, None elsewhere.
    Blocks are padded to n_g so that a single fixed-size gather suffices."""
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    n_g = (n_total + world - 1) // world
    t = torch.zeros(max(n_g, 1), dtype=torch.uint8, device=device)
    if isinstance(hit_local, torch.Tensor):
        t[:hit_local.numel()] = hit_local
    else:
        t[:len(hit_local)] = torch.from_numpy(np.ascontiguousarray(hit_local, dtype=np.uint8)).to(t.device)
    out = [torch.empty_like(t) for _ in range(world)] if rank == dst else None
    dist.gather(t, out, dst=dst)
    if rank != dst:
        return None
    full = torch.cat(out)[:n_total] if n_g * world >= n_total else torch.cat(out)
    return full.cpu().numpy()


def check_discrete_sharded(checker, q, dist, test_type=0, device=None):
    """Each rank checks its block of `q` (same array on every rank, e.g. seeded synthetic states) and rank 0 receives
    all flags.  `checker` is a tesseract_b200.managers.BatchChecker built on this rank's device."""
    lo, hi = shard_range(len(q), dist.get_rank(), dist.get_world_size())
    if hi > lo:
        hit, _, _ = checker.check_discrete(q[lo:hi], test_type, want_contacts=False)
    else:
        hit = np.zeros(0, dtype=np.uint8)
    return gather_hits(hit, len(q), dist, device)
