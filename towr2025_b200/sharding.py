"""Sharding of a problem batch over the ranks of one job (one process per GPU).

Problems are independent: each rank evaluates a contiguous range of the global problem index with
no collective on the data path.  The only exchange is the gather of solution vectors at the end
(`gather_solutions`), over NCCL on GPUs (NVLink/NVSwitch) or gloo in the CPU tests."""
import numpy as np


def shard_range(n_problems, rank, world):
    """Contiguous [lo, hi) of the global problem index owned by `rank`; sizes differ by at most one."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    base, rem = divmod(n_problems, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_solutions(x_local, n_problems, device=None):
    """all_gather the per-rank solution blocks (n_local x n_vars, float64) into the global (n_problems x n_vars)
    array, in global problem order, on every rank.  Uneven shards are padded to the largest shard for the
    collective and trimmed afterwards."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(), dist.get_rank()
    lo, hi = shard_range(n_problems, rank, world)
    t = torch.as_tensor(np.ascontiguousarray(x_local) if isinstance(x_local, np.ndarray) else x_local)
    assert t.shape[0] == hi - lo
    if device is not None:
        t = t.to(device)
    n_max = -(-n_problems // world)
    pad = torch.zeros((n_max, t.shape[1]), dtype=t.dtype, device=t.device)
    pad[: t.shape[0]] = t
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad)
    parts = []
    for r in range(world):
        a, b = shard_range(n_problems, r, world)
        parts.append(out[r][: b - a])
    return torch.cat(parts, dim=0)
