"""Multi-rank host logic on CPU (gloo, world_size 2): the batch shards without overlap and the solution gather
restores global problem order.  The data path itself has no collective."""
import os
import socket

import numpy as np
import pytest


def test_shard_ranges_partition_the_batch():
    from towr2025_b200.sharding import shard_range
    for n in (1, 7, 4096, 4097, 65536):
        for world in (1, 2, 3, 4, 8):
            cover = []
            for r in range(world):
                lo, hi = shard_range(n, r, world)
                assert 0 <= lo <= hi <= n
                cover += list(range(lo, hi))
            assert cover == list(range(n))
            sizes = [shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _worker(rank, world, port, n_problems, n_vars, q):
    import torch.distributed as dist
    from towr2025_b200.sharding import gather_solutions, shard_range
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_range(n_problems, rank, world)
    # the "solution" of global problem i is a vector filled with i + j/1000: checks order and trimming
    x = np.arange(lo, hi, dtype=np.float64)[:, None] + np.arange(n_vars, dtype=np.float64)[None, :] / 1000.0
    full = gather_solutions(x, n_problems).numpy()
    q.put((rank, full))
    dist.barrier()
    dist.destroy_process_group()


def test_gather_solutions_gloo_world2():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    n_problems, n_vars, world = 7, 5, 2
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_problems, n_vars, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    expect = np.arange(n_problems, dtype=np.float64)[:, None] + np.arange(n_vars, dtype=np.float64)[None, :] / 1000.0
    for r in range(world):
        assert np.array_equal(got[r], expect)
