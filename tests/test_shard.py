"""N > 1 path of the batched reroute on CPU: two gloo ranks shard the origin list, compute their trees (with the
C oracle standing in for the GPU kernel -- this test is about the sharding and the collective, not the kernel) and
all-gather the next-hop tables; every rank must end up with exactly the single-process result."""
import os
import socket

import numpy as np
import pytest

from conftest import grid_graph
from spatial_queue_b200 import shard


def test_shard_bounds_cover_everything():
    for n in (0, 1, 7, 64, 1001):
        for world in (1, 2, 3, 8):
            b = [shard.shard_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, n_origins, q):
    import torch
    import torch.distributed as dist
    from oracle import sssp as osssp
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    s, e, w = grid_graph(12, seed=3)
    origins = np.random.default_rng(5).permutation(144)[:n_origins].astype(np.int32)
    mine = shard.shard_origins(origins, rank, world)
    _, parent = osssp.OracleGraph(s, e, w).sssp_batch(mine)
    full = shard.all_gather_tables(torch.from_numpy(parent), len(origins))
    only0 = shard.gather_tables_to(torch.from_numpy(parent), len(origins), dst=0)
    assert (only0 is None) == (rank != 0)
    if rank == 0:
        assert torch.equal(only0, full)
    q.put((rank, full.numpy().copy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_origins", [16, 17])
def test_two_ranks_gather_the_single_process_result(n_origins):
    import torch.multiprocessing as mp
    from oracle import sssp as osssp
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_origins, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    s, e, w = grid_graph(12, seed=3)
    origins = np.random.default_rng(5).permutation(144)[:n_origins].astype(np.int32)
    _, want = osssp.OracleGraph(s, e, w).sssp_batch(origins)
    for r in range(2):
        assert got[r].shape == want.shape and (got[r] == want).all()
