"""Multi-GPU plumbing of the batched reroute: trees are independent, so the origin (or destination)
list is split into contiguous chunks, one per rank, every rank holds a replica of the routing graph, and one
collective returns the int32 next-hop tables to every rank (NCCL over NVLink on GPUs; gloo in the CPU tests).

The reference's only parallel strategy is the same split over a multiprocessing.Pool (util/dta_meso.py:513-536);
there is no other exchange step on this path, hence no other collective."""
import numpy as np


def shard_bounds(n, rank, world):
    """[lo, hi) of rank's contiguous chunk of n items; chunk sizes differ by at most one."""
    base, extra = divmod(int(n), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_origins(origins, rank, world):
    lo, hi = shard_bounds(len(origins), rank, world)
    return origins[lo:hi]


def all_gather_tables(local, n_total, dist=None, group=None):
    """Concatenate the per-rank [n_rank, V] tables in rank order -> [n_total, V] on every rank.

    `local` is a torch tensor (CUDA with NCCL, CPU with gloo).  Ranks may hold chunks that differ by one row
    (shard_bounds); rows are padded to the largest chunk for the collective and trimmed afterwards."""
    import torch
    import torch.distributed as tdist
    dist = dist or tdist
    world = dist.get_world_size(group)
    if world == 1:
        return local
    V = local.shape[1]
    rows = max(shard_bounds(n_total, r, world)[1] - shard_bounds(n_total, r, world)[0] for r in range(world))
    send = local
    if local.shape[0] != rows:
        send = torch.empty((rows, V), dtype=local.dtype, device=local.device)
        send[: local.shape[0]] = local
    out = torch.empty((world * rows, V), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, send.contiguous(), group=group)
    if n_total == world * rows:
        return out
    parts = []
    for r in range(world):
        lo, hi = shard_bounds(n_total, r, world)
        parts.append(out[r * rows: r * rows + (hi - lo)])
    return torch.cat(parts, 0)
