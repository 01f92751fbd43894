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


def gather_tables_to(local, n_total, dst=0, dist=None, group=None):
    """The north-star form of the exchange: the tables go to the SIMULATING rank only (`dst`), which gets the
    [n_total, V] table in rank order; every other rank gets None.  Same padding rule as all_gather_tables."""
    import torch
    import torch.distributed as tdist
    dist = dist or tdist
    world = dist.get_world_size(group)
    if world == 1:
        return local
    rank = dist.get_rank(group)
    V = local.shape[1]
    rows = max(shard_bounds(n_total, r, world)[1] - shard_bounds(n_total, r, world)[0] for r in range(world))
    send = local
    if local.shape[0] != rows:
        send = torch.empty((rows, V), dtype=local.dtype, device=local.device)
        send[: local.shape[0]] = local
    parts = [torch.empty((rows, V), dtype=local.dtype, device=local.device) for _ in range(world)] if rank == dst else None
    dist.gather(send.contiguous(), parts, dst=dst, group=group)
    if rank != dst:
        return None
    out = []
    for r in range(world):
        lo, hi = shard_bounds(n_total, r, world)
        out.append(parts[r][: hi - lo])
    return torch.cat(out, 0)


class PeerTables:
    """All-gather of next-hop tables without SMs: every rank owns a [slots, world * rows, V] buffer in symmetric
    memory (torch.distributed._symmetric_memory: the peers' buffers are mapped into this process over
    NVLink / NVSwitch) and *pushes* its finished rows into everybody's copy with plain device-to-device
    copies, which the driver runs on the copy engines.  The tree kernel keeps every SM; an NCCL all-gather
    running next to it fights it for SM resources and stretched the 8-GPU step by 28 %.

    push() is stream-ordered after the trees it ships; fence() makes every rank's earlier pushes visible to every
    rank (a device-side barrier on the communication stream)."""

    def __init__(self, slots, rows, V, dtype, group=None):
        import torch
        import torch.distributed as tdist
        import torch.distributed._symmetric_memory as symm
        self.group = group or tdist.group.WORLD
        self.world = tdist.get_world_size(self.group)
        self.rank = tdist.get_rank(self.group)
        self.rows = rows
        shape = (slots, self.world * rows, V)
        self.local = symm.empty(shape, dtype=dtype, device=torch.device("cuda", torch.cuda.current_device()))
        self.handle = symm.rendezvous(self.local, self.group)
        self.peers = [self.local if r == self.rank else self.handle.get_buffer(r, shape, dtype)
                      for r in range(self.world)]
        self.stream = torch.cuda.Stream()

    def push(self, slot, src, after=None, to=None):
        """Copy `src` ([rows, V], this rank's finished chunk) into row block `rank` of slot `slot` on every rank, or,
        with `to` = a rank, into that rank's table only (the tables are needed where the simulation runs)."""
        import torch
        if after is not None:
            self.stream.wait_event(after)
        lo = self.rank * self.rows
        with torch.cuda.stream(self.stream):
            if to is not None:
                self.peers[to][slot, lo: lo + src.shape[0]].copy_(src, non_blocking=True)
                return
            for k in range(self.world):
                r = (self.rank + k) % self.world          # start with the own copy, then walk the ring
                self.peers[r][slot, lo: lo + src.shape[0]].copy_(src, non_blocking=True)

    def fence(self):
        import torch
        with torch.cuda.stream(self.stream):
            self.handle.barrier(channel=0)
        ev = torch.cuda.Event()
        ev.record(self.stream)
        return ev

    def table(self, slot):
        return self.local[slot]
