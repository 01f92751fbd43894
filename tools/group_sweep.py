"""Tuning sweep of the group kernel (sp_group.cuh) on the config-5 workload of bench.py (run on the GPU box).
Prints one JSON line per configuration: whole step (sort + trees + un-interleave / predecessors), trees only, pops."""
import json
import os
import sys

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import numpy as np
import torch

import bench
from spatial_queue_b200 import sp_interface as spi

n = int(os.environ.get("GRID", 224))
B = int(os.environ.get("BATCH", 18944))
s, e, w, V = bench.sssp_workload(n)
g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
origins = bench.origins_for(0, B, V, step=0)
o = torch.from_numpy(origins).cuda()
dist = torch.empty((B, V), dtype=torch.float64, device="cuda")
par = torch.empty((B, V), dtype=torch.int32, device="cuda")
btree = 20 * len(s) + 24 * V
peak = bench.measured_peaks()[0]


def timed(fn, reps=3):
    fn(); torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best


# reference tables from the warp kernel (pinned to the oracle by tests/test_sssp_gpu.py) for a bitwise check
g.set_algo(2)
chk = 1024
dref = torch.empty((chk, V), dtype=torch.float64, device="cuda")
pref = torch.empty((chk, V), dtype=torch.int32, device="cuda")
g.sssp_batch_device(o[:chk], dref, pref)
ms_warp = timed(lambda: g.sssp_batch_device(o, dist, par))
print(json.dumps({"algo": "warp", "ms": round(ms_warp, 2), "trees_per_s": round(B / ms_warp * 1e3),
                  "frac": round(B / ms_warp * 1e3 * btree / (peak * 1e9), 4)}), flush=True)

g.set_algo(4)
configs = json.loads(os.environ.get("CONFIGS", "null")) or [
    (16, 16, 1, 4.0), (16, 16, 0, 4.0), (16, 32, 1, 4.0), (16, 8, 1, 4.0), (8, 32, 1, 4.0), (8, 16, 1, 4.0), (12, 16, 1, 4.0),
    (24, 16, 1, 4.0), (32, 16, 1, 4.0), (32, 8, 1, 4.0), (16, 16, 1, 2.0), (16, 16, 1, 8.0), (16, 32, 0, 4.0), (32, 32, 0, 4.0)]
for wps, pb, pf, df in configs:
    try:
        g.set_group_tuning(warps_per_sm=wps, batch=pb, prefetch=pf)
        g.set_tuning(delta_factor=df)
        dist.zero_(); par.zero_()
        ms = timed(lambda: g.sssp_batch_device(o, dist, par))
        ok = bool(torch.equal(dist[:chk].view(torch.int64), dref.view(torch.int64))) and bool(torch.equal(par[:chk], pref))
        ms_tree = timed(lambda: g.sssp_batch_device(o, None, None), reps=2)
        g.sssp_batch_device(o[:2048], None, None); torch.cuda.synchronize()
        st = g.last_stats() if False else None
        tps = B / ms * 1e3
        print(json.dumps({"algo": "group", "warps_per_sm": wps, "batch": pb, "prefetch": pf, "delta": df, "ms": round(ms, 2),
                          "ms_trees_only": round(ms_tree, 2), "trees_per_s": round(tps), "frac": round(tps * btree / (peak * 1e9), 4),
                          "bit_exact_vs_warp": ok}), flush=True)
    except Exception as ex:
        print("ERR", wps, pb, pf, df, ex, flush=True)
