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
# (mode, warps per SM, a, b, delta factor): mode 0 = grouped warp kernel, a = ILP, b = ring entries per tree;
#                                           mode 1 = sssp_group_kernel, a = pops per batch, b = prefetch
configs = json.loads(os.environ.get("CONFIGS", "null")) or [
    (0, 16, 4, 1024, 4.0, 0), (2, 8, 16, 0, 4.0, 0), (2, 16, 8, 0, 4.0, 0), (2, 12, 8, 0, 4.0, 0), (2, 8, 8, 0, 4.0, 0),
    (2, 16, 4, 0, 4.0, 0), (2, 4, 16, 0, 4.0, 0), (2, 8, 16, 0, 2.0, 0), (2, 8, 16, 0, 8.0, 0), (2, 16, 8, 0, 8.0, 0)]
for mode, wps, a, b, df, rm in configs:
    try:
        if mode == 0:
            g.set_group_mode(0, b)
            g.set_warp_tuning(ilp=a)
            g.set_group_tuning(warps_per_sm=wps)
        elif mode == 2:
            g.set_group_mode(2)
            g.set_group_tuning(warps_per_sm=wps, batch=a)
        else:
            g.set_group_mode(1)
            g.set_group_tuning(warps_per_sm=wps, batch=a, prefetch=b)
        g.set_tuning(delta_factor=df)
        dist.zero_(); par.zero_()
        ms = timed(lambda: g.sssp_batch_device(o, dist, par))
        ok = bool(torch.equal(dist[:chk].view(torch.int64), dref.view(torch.int64))) and bool(torch.equal(par[:chk], pref))
        ms_tree = timed(lambda: g.sssp_batch_device(o, None, None), reps=2)
        g.sssp_batch(origins[:64], want_dist=False)
        st = g.last_stats()
        tps = B / ms * 1e3
        print(json.dumps({"algo": "group", "mode": mode, "warps_per_sm": wps, "a": a, "b": b, "delta": df, "relax_mode": rm, "ms": round(ms, 2),
                          "ms_trees_only": round(ms_tree, 2), "trees_per_s": round(tps), "frac": round(tps * btree / (peak * 1e9), 4),
                          "bit_exact_vs_warp": ok, "overflows": st["overflows"]}), flush=True)
    except Exception as ex:
        print("ERR", mode, wps, a, b, df, rm, ex, flush=True)
