"""Warm timeline of the two phase kernels of a tick (run on the GPU box): CASE = c4 (default) / c2 / c3 of
bench.tick_cases, T0 = first tick of the window, W = ticks recorded.  Uses sq_enable_timeline (%globaltimer at entry and
exit of every warp), so the numbers come from an ordinary graph-replayed run, not from a profiler.  One JSON line."""
import argparse
import json
import os
import sys

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import numpy as np
import torch

import bench

T0 = int(os.environ.get("T0", 3000))
W = int(os.environ.get("W", 256))
args = argparse.Namespace(tick_grid=int(os.environ.get("GRID", 159)), tick_agents=int(os.environ.get("AGENTS", 1000000)),
                          tick_ticks=T0 + W, cpu_window=1800, no_configs23=False)
name = os.environ.get("CASE", "c4")
case = {c["key"]: c for c in bench.tick_cases(args)}[{"c4": "tick", "c2": "tick_c2", "c3": "tick_c3"}[name]]
flat = case["flat"]()
sim = bench.make_tick_sim(flat, case, torch.cuda.current_stream().cuda_stream)
if os.environ.get("DELTA") and getattr(sim, "_router_graph", None) is not None:
    sim._router_graph.set_tuning(delta_factor=float(os.environ["DELTA"]))      # bucket width of the router's trees
sim.run(0, T0)
sim.sync()
sim.enable_timeline(T0, W)
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ev0.record(); sim.run(T0, W); ev1.record(); sim.sync()
tl = sim.timeline().astype(np.int64)            # [W, 2, warps, 2]
used = tl[..., 1] > 0
big = np.iinfo(np.int64).max
start = np.where(used, tl[..., 0], big).min(axis=2)          # [W, 2]
end = np.where(used, tl[..., 1], 0).max(axis=2)
span = (end - start) / 1e3                                   # us
gap_ln = (start[:, 1] - end[:, 0]) / 1e3
gap_nl = (start[1:, 0] - end[:-1, 1]) / 1e3
ticks = np.arange(T0, T0 + W)
wf = int(case["router"][1]) if case["router"] else 10 ** 9
refresh = (ticks[1:] % wf == 0)
dur = (tl[..., 1] - tl[..., 0]) / 1e3
uniq = np.unique(tl[used])
res = int(np.diff(uniq).min()) if len(uniq) > 1 else 0


def pct(a, q):
    return round(float(np.percentile(a, q)), 2)


def at(phase, which, q):
    """us after the kernel's first warp entered until a fraction q of its warps had entered (0) / left (1)."""
    out = []
    for i in range(W):
        e = tl[i, phase, used[i, phase], which]
        out.append((np.percentile(e, q) - start[i, phase]) / 1e3)
    return round(float(np.mean(out)), 2)


def spans(a):
    return {"mean": round(float(a.mean()), 2), "p50": pct(a, 50), "p95": pct(a, 95)}


def warp_dur(ph):
    d, u = dur[:, ph], used[:, ph]
    return {"p50": pct(d[u], 50), "p99": pct(d[u], 99), "max_mean": round(float(np.where(u, d, 0).max(axis=1).mean()), 2)}


line = {
    "case": name, "t0": T0, "ticks": W, "us_per_tick_events": round(ev0.elapsed_time(ev1) * 1e3 / W, 2),
    "timer_resolution_ns": res,
    "link_span_us": spans(span[:, 0]), "node_span_us": spans(span[:, 1]),
    "gap_link_to_node_us": {"mean": round(float(gap_ln.mean()), 2), "p50": pct(gap_ln, 50)},
    "gap_node_to_link_us": {"mean_plain": round(float(gap_nl[~refresh].mean()), 2), "p50_plain": pct(gap_nl[~refresh], 50),
                            "mean_at_refresh": round(float(gap_nl[refresh].mean()), 2) if refresh.any() else None,
                            "refreshes": int(refresh.sum())},
    "warps": {"link": int(used[0, 0].sum()), "node": int(used[0, 1].sum())},
    "warp_duration_us": {"link": warp_dur(0), "node": warp_dur(1)},
    "all_warps_entered_us": {"link": at(0, 0, 100), "node": at(1, 0, 100)},
    "warps_done_us": {"link": {f"p{q}": at(0, 1, q) for q in (50, 90, 99, 100)},
                      "node": {f"p{q}": at(1, 1, q) for q in (50, 90, 99, 100)}},
}
print(json.dumps(line))
