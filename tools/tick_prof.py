"""A tick run for ncu (run on the GPU box): CASE = c4 (default) / c2 / c3 of bench.tick_cases, TICKS ticks from tick 0.
GRAPH=1 replays CUDA graphs (sq_run); otherwise one sq_step per tick (every launch visible to ncu)."""
import argparse
import os
import sys
import time

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch

import bench

T = int(os.environ.get("TICKS", 3000))
args = argparse.Namespace(tick_grid=int(os.environ.get("GRID", 159)), tick_agents=int(os.environ.get("AGENTS", 1000000)),
                          tick_ticks=T, cpu_window=1800, no_configs23=False)
case = {c["key"]: c for c in bench.tick_cases(args)}[{"c4": "tick", "c2": "tick_c2", "c3": "tick_c3"}[os.environ.get("CASE", "c4")]]
flat = case["flat"]()
sim = bench.make_tick_sim(flat, case, torch.cuda.current_stream().cuda_stream)
if os.environ.get("GRAPH", "0") == "1":
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(); sim.run(0, T); ev1.record(); sim.sync()
    print("graph run ms/tick", ev0.elapsed_time(ev1) / T)
else:
    t0 = time.time()
    for t in range(T):
        sim.step(t)
    sim.sync()
    print("stepped ms/tick", (time.time() - t0) * 1e3 / T)
print(sim.counters())
