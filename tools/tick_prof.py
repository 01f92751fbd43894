"""Config-4 tick run for ncu (run on the GPU box): device router attached, TICKS ticks from tick 0.
GRAPH=1 replays CUDA graphs (sq_run); otherwise one sq_step per tick (every launch visible to ncu)."""
import sys, os, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import numpy as np, torch
from spatial_queue_b200 import synth
from spatial_queue_b200.tick import DeviceSim
n = int(os.environ.get("GRID", 159)); A = int(os.environ.get("AGENTS", 1000000)); T = int(os.environ.get("TICKS", 3000))
flat = synth.grid_scenario(n, A, horizon=7200, seed=1, device_routes=True)
sim = DeviceSim(flat, seed=7, routes_from_router=True)
sim.set_stream(torch.cuda.current_stream().cuda_stream)
sim.attach_router(synth.routing_graph(flat), flat, reroute_freq=60, weight_freq=100)
sim.set_shelter_after(3 * 3600)
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
