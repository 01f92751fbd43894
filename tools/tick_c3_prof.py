"""Where the time of the config-3 run goes (run on the GPU box): the full run in 300-tick segments, then the same run
with one ingredient removed at a time (rules, closures, weight refresh)."""
import json
import os
import sys

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch

import bench
from spatial_queue_b200 import synth
from spatial_queue_b200.tick import DeviceSim

flat = synth.butte_standin(device_routes=True)
ev = synth.butte_closure_schedule(flat)
T = int(os.environ.get("TICKS", 18001))


def make(rules=True, closures=True, wfreq=100, rfreq=60):
    sim = DeviceSim(flat, seed=7, routes_from_router=True)
    sim.set_stream(torch.cuda.current_stream().cuda_stream)
    sim.attach_router(synth.routing_graph(flat), flat, reroute_freq=rfreq, weight_freq=wfreq)
    if rules:
        sim.set_reroute_rules(2.0, 10)
    if closures:
        sim.set_closure_schedule([e[0] for e in ev], [e[1] for e in ev], [e[2] for e in ev])
    sim.set_shelter_after(9000)
    return sim


def run(sim, seg=300):
    out = []
    t = 0
    while t < T:
        n = min(seg, T - t)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); sim.run(t, n); b.record(); sim.sync()
        out.append(a.elapsed_time(b) / n * 1e3)
        t += n
    return out


w = make(); w.run(0, 256); w.sync(); w.close()
for name, kw in (("full", {}), ("no rules", {"rules": False}), ("no closures", {"closures": False}),
                 ("no weight refresh", {"wfreq": 10 ** 6}), ("neither rules nor closures", {"rules": False, "closures": False})):
    sim = make(**kw)
    us = run(sim)
    c = sim.counters()
    print(json.dumps({"case": name, "us_per_tick": round(sum(us) / len(us), 2), "segments_us_per_tick": [round(x, 1) for x in us[::4]],
                      "agent_steps": c["agent_steps"], "summary": sim.summary()}), flush=True)
    sim.close()
