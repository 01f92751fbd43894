"""Config-4 tick: ms per tick by segment of the run (run on the GPU box)."""
import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import torch
from spatial_queue_b200 import synth
from spatial_queue_b200.tick import DeviceSim
flat = synth.grid_scenario(159, 1000000, horizon=7200, seed=1, device_routes=True)
sim = DeviceSim(flat, seed=7, routes_from_router=True)
sim.set_stream(torch.cuda.current_stream().cuda_stream)
sim.attach_router(synth.routing_graph(flat), flat, reroute_freq=60, weight_freq=100)
sim.set_shelter_after(3 * 3600)
seg = int(os.environ.get("SEG", 1200))
prev = sim.counters()
for t0 in range(0, 14400, seg):
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(); sim.run(t0, seg); ev1.record(); sim.sync()
    c = sim.counters()
    print(f"ticks {t0:5d}-{t0 + seg:5d}: {ev0.elapsed_time(ev1) / seg * 1e3:6.1f} us/tick  on-link {(c['agent_steps'] - prev['agent_steps']) / seg:9.0f}  "
          f"moves/tick {(c['moves'] - prev['moves']) / seg:7.0f}  active nodes/tick {(c['active_nodes'] - prev['active_nodes']) / seg:7.0f}")
    prev = c
