"""Time of the device router's tree batch alone (run on the GPU box): the eight reverse trees of config 4's routing graph,
CUDA events around sp_sssp_batch_device_ex, warm.  One JSON line."""
import json, os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import numpy as np
import torch
from spatial_queue_b200 import synth

f = synth.flat_grid(int(os.environ.get("GRID", 159)), seed=1)
g = synth.routing_graph(f)
V = len(f.node_key)
rng = np.random.default_rng(3)
# weights like a loaded network: free-flow time plus a queue term on a third of the links
w = np.asarray(f.edge_fft, np.float64) * (1.0 + (rng.random(len(f.edge_fft)) < 0.33) * rng.random(len(f.edge_fft)) * 6.0)
g.set_weights(w)
n = int(os.environ.get("TREES", 8))
dests = torch.from_numpy(rng.integers(0, V, n).astype(np.int32)).cuda()
d = torch.empty((n, V), dtype=torch.float64, device="cuda")
p = torch.empty((n, V), dtype=torch.int32, device="cuda")
for _ in range(5):
    g.sssp_batch_device(dests, d, p, reverse=True)
torch.cuda.synchronize()
ts = []
for _ in range(20):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); g.sssp_batch_device(dests, d, p, reverse=True); b.record(); torch.cuda.synchronize()
    ts.append(a.elapsed_time(b) * 1e3)
g.sssp_batch(dests.cpu().numpy()[:n], want_dist=False, reverse=True)
print(json.dumps({"trees": n, "V": V, "us_median": round(float(np.median(ts)), 1), "us_min": round(float(min(ts)), 1), "stats": g.last_stats()}))
