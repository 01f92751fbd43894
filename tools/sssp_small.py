import sys, os, json
R = "/root/repo"
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
from conftest import grid_graph
from spatial_queue_b200 import sp_interface as spi
n = int(os.environ.get("GRID", 159)); B = int(os.environ.get("BATCH", 8))
s, e, w = grid_graph(n, seed=1)
g = spi.from_arrays(s, e, w); V = n * n
if os.environ.get("DELTA"):
    g.set_tuning(delta_factor=float(os.environ["DELTA"]))
origins = np.random.default_rng(0).permutation(V)[:B].astype(np.int32)
o = torch.from_numpy(origins).cuda()
dist = torch.empty((B, V), dtype=torch.float64, device="cuda"); par = torch.empty((B, V), dtype=torch.int32, device="cuda")
for _ in range(3):
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(); g.sssp_batch_device(o, dist, par); ev1.record(); torch.cuda.synchronize()
    print("ms", ev0.elapsed_time(ev1))
dd, pp = g.sssp_batch(origins)
st = g.last_stats(); print(st, "rounds/tree", st["rounds"] / B, "trans/tree", st["transitions"] / B)
