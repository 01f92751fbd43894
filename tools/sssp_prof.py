"""One launch of the SSSP tree kernel on the C5 grid, for ncu (run on the GPU box)."""
import sys, os, json
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np, torch
from conftest import grid_graph
from spatial_queue_b200 import sp_interface as spi
n = int(os.environ.get("GRID", 224)); B = int(os.environ.get("BATCH", 2048))
ALGO = int(os.environ.get("ALGO", 2))
t, c, d, q = json.loads(os.environ.get("CFG", "[4,0,2.0,1024]" if ALGO == 2 else "[256,6,2.0,0]"))
s, e, w = grid_graph(n, seed=1)
g = spi.from_arrays(s, e, w); V = n * n
g.set_algo(ALGO)
if ALGO == 4:   # group path: GMODE 0 = warp kernel on the group layout, 1 = sssp_group_kernel, 2 = sssp_union_kernel
    g.set_group_mode(int(os.environ.get("GMODE", 2)))
    g.set_group_tuning(warps_per_sm=int(os.environ.get("GWARPS", 16)), batch=int(os.environ.get("GBATCH", 8)))
    g.set_tuning(delta_factor=d)
elif ALGO == 2:   # CFG = [ilp, CTAs of four warps per SM, delta factor, ring entries]
    g.set_warp_tuning(ring_cap=q, ilp=t, ctas_per_sm=c)
    g.set_tuning(delta_factor=d)
else:
    g.set_tuning(threads=t, ctas_per_sm=c, delta_factor=d, smem_queue=q)
origins = np.random.default_rng(0).permutation(V)[:B].astype(np.int32)
o = torch.from_numpy(origins).cuda()
dist = torch.empty((B, V), dtype=torch.float64, device="cuda"); par = torch.empty((B, V), dtype=torch.int32, device="cuda")
for _ in range(2):
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(); g.sssp_batch_device(o, dist, par); ev1.record(); torch.cuda.synchronize()
    print("ms", ev0.elapsed_time(ev1), "trees/s", B / ev0.elapsed_time(ev1) * 1e3)
dd, pp = g.sssp_batch(origins[:64])
st = g.last_stats(); print(st, "relax/tree", st["relaxations"] / 64, "rounds/tree", st["rounds"] / 64, "trans/tree", st["transitions"] / 64, "E", len(s))
