"""Tuning sweep for the SSSP tree kernel on the C5 grid (run on the GPU box)."""
import sys, os, time, json, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import torch
from conftest import grid_graph
from spatial_queue_b200 import sp_interface as spi

n = int(os.environ.get("GRID", 224))
B = int(os.environ.get("BATCH", 4096))
s, e, w = grid_graph(n, seed=1)
TILE = int(os.environ.get("TILE", 0))     # experiment: renumber the grid into TILE x TILE tiles before building the graph
if TILE:
    ii, jj = np.divmod(np.arange(n * n), n)
    nt = (n + TILE - 1) // TILE
    key = ((ii // TILE) * nt + (jj // TILE)) * (TILE * TILE) + (ii % TILE) * TILE + (jj % TILE)
    perm = np.argsort(np.argsort(key, kind="stable"), kind="stable").astype(np.int32)
    s, e = perm[s], perm[e]
g = spi.from_arrays(s, e, w)
V = n * n
origins = torch.from_numpy(np.random.default_rng(0).permutation(V)[:B].astype(np.int32)).cuda()
dist = torch.empty((B, V), dtype=torch.float64, device="cuda")
par = torch.empty((B, V), dtype=torch.int32, device="cuda")
btree = 20 * len(s) + 24 * V
ALGO = int(os.environ.get("ALGO", 2))
g.set_algo(ALGO)
configs = json.loads(os.environ.get("CONFIGS", "null")) or (
    [(t, c, d, 0) for t in (32, 64, 128, 256) for c in (4, 8, 12, 16, 24, 32) for d in (1.0, 2.0, 4.0) if t * c <= 2048]
    if ALGO == 0 else
    [(i, c, d, q) for i in (1, 2, 4) for q in (1024, 2048) for c in (3, 4, 5, 6) for d in (1.0, 2.0, 4.0)]
    if ALGO == 2 else
    [(t, c, d, q) for t in (128, 256, 512) for c in (0, 2, 4) for d in (1.0, 2.0, 4.0, 8.0) for q in (1024,)])
for cfg in configs:
    t, c, d, q = cfg[:4]
    try:
        if ALGO == 2:   # (ilp, CTAs of four warps per SM, delta factor, ring entries)
            g.set_warp_tuning(ring_cap=q, ilp=t, ctas_per_sm=c)
            g.set_tuning(delta_factor=d)
        else:
            g.set_tuning(threads=t, ctas_per_sm=c, delta_factor=d, smem_queue=q)
        g.sssp_batch_device(origins[:256], dist, par); torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ms = 1e30
        for _ in range(3):
            ev0.record(); g.sssp_batch_device(origins, dist, par); ev1.record(); torch.cuda.synchronize()
            ms = min(ms, ev0.elapsed_time(ev1))
        tps = B / ms * 1e3
        ev0.record(); g.sssp_batch_device(origins, dist, None); ev1.record(); torch.cuda.synchronize()
        ms_tree = ev0.elapsed_time(ev1)
        st = g.last_stats()
        g.sssp_batch(np.random.default_rng(0).permutation(V)[:64].astype(np.int32), want_dist=False)
        st = g.last_stats()
        print(json.dumps({"algo": ALGO, "threads": t, "ctas": c, "delta": d, "q": q, "ms": round(ms, 2), "ms_trees_only": round(ms_tree, 2), "trees_per_s": round(tps),
                          "frac_hbm": round(tps * btree / 6547.2e9, 4), "relax_per_tree": st["relaxations"] / 64,
                          "rounds_per_tree": st["rounds"] / 64, "overflows": st["overflows"]}), flush=True)
    except Exception as ex:
        print("ERR", t, c, d, q, ex, flush=True)
