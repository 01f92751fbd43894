"""Where does a node-phase warp spend its time?  Experiment build only (SQ_NVCC_EXTRA=-DSQ_TICK_PHASES python -m
spatial_queue_b200.build --force): checkpoints of %globaltimer inside k_node_phase, per warp, for W ticks from T0.
Prints the mean time between consecutive checkpoints for all warps and for the slowest 2 % of warps of each tick."""
import argparse, ctypes as C, json, os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import numpy as np
import torch
import bench
from spatial_queue_b200._lib import lib

T0, W = int(os.environ.get("T0", 7000)), int(os.environ.get("W", 64))
args = argparse.Namespace(tick_grid=159, tick_agents=1000000, tick_ticks=T0 + W, cpu_window=1800, no_configs23=True)
case = {c["key"]: c for c in bench.tick_cases(args)}["tick"]
flat = case["flat"]()
sim = bench.make_tick_sim(flat, case, torch.cuda.current_stream().cuda_stream)
sim.run(0, T0); sim.sync()
sim.enable_timeline(T0, W)
sim.run(T0, W); sim.sync()
tl = sim.timeline()
stride = tl.shape[2]
ph = np.zeros((W, stride, 16), np.uint64)
L = lib(); L.sq_get_phases.restype = C.c_int; L.sq_get_phases.argtypes = [C.c_void_p, C.c_void_p]
assert L.sq_get_phases(sim._h, ph.ctypes.data) == 0
ph = ph.astype(np.int64)
names = ["wait->cand", "cand->primary", "primary->win(ring issued)", "ring->agent issued", "agent->hop issued", "hop->tbl issued",
         "tbl->admit_set p", "admit_finish p", "admit_set s", "admit_finish s", "flush", "to end"]
have = ph[..., 12] > 0
dur = (ph[..., 12] - ph[..., 0]) / 1e3
out = {"t0": T0, "ticks": W, "warps": int(have[0].sum()), "warp_us": {"mean": round(float(dur[have].mean()), 2)}}
thr = np.array([np.percentile(dur[i][have[i]], 98) for i in range(W)])
slow = have & (dur >= thr[:, None])


def steps(mask):
    res = {}
    prev = ph[..., 0].copy()
    for k in range(1, 13):
        cur = ph[..., k]
        ok = mask & (cur > 0)
        d = np.where(ok, cur - prev, 0) / 1e3
        res[names[k - 1]] = [round(float(d[mask].mean()), 2), round(float(ok[mask].mean()), 2)]   # mean us over all warps of the class, share of warps that passed
        prev = np.where(cur > 0, cur, prev)
    return res


out["all_warps"] = steps(have)
out["slowest_2pct"] = steps(slow)
out["slowest_2pct_warp_us"] = round(float(dur[slow].mean()), 2)
print(json.dumps(out))
