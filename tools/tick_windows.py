"""How long are the node model's windows?  (run on the GPU box)  Runs config 4 to tick T and prints, per class of
in-link, the distribution of k = min(whole lanes, queue length) -- the vehicles one node thread walks in a tick."""
import argparse
import json
import os
import sys

R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R)
import numpy as np
import torch

import bench

args = argparse.Namespace(tick_grid=159, tick_agents=1000000, tick_ticks=14400, cpu_window=1800, no_configs23=True)
case = {c["key"]: c for c in bench.tick_cases(args)}["tick"]
flat = case["flat"]()
sim = bench.make_tick_sim(flat, case, torch.cuda.current_stream().cuda_stream)
prev = 0
for T in [int(x) for x in os.environ.get("TS", "3000,7000,10000").split(",")]:
    sim.run(prev, T - prev); sim.sync(); prev = T
    st = sim.link_state()
    lanes = np.asarray(flat.link_lanes_floor)
    virt = np.asarray(flat.link_is_virtual).astype(bool)
    k = np.minimum(lanes, st["q"])
    endn = np.asarray(flat.link_end_node)
    per_node = np.zeros(len(flat.node_key), np.int64)
    np.maximum.at(per_node, endn, k)
    out = {"t": T, "links_with_queue": int((st["q"] > 0).sum()), "virtual_with_queue": int(((st["q"] > 0) & virt).sum()),
           "queue_total": int(st["q"].sum()), "queue_on_virtual": int(st["q"][virt].sum()),
           "k_hist_virtual": np.bincount(np.minimum(k[virt], 101))[:102].tolist()[:12] + ["..."] +
                             [int((k[virt] > 11).sum()), int((k[virt] > 32).sum()), int((k[virt] >= 100).sum())],
           "k_hist_real": np.bincount(k[~virt])[:8].tolist(),
           "nodes_max_k": {f">{x}": int((per_node > x).sum()) for x in (0, 4, 12, 20, 36, 68, 99)},
           "blocked_real_links_storage_lt_8": int(((st["storage_c"] < 8.0) & ~virt).sum())}
    print(json.dumps(out), flush=True)
