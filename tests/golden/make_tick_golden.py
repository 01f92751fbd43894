"""Generate the HP-1 golden vectors by running the UNMODIFIED reference tick (build container only).

For each scenario the reference's own Network loader, `one_step` driver (run_traffic_simulation.py:71-139) and
queue_class tick run on the stubs of oracle/refstubs with the counter-based RNG shim (oracle/ref_harness.py).
Written per scenario to tests/golden/tick_<name>.npz:
  flat_*      the flattened network + agents + t=0 routes (spatial_queue_b200.flatten applied to the reference objects)
  ev_*        every later route change (tick, agent, new link sequence) so the GPU side replays identical routes
  digests     one 64-bit digest per tick of (queue/run lengths, vehicles left, vehicle order on every link,
              agent status / link / next link / enter time, node_move)
  snap<T>_*   full state arrays at selected ticks; arrive_tick per agent; per-tick on-link and move counts
Run:  python tests/golden/make_tick_golden.py [scenario ...]
"""
import json
import os
import sys
import time

import numpy as np
import pandas as pd

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_harness as rh  # noqa: E402
from spatial_queue_b200 import flatten as fl, synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
SEED = 12345


def run(name, network, special_nodes, qc, rts, max_ticks, dept_override=None, snapshot_every=100):
    data, config = {"network": network}, {"scen_nm": name, "simulation_outputs": "/simulation_outputs",
                                          "special_nodes": special_nodes}
    np.random.seed(0)
    rts.one_step(0, data, config, {})
    if dept_override is not None:
        for aid, a in network.agents.items():
            a.departure_time = int(dept_override[aid])
    flat = fl.flatten_network(network, special_nodes)
    snaps = set(range(0, max_ticks, snapshot_every)) | set(range(1, 40))
    rec = rh.Recorder(network, flat, snaps)
    rlog = rh.RouteLog(qc, network, flat)
    rec.record(0)
    t0 = time.time()
    last = 0
    for t in range(1, max_ticks):
        _, status = rts.one_step(t, data, config, {})
        rec.record(t)
        last = t
        if status == "stop":
            break
    rlog.close()
    out = {"flat_" + k: v for k, v in flat.as_dict().items()}
    out.update(rec.to_npz_dict())
    out.update(rlog.to_npz_dict())
    out["seed"] = np.uint64(SEED)
    out["n_ticks"] = np.int32(last + 1)
    np.savez_compressed(os.path.join(OUT, f"tick_{name}.npz"), **out)
    arrived = int((rec.arrive_tick >= 0).sum())
    print(f"{name}: links {len(flat.link_ids)} agents {len(flat.agent_ids)} ticks {last + 1} arrived {arrived} "
          f"route events {len(rlog.events)} max on-link {max(rec.on_link)} ref wall {time.time() - t0:.1f}s "
          f"agent-steps {sum(rec.on_link)}")


def bolinas(qc, rts):
    wp = rh.make_write_path()
    rng = np.random.default_rng(1)
    res = pd.read_csv(rh.REF_PROJECT + "/network_inputs/osm_nodes.csv")
    vis = pd.read_csv(rh.REF_PROJECT + "/network_inputs/osm_nodes_visitor_parking.csv")
    nodes_sim = pd.read_csv(rh.REF_PROJECT + "/network_inputs/bolinas_nodes_sim.csv")
    ok = set(nodes_sim.osmid)
    r = rng.choice([o for o in res.node_osmid if o in ok and o != 110397253], 900)
    v = rng.choice([o for o in vis.node_osmid if o in ok and o != 110397253], 300)
    od = pd.DataFrame({"origin_osmid": np.concatenate([r, v]), "destin_osmid": 110397253, "dept_time": 0,
                       "type": ["local"] * 900 + ["visitor"] * 300})
    od.to_csv(wp + "/demand_inputs/od_csv/resident_visitor_od_vphh1.5_visitor300.csv", index=False)
    np.random.seed(0)
    data, config, _ = rts.preparation(vphh=1.5, visitor_cnts=300, scen_nm="golden", write_path=wp)
    run("bolinas", data["network"], config["special_nodes"], qc, rts, 5400)


def grid(qc, rts, name, n, n_agents, special=None, horizon=100, max_ticks=3000, arterial_every=4, exits=None,
         snapshot_every=50):
    wp = rh.make_write_path()
    edges, nodes = synth.grid_frames(n, seed=1, arterial_every=arterial_every)
    edges.to_csv(wp + "/edges.csv", index=False)
    nodes.to_csv(wp + "/nodes.csv", index=False)
    od = synth.grid_od(n, n_agents, seed=2, horizon=horizon, exits=exits)
    od.to_csv(wp + "/od.csv", index=False)
    special_nodes = {"signal": {}, "priority": {}, "prohibition": {}, "closure": []}
    special_nodes.update(special or {})
    old_abs = qc.abs_path
    qc.abs_path = ""
    np.random.seed(0)
    net = qc.Network()
    net.dataframe_to_network(project_location="", network_file_edges=wp + "/edges.csv", network_file_nodes=wp + "/nodes.csv",
                             special_nodes=special_nodes, scen_nm=name, write_path=wp)
    net.add_connectivity()
    net.add_demand(demand_files=["/od.csv"], write_path=wp)
    qc.abs_path = old_abs
    dept = {aid: int(a.departure_time) for aid, a in net.agents.items()}
    run(name, net, special_nodes, qc, rts, max_ticks, dept_override=dept, snapshot_every=snapshot_every)


def main():
    qc, rts = rh.import_reference(seed=SEED)
    which = sys.argv[1:] or ["bolinas", "grid8", "grid7s", "grid5x", "grid6p"]
    if "bolinas" in which:
        bolinas(qc, rts)
    if "grid8" in which:
        grid(qc, rts, "grid8", 8, 1500, horizon=60, exits=[0, 63])
    if "grid7s" in which:
        n = 7
        edges, _ = synth.grid_frames(n, seed=1, arterial_every=4)
        def ins(node):
            return [int(x) for x in edges.loc[edges.nid_e == node, "eid"]]
        sig = {}
        for node in (8, 24, 40):
            il = ins(node)
            sig[str(node)] = {"phase_1": il[0::2], "phase_2": il[1::2]}
        pri = {str(node): list(reversed(ins(node))) for node in (16, 32)}
        grid(qc, rts, "grid7s", n, 1200, special={"signal": sig, "priority": pri, "closure": [5, 30]}, horizon=80,
             exits=[0, 48, 6])


    if "grid6p" in which:
        grid(qc, rts, "grid6p", 6, 900, special=grid6p_special(), horizon=60, exits=[0, 35, 5], max_ticks=2500)
    if "grid5x" in which:
        # a signal whose phases never serve one approach: the vehicles queued there wait for ever and give up after
        # three hours (status shelter_a1, run_traffic_simulation.py:100-104, 113-118), as do the ones stuck behind
        n = 5
        edges, _ = synth.grid_frames(n, seed=1, arterial_every=4)
        served = [int(x) for x in edges.loc[(edges.nid_e == 1) & (edges.nid_s != 2), "eid"]]
        sig = {"1": {"phase_1": ["n1_vl"] + served[0:1], "phase_2": ["n1_vl"] + served[1:2]}}   # 2 -> 1 is never served
        grid(qc, rts, "grid5x", n, 500, special={"signal": sig}, horizon=100, exits=[0, 24], max_ticks=11500,
             snapshot_every=900)


def grid6p_special(n=6):
    """Turn prohibitions (queue_class.py:76-131) at two interior nodes of the 6 x 6 grid: the loader splits each node
    into one node per link end and joins them by zero-length connector links, leaving the forbidden pairs out."""
    edges, _ = synth.grid_frames(n, seed=1, arterial_every=4)
    proh = {}
    for node, skip in ((14, (0, 5, 10)), (21, (3, 6))):
        ins = [int(x) for x in edges.loc[edges.nid_e == node, "eid"]]
        outs = [int(x) for x in edges.loc[edges.nid_s == node, "eid"]]
        pairs = [[a, b] for a in ins for b in outs]
        proh[str(node)] = [pairs[k] for k in skip]
    return {"signal": {}, "priority": {}, "prohibition": proh, "closure": []}


if __name__ == "__main__":
    main()
