"""Times the UNMODIFIED Python reference tick (build container only: /root/reference must exist).

VERDICT r1 item 8: the thing north_star names as the baseline is the Python reference itself, so it is timed here
on one core, through oracle/ref_harness.py (stubs + RNG shim):
  * Bolinas as shipped (861 links, 1,200 agents, reroute every 300 s), run to completion;
  * the 50 x 50 synthetic grid of SURVEY.md section 6 (12,300 links, 20,000 agents departing over 600 s), ticks 1..600.
Prints one JSON line per case; BASELINE.md section 3 and bench.py's tick cpu_baseline quote the result."""
import json
import os
import platform
import sys
import time

import numpy as np
import pandas as pd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_harness as rh  # noqa: E402
from spatial_queue_b200 import synth  # noqa: E402


def run(name, network, special_nodes, rts, max_ticks, dept_override=None):
    data, config = {"network": network}, {"scen_nm": name, "simulation_outputs": "/simulation_outputs", "special_nodes": special_nodes}
    np.random.seed(0)
    t0 = time.perf_counter()
    rts.one_step(0, data, config, {})
    t_route0 = time.perf_counter() - t0
    if dept_override is not None:
        for aid, a in network.agents.items():
            a.departure_time = int(dept_override[aid])
    resident = on_link = 0
    t0 = time.perf_counter()
    last = 0
    for t in range(1, max_ticks):
        resident += len(network.agents)
        _, status = rts.one_step(t, data, config, {})
        on_link += sum(len(l.queue_vehicles) + len(l.run_vehicles) for l in network.links.values())
        last = t
        if status == "stop":
            break
    wall = time.perf_counter() - t0
    print(json.dumps({"case": name, "links": len(network.links), "ticks": last, "t0_routing_s": round(t_route0, 2),
                      "wall_s": round(wall, 2), "s_per_tick": round(wall / last, 5),
                      "resident_agent_steps_per_s": round(resident / wall), "on_link_agent_steps_per_s": round(on_link / wall),
                      "cores": 1, "machine": platform.processor() or platform.machine(), "cpu_count": os.cpu_count()}), flush=True)


def main():
    qc, rts = rh.import_reference(seed=12345)
    which = sys.argv[1:] or ["bolinas", "grid50"]
    if "bolinas" in which:
        wp = rh.make_write_path()
        rng = np.random.default_rng(1)
        res = pd.read_csv(rh.REF_PROJECT + "/network_inputs/osm_nodes.csv")
        vis = pd.read_csv(rh.REF_PROJECT + "/network_inputs/osm_nodes_visitor_parking.csv")
        ok = set(pd.read_csv(rh.REF_PROJECT + "/network_inputs/bolinas_nodes_sim.csv").osmid)
        r = rng.choice([o for o in res.node_osmid if o in ok and o != 110397253], 900)
        v = rng.choice([o for o in vis.node_osmid if o in ok and o != 110397253], 300)
        od = pd.DataFrame({"origin_osmid": np.concatenate([r, v]), "destin_osmid": 110397253, "dept_time": 0,
                           "type": ["local"] * 900 + ["visitor"] * 300})
        od.to_csv(wp + "/demand_inputs/od_csv/resident_visitor_od_vphh1.5_visitor300.csv", index=False)
        np.random.seed(0)
        data, config, _ = rts.preparation(vphh=1.5, visitor_cnts=300, scen_nm="timing", write_path=wp)
        run("bolinas (861 links, 1200 agents, to completion)", data["network"], config["special_nodes"], rts, 5400)
    if "grid50" in which:
        wp = rh.make_write_path()
        n, agents = 50, 20000
        edges, nodes = synth.grid_frames(n, seed=1)
        edges.to_csv(wp + "/edges.csv", index=False)
        nodes.to_csv(wp + "/nodes.csv", index=False)
        synth.grid_od(n, agents, seed=2, horizon=600).to_csv(wp + "/od.csv", index=False)
        special = {"signal": {}, "priority": {}, "prohibition": {}, "closure": []}
        old, qc.abs_path = qc.abs_path, ""
        np.random.seed(0)
        net = qc.Network()
        net.dataframe_to_network(project_location="", network_file_edges=wp + "/edges.csv", network_file_nodes=wp + "/nodes.csv",
                                 special_nodes=special, scen_nm="grid50", write_path=wp)
        net.add_connectivity()
        net.add_demand(demand_files=["/od.csv"], write_path=wp)
        qc.abs_path = old
        dept = {aid: int(a.departure_time) for aid, a in net.agents.items()}
        run("50x50 grid (12300 links, 20000 agents over 600 s), ticks 1..600", net, special, rts, 601, dept_override=dept)


if __name__ == "__main__":
    main()
