"""The object mirror (spatial_queue_b200/queue_class.py) against the reference's own objects.

CPU part: the mirror's loader + add_connectivity + add_demand, flattened, must equal the flattening of the
REFERENCE's objects stored in the golden files (tests/golden/tick_*.npz `flat_*`, written by make_tick_golden.py
from the unmodified reference loader on the same CSV inputs).  GPU part: `use_gpu()` + `step(t)` + device-side
batched rerouting replays the reference's whole run, digest for digest.
"""
import json
import os

import numpy as np
import pandas as pd
import pytest

from tickcheck import load_golden, state_digest

from spatial_queue_b200 import flatten as fl, queue_class as qc, synth

REF_BOLINAS = "/root/reference/projects/bolinas_civic"
EMPTY = {"signal": {}, "priority": {}, "prohibition": {}, "closure": []}

STATIC_FIELDS = ["node_key", "link_start_node", "link_end_node", "link_is_virtual", "link_lanes_floor", "link_cap_per_s",
                 "link_storage_cap", "link_fft", "link_closed", "link_reverse", "link_capacity", "link_maxmph",
                 "node_type", "node_in_ptr", "node_in_links", "node_in_phase", "node_in_prio", "node_out_ptr",
                 "node_out_links", "node_turn_ptr", "turn_angle", "pair_start", "pair_end", "pair_link",
                 "agent_key", "agent_origin", "agent_destin", "agent_dept", "agent_vlen"]


def grid7s_special(n=7):
    edges, _ = synth.grid_frames(n, seed=1, arterial_every=4)

    def ins(node):
        return [int(x) for x in edges.loc[edges.nid_e == node, "eid"]]
    sig = {}
    for node in (8, 24, 40):
        il = ins(node)
        sig[str(node)] = {"phase_1": il[0::2], "phase_2": il[1::2]}
    pri = {str(node): list(reversed(ins(node))) for node in (16, 32)}
    return {"signal": sig, "priority": pri, "prohibition": {}, "closure": [5, 30]}


def grid6p_special(n=6):
    """tests/golden/make_tick_golden.py:grid6p_special: turn prohibitions at two interior nodes."""
    edges, _ = synth.grid_frames(n, seed=1, arterial_every=4)
    proh = {}
    for node, skip in ((14, (0, 5, 10)), (21, (3, 6))):
        ins = [int(x) for x in edges.loc[edges.nid_e == node, "eid"]]
        outs = [int(x) for x in edges.loc[edges.nid_s == node, "eid"]]
        pairs = [[a, b] for a in ins for b in outs]
        proh[str(node)] = [pairs[k] for k in skip]
    return {"signal": {}, "priority": {}, "prohibition": proh, "closure": []}


SCENARIOS = {
    "grid6p": dict(n=6, agents=900, horizon=60, exits=[0, 35, 5], special=grid6p_special()),
    "grid8": dict(n=8, agents=1500, horizon=60, exits=[0, 63], special=EMPTY),
    "grid7s": dict(n=7, agents=1200, horizon=80, exits=[0, 48, 6], special=grid7s_special()),
}


def build_grid(name, tmp_path):
    """Same inputs and call sequence as tests/golden/make_tick_golden.py:grid, through the mirror."""
    sc = SCENARIOS[name]
    edges, nodes = synth.grid_frames(sc["n"], seed=1, arterial_every=4)
    od = synth.grid_od(sc["n"], sc["agents"], seed=2, horizon=sc["horizon"], exits=sc["exits"])
    wp = str(tmp_path)
    edges.to_csv(wp + "/edges.csv", index=False)
    nodes.to_csv(wp + "/nodes.csv", index=False)
    od.to_csv(wp + "/od.csv", index=False)
    np.random.seed(0)
    net = qc.Network()
    net.dataframe_to_network(project_location="", network_file_edges=wp + "/edges.csv",
                             network_file_nodes=wp + "/nodes.csv", special_nodes=sc["special"], scen_nm=name)
    net.add_connectivity()
    net.add_demand(demand_files=["/od.csv"], write_path=wp)
    return net, sc["special"]


def assert_static_equal(flat, gold):
    for k in STATIC_FIELDS:
        a, b = np.asarray(getattr(flat, k)), np.asarray(getattr(gold, k))
        assert a.shape == b.shape, k
        assert a.tobytes() == b.astype(a.dtype).tobytes(), f"{k} differs from the reference's objects"


@pytest.mark.parametrize("name", ["grid8", "grid7s", "grid6p"])
def test_mirror_loader_matches_reference_objects(name, tmp_path):
    """grid6p: turn prohibitions (model/queue_class.py:69-120) -- split nodes, trimmed link geometry (turn angles),
    zero-length connector links, the routing graph over the new ids; agents that start at or head for a node the split
    left without links shelter at t = 0 like in the reference."""
    z, gold = load_golden(name)
    net, special = build_grid(name, tmp_path)
    if name == "grid6p":
        # the reference's first one_step routes everyone: unreachable pairs become shelter_p (queue_class.py:494-499)
        dests = sorted({a.destin_nid for a in net.agents.values()})
        assert len(net.links) == len(gold.link_end_node) and len(net.nodes) == 2 * 52
    flat = fl.flatten_network(net, special)
    if name == "grid6p":
        keep = [k for k in STATIC_FIELDS if not k.startswith("agent_")]
        for k in keep:
            a, b = np.asarray(getattr(flat, k)), np.asarray(getattr(gold, k))
            assert a.shape == b.shape and a.tobytes() == b.astype(a.dtype).tobytes(), f"{k} differs from the reference's objects"
        return
    assert_static_equal(flat, gold)
    # string ids of the virtual elements (queue_class.py:210, 213)
    assert net.links["n3_vl"].start_nid == "vn3" and net.links["n3_vl"].end_nid == 3
    assert net.node2link_dict[("vn3", 3)] == "n3_vl"


@pytest.mark.skipif(not os.path.isdir(REF_BOLINAS), reason="Bolinas inputs live in /root/reference (build container only)")
def test_mirror_loader_bolinas_matches_reference_objects(tmp_path):
    """The shipped Bolinas network through the mirror == through the reference's loader (golden flat arrays)."""
    import pandas as pd
    z, gold = load_golden("bolinas")
    rng = np.random.default_rng(1)
    res = pd.read_csv(REF_BOLINAS + "/network_inputs/osm_nodes.csv")
    vis = pd.read_csv(REF_BOLINAS + "/network_inputs/osm_nodes_visitor_parking.csv")
    ok = set(pd.read_csv(REF_BOLINAS + "/network_inputs/bolinas_nodes_sim.csv").osmid)
    r = rng.choice([o for o in res.node_osmid if o in ok and o != 110397253], 900)
    v = rng.choice([o for o in vis.node_osmid if o in ok and o != 110397253], 300)
    od = pd.DataFrame({"origin_osmid": np.concatenate([r, v]), "destin_osmid": 110397253, "dept_time": 0,
                       "type": ["local"] * 900 + ["visitor"] * 300})
    od.to_csv(str(tmp_path) + "/od.csv", index=False)
    special = json.load(open(REF_BOLINAS + "/network_inputs/bolinas_special_nodes.json"))
    np.random.seed(0)
    net = qc.Network()
    old, qc.abs_path = qc.abs_path, REF_BOLINAS
    try:
        net.dataframe_to_network(project_location="", network_file_edges="/network_inputs/bolinas_edges_sim.csv",
                                 network_file_nodes="/network_inputs/bolinas_nodes_sim.csv", special_nodes=special)
    finally:
        qc.abs_path = old
    net.add_connectivity()
    net.add_demand(demand_files=["/od.csv"], write_path=str(tmp_path))
    # run_traffic_simulation.py:86-88 overwrites the departure times at t == 0 from numpy's global stream
    # (the golden generator reseeds right before that first one_step call)
    np.random.seed(0)
    for a in net.agents.values():
        a.departure_time = np.random.randint(1, 100)
    flat = fl.flatten_network(net, special)
    assert_static_equal(flat, gold)


def test_host_tick_methods_refuse_to_run(tmp_path):
    """No CPU tick: the per-object methods of the reference API fail loudly instead of computing on the host."""
    from spatial_queue_b200._lib import B200LibraryError
    net, _ = build_grid("grid8", tmp_path)
    link = next(iter(net.links.values()))
    node = net.nodes[0]
    agent = next(iter(net.agents.values()))
    for call in (lambda: link.run_link_model(0, agent_id_dict=net.agents),
                 lambda: node.run_node_model(0, node_id_dict=net.nodes, link_id_dict=net.links),
                 lambda: agent.load_vehicle(0, node2link_dict=net.node2link_dict, link_id_dict=net.links),
                 lambda: net.step(0)):
        with pytest.raises(B200LibraryError):
            call()


@pytest.mark.gpu
@pytest.mark.parametrize("route_mode", ["host", "device"])
@pytest.mark.parametrize("name", ["grid8", "grid7s"])
def test_mirror_on_gpu_replays_reference_run(name, route_mode, tmp_path):
    """One call per tick (`network.step`) with the reroutes computed by the GPU library reproduces the
    reference's run: initial routes, every later route change, and the per-tick state digests.  In "host" mode the
    routes are forward trees walked into `agent.route` dicts; in "device" mode nothing but `sq_step` runs per tick
    (reverse trees per destination, next-link tables and the reroute trigger all on the GPU)."""
    z, gold = load_golden(name)
    net, special = build_grid(name, tmp_path)
    net.use_gpu(seed=int(z["seed"]), special_nodes=special, reroute_freq=300, weight_freq=100, route_mode=route_mode)
    if route_mode == "host":
        assert net._flat.route_ptr.tobytes() == gold.route_ptr.tobytes()
        assert net._flat.route_links.tobytes() == gold.route_links.tobytes()
    n = int(z["n_ticks"])
    for t in range(n):
        net.step(t)
        assert state_digest(net.sim.state()) == int(z["digests"][t]), f"tick {t}"
    net.sync_to_host()
    arrived = {aid: v[1] - 1 for aid, v in net.agents_stopped.items()}
    gold_arr = {int(gold.agent_key[i]): int(t) for i, t in enumerate(z["arrive_tick"]) if t >= 0}
    assert arrived == gold_arr
    assert sum(len(l.queue_vehicles) + len(l.run_vehicles) for l in net.links.values()) == int(z["on_link"][n - 1])
    assert sum(nd.node_move for nd in net.nodes.values() if nd.node_type == "real") == int(z["moves"][:n].sum())


@pytest.mark.gpu
def test_closures_host_and_device_routing_agree(tmp_path):
    """Fire-style closures (model/dta_meso_butte.py:88-115): `link.close_link_to_newcomers(g)` + status 'closed' on
    both approaches of an exit makes it unreachable, so the agents heading there keep their old routes
    (queue_class.py:501-503) and pile up; 900 s later the links reopen.  The two routing modes of the mirror --
    per-agent route dicts recomputed on the host vs next-link tables on the device, including the hand-over of
    routes that outlive their table epoch -- must produce the same run, tick for tick."""
    runs = []
    for mode in ("host", "device"):
        net, special = build_grid("grid8", tmp_path)
        net.use_gpu(seed=11, special_nodes=special, reroute_freq=300, weight_freq=100, route_mode=mode)
        runs.append(net)
    closing = [lid for lid, l in runs[0].links.items() if l.link_type != "v" and l.end_nid == 0]
    assert len(closing) == 2
    digests = [[], []]
    for k, net in enumerate(runs):
        for t in range(1400):
            if t == 120:
                for lid in closing:
                    net.links[lid].close_link_to_newcomers(g=net.g)
                    net.links[lid].status = "closed"
            if t == 1020:
                for lid in closing:
                    net.links[lid].open_link_to_newcomers(g=net.g)
                    net.links[lid].status = "burnt_over"
            net.step(t)
            if t % 10 == 0 or t in (119, 120, 121, 1019, 1020, 1021):
                digests[k].append(state_digest(net.sim.state()))
    assert digests[0] == digests[1]
    a, b = runs[0].sim.agent_state(), runs[1].sim.agent_state()
    assert (a["arrive_tick"] == b["arrive_tick"]).all()
    # the closure really bit: once the vehicles already on the two links are through, arrivals at node 0 pause
    arr = a["arrive_tick"][runs[0]._flat.agent_destin == runs[0]._flat.node_index[0]]
    assert ((arr > 400) & (arr < 1020)).sum() == 0 and (arr >= 1020).sum() > 0


@pytest.mark.gpu
def test_closure_schedule_equals_per_tick_calls(tmp_path):
    """Row N3: the same closures given once as a device event table (Network.schedule_closures ->
    sq_set_closure_schedule) and replayed with sq_run in long stretches == the driver closing and reopening the
    links by hand before the ticks in question (the run of test_closures_host_and_device_routing_agree)."""
    nets = []
    for _ in range(2):
        net, special = build_grid("grid8", tmp_path)
        net.use_gpu(seed=11, special_nodes=special, reroute_freq=300, weight_freq=100, route_mode="device")
        nets.append(net)
    manual, sched = nets
    closing = [lid for lid, l in manual.links.items() if l.link_type != "v" and l.end_nid == 0]
    sched.schedule_closures([(120, lid, True) for lid in closing] + [(1020, lid, False) for lid in closing]
                            + [(700, closing[0], True), (700, closing[0], True)])        # duplicates are harmless
    stops = [119, 120, 121, 500, 1019, 1020, 1021, 1399]
    dm, ds = [], []
    t = 0
    for stop in stops:
        while t <= stop:
            if t == 120:
                for lid in closing:
                    manual.links[lid].close_link_to_newcomers(g=manual.g)
                    manual.links[lid].status = "closed"
            if t == 1020:
                for lid in closing:
                    manual.links[lid].open_link_to_newcomers(g=manual.g)
                    manual.links[lid].status = "burnt_over"
            manual.step(t)
            t += 1
        dm.append(state_digest(manual.sim.state()))
    t = 0
    for stop in stops:
        sched.run(t, stop + 1 - t)
        t = stop + 1
        ds.append(state_digest(sched.sim.state()))
    assert dm == ds
    assert (manual.sim.agent_state()["arrive_tick"] == sched.sim.agent_state()["arrive_tick"]).all()


@pytest.mark.gpu
def test_get_partial_path_matches_the_reference_class():
    """Agent.get_partial_path of the mirror (look-ahead re-plan + splice, model/queue_class.py:514-575) on the GPU
    `sp` library against tests/golden/partial_path.json, which the unmodified reference class produced on its own
    sp library: returned tuple, route dict (content and order), route distance and status, exactly."""
    import types
    from ctypes import c_double
    from spatial_queue_b200 import sp_interface
    z = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "partial_path.json")))
    n = z["n"]
    lon, lat = synth.grid_coords(n, seed=z["seed"])
    s, e = synth.grid_edges(n)
    length, _, maxmph = synth.grid_attributes(n, s, e, lon, lat, z["arterial_every"])
    w = length / maxmph * 2.237
    node2link = {(int(a), int(b)): i for i, (a, b) in enumerate(zip(s, e))}
    for k, c in enumerate(z["cases"]):
        g = sp_interface.from_arrays(s, e, w, max_vertex_id=n * n - 1)
        links = {i: types.SimpleNamespace(travel_time=float(w[i])) for i in range(len(s))}
        a = qc.Agent(k, c["origin"], c["destin"], 0, 8, 1, None)
        a.get_path(0, g=g)
        assert [[int(x), int(y)] for x, y in a.route.items()] == c["route_before"], k
        cl = c["closed"]
        g.update_edge(int(s[cl]), int(e[cl]), c_double(1e8))
        links[cl].travel_time = 1e8
        ret = a.get_partial_path(5, g=g, link_id_dict=links, node2link_dict=node2link)
        assert (None if ret is None else [float(x) for x in ret]) == c["ret"], k
        assert [[int(x), int(y)] for x, y in a.route.items()] == c["route_after"], k
        assert float(a.route_distance) == c["route_distance"] and a.status == c["status"], k
        g.free()


@pytest.mark.gpu
def test_detour_rule_shelters_agents_on_the_device(tmp_path):
    """Butte's shelter_a2 rule in host route mode (model/dta_meso_butte.py:199-204, 256-263): an enroute agent whose
    new route is more than `detour_factor` times its previous one shelters in place -- status, stop tick, off its
    link's lists at once (sq_stop_agents), into agents_stopped.  An exit is closed so that the reroutes lengthen."""
    net, special = build_grid("grid8", tmp_path)
    net.use_gpu(seed=5, special_nodes=special, reroute_freq=60, weight_freq=50, route_mode="host", detour_factor=1.25)
    closing = [lid for lid, l in net.links.items() if l.link_type != "v" and l.end_nid == 0]
    n0 = len(net.agents)
    for t in range(500):
        if t == 100:
            for lid in closing:
                net.links[lid].close_link_to_newcomers(g=net.g)
                net.links[lid].status = "closed"
        net.step(t)
    net.sync_to_host()
    a2 = {k: v for k, v in net.agents_stopped.items() if v[0] == "shelter_a2"}
    assert len(a2) > 5 and all(60 < v[1] < 500 for v in a2.values())
    on_links = {v for l in net.links.values() for v in l.queue_vehicles + l.run_vehicles}
    assert not (set(a2) & on_links) and not (set(a2) & set(net.agents))
    st = net.sim.agent_state()["status"]
    assert int((st == 5).sum()) == len(a2)
    assert len(net.agents) + len(net.agents_stopped) == n0


@pytest.mark.gpu
def test_reroute_upon_closure_rule(tmp_path):
    """Butte's reroute-upon-closure rule in host route mode (model/dta_meso_butte.py:205-209): an agent whose next
    link is closed is rerouted when t % 10 == agent_id % 10, so twelve ticks after a closure the only agents still
    heading into it are those that reached the link before it less than ten ticks ago (there is a way round);
    without the rule the regular 300 s stagger leaves several times as many waiting at the closed link."""
    waiting = {}
    for freq in (10, None):
        net, special = build_grid("grid8", tmp_path)
        net.use_gpu(seed=5, special_nodes=special, reroute_freq=300, weight_freq=100, route_mode="host",
                    closure_reroute_freq=freq)
        for t in range(60):
            net.step(t)
        # close the street most agents are about to turn into (not one that ends at an exit: there must be a way round)
        st = net.sim.agent_state()
        exits = set(int(x) for x in net._flat.agent_destin)
        nxt = st["nxt"][(st["status"] <= 1) & (st["nxt"] >= 0)]
        cnt = np.bincount(nxt, minlength=len(net._link_objs))
        cnt[[i for i, l in enumerate(net._link_objs) if l.link_type == "v" or net._flat.node_index[l.end_nid] in exits]] = 0
        busiest = int(cnt.argmax())
        assert cnt[busiest] > 5
        link = net._link_objs[busiest]
        link.close_link_to_newcomers(g=net.g)
        link.status = "closed"
        for t in range(60, 72):
            net.step(t)
        st = net.sim.agent_state()
        head_in = (st["status"] <= 1) & (st["nxt"] == busiest)
        waiting[freq] = int(head_in.sum())
        if freq:
            # whoever still heads into it only just got there: moved onto the link before it within the last ten ticks
            assert ((st["status"][head_in] == 1) & (st["enter"][head_in] > 71 - 10)).all()
    assert waiting[10] * 3 < waiting[None], waiting


def build_town(seed, tmp_path, n=9, agents=900):
    """Irregular network through the mirror's CSV loader; origins anywhere (some cannot reach their exit: shelter_p at
    t=0, queue_class.py:494-499), origins only at nodes that have a way out (the reference crashes otherwise)."""
    edges, nodes = synth.town_frames(n, seed, hubs=1, hub_links=6)
    rng = np.random.default_rng(seed)
    has_out = np.unique(edges.nid_s.values)
    exits = rng.choice(np.unique(edges.nid_e.values), 3, replace=False)
    origin = rng.choice(has_out, agents)
    destin = exits[rng.integers(0, 3, agents)]
    clash = origin == destin
    origin[clash] = has_out[(np.searchsorted(has_out, origin[clash]) + 1) % len(has_out)]
    origin[origin == destin] = has_out[0] if has_out[0] not in exits else has_out[1]
    od = pd.DataFrame({"origin_osmid": 10_000_000 + origin, "destin_osmid": 10_000_000 + destin,
                       "dept_time": 1 + (np.arange(agents) * 50) // agents, "type": "local"})
    wp = str(tmp_path)
    edges.to_csv(wp + "/edges.csv", index=False)
    nodes.to_csv(wp + "/nodes.csv", index=False)
    od.to_csv(wp + "/od.csv", index=False)
    np.random.seed(seed)
    net = qc.Network()
    net.dataframe_to_network(project_location="", network_file_edges=wp + "/edges.csv",
                             network_file_nodes=wp + "/nodes.csv", special_nodes=EMPTY, scen_nm="town")
    net.add_connectivity()
    net.add_demand(demand_files=["/od.csv"], write_path=wp)
    return net


@pytest.mark.gpu
@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_host_and_device_routing_agree_on_irregular_networks(seed, tmp_path):
    """The two routing modes of the mirror on irregular networks (one-ways, dead ends, unreachable exits -> shelter_p
    at t=0): per-agent forward paths from `sp_route_batch` on the host side vs reverse-tree next-link tables on the
    device, rerouting every 40 s on weights refreshed every 25 s, one street closed for a while."""
    runs = []
    for mode in ("host", "device"):
        (tmp_path / mode).mkdir()
        net = build_town(seed, tmp_path / mode)
        net.use_gpu(seed=seed, special_nodes=EMPTY, reroute_freq=40, weight_freq=25, route_mode=mode, shelter_after=200)
        runs.append(net)
    digests = [[], []]
    for k, net in enumerate(runs):
        closing = sorted(lid for lid, l in net.links.items() if l.link_type != "v")[5::40]
        for t in range(330):
            if t in (60, 200):
                for lid in closing:
                    if t == 60:
                        net.links[lid].close_link_to_newcomers(g=net.g)
                        net.links[lid].status = "closed"
                    else:
                        net.links[lid].open_link_to_newcomers(g=net.g)
                        net.links[lid].status = "open"
            net.step(t)
            if t % 10 == 0:
                digests[k].append(state_digest(net.sim.state()))
    bad = [i for i, (a, b) in enumerate(zip(*digests)) if a != b]
    assert not bad, f"first difference at tick {bad[0] * 10}"
    assert (runs[0].sim.agent_state()["arrive_tick"] == runs[1].sim.agent_state()["arrive_tick"]).all()
    assert (runs[0].sim.agent_state()["status"] == 2).sum() > 100


@pytest.mark.gpu
def test_butte_rules_inside_the_device_router_equal_host_mode(tmp_path):
    """Rows N1 / N2: Butte's shelter_a2 detour rule and reroute-upon-closure (model/dta_meso_butte.py:193-209, 256-263)
    inside the device router (sq_set_reroute_rules: per-agent route distance, closure stagger slot, the sheltering
    vehicle stripped from the middle of its ring by the link phase of the same tick) against the same rules in host
    route mode (Python over sp_route_batch + sq_stop_agents), tick for tick, on four irregular networks with streets
    closed at t = 60 and reopened at t = 200.  A low detour factor makes the rule fire."""
    total_a2 = 0
    for seed in (1, 2, 3, 4):
        runs = []
        for mode in ("host", "device"):
            d = tmp_path / f"{mode}{seed}"
            d.mkdir()
            net = build_town(seed, d)
            net.use_gpu(seed=seed, special_nodes=EMPTY, reroute_freq=40, weight_freq=25, route_mode=mode, shelter_after=200,
                        detour_factor=1.2, closure_reroute_freq=10)
            runs.append(net)
        digests = [[], []]
        for k, net in enumerate(runs):
            closing = sorted(lid for lid, l in net.links.items() if l.link_type != "v")[5::12]
            for t in range(330):
                if t in (60, 200):
                    for lid in closing:
                        if t == 60:
                            net.links[lid].close_link_to_newcomers(g=net.g)
                            net.links[lid].status = "closed"
                        else:
                            net.links[lid].open_link_to_newcomers(g=net.g)
                            net.links[lid].status = "open"
                net.step(t)
                if t % 5 == 0:
                    digests[k].append(state_digest(net.sim.state()))
        bad = [i for i, (a, b) in enumerate(zip(*digests)) if a != b]
        assert not bad, f"seed {seed}: first difference at tick {bad[0] * 5}"
        sa, sb = runs[0].sim.agent_state(), runs[1].sim.agent_state()
        assert (sa["status"] == sb["status"]).all() and (sa["arrive_tick"] == sb["arrive_tick"]).all()
        sh0, sh1 = runs[0].sim.shelter_state(), runs[1].sim.shelter_state()
        assert (sh0["stop_tick"] == sh1["stop_tick"]).all() and (sh0["link_shelter"] == sh1["link_shelter"]).all()
        total_a2 += int((sb["status"] == 5).sum())
        # the device run made no per-tick host call beyond step(): its mirror objects only move at sync_to_host
        runs[1].sync_to_host()
        a2 = {k for k, v in runs[1].agents_stopped.items() if v[0] == "shelter_a2"}
        assert len(a2) == int((sb["status"] == 5).sum())
    assert total_a2 > 10
