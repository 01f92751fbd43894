"""GPU parity for HP-1: the CUDA tick, called through the C ABI, replays the reference's golden runs tick by tick
(bit-exact integer state), and matches the C oracle on capacities (f64, bit-exact) along the way."""
import numpy as np
import pytest

from tickcheck import load_golden, replay, route_events, state_digest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name,every", [("grid8", 1), ("grid7s", 1), ("bolinas", 1), ("grid5x", 5), ("grid6p", 1)])
def test_device_replays_reference_golden(name, every):
    from spatial_queue_b200.tick import DeviceSim
    z, flat = load_golden(name)
    sim = DeviceSim(flat, seed=int(z["seed"]))
    sim.set_shelter_after(3 * 3600)
    n = replay(sim, z, every=every)
    ag = sim.agent_state()
    assert (ag["arrive_tick"] == z["arrive_tick"]).all()
    c = sim.counters()
    assert c["ticks"] == n
    assert c["agent_steps"] == int(z["on_link"][:n].astype(np.int64).sum())
    assert c["moves"] == int(z["moves"][:n].astype(np.int64).sum())
    assert c["arrivals"] == int((z["arrive_tick"] >= 0).sum())
    sim.close()


@pytest.mark.parametrize("name", ["grid8", "grid7s"])
def test_device_matches_oracle_capacities_and_multi_tick_run(name):
    """sq_run (CUDA-graph replay of many ticks) == stepping; f64 capacities equal the oracle's bit for bit."""
    from oracle.tick import OracleSim
    from spatial_queue_b200.tick import DeviceSim
    z, flat = load_golden(name)
    seed = int(z["seed"])
    dev, orc = DeviceSim(flat, seed=seed), OracleSim(flat, seed=seed)
    events = route_events(z)
    ticks = sorted(t for t in events if t < int(z["n_ticks"]))
    t = 0
    for stop in ticks + [int(z["n_ticks"])]:
        if stop > t:
            dev.run(t, stop - t)
            orc.run(t, stop - t)
            t = stop
        sd, so = dev.state(), orc.state()
        assert state_digest(sd) == state_digest(so) == int(z["digests"][t - 1]), f"tick {t - 1}"
        for k in ("in_c", "out_c", "storage_c"):
            assert sd[k].tobytes() == so[k].tobytes(), (k, t)
        if stop in events:
            dev.set_routes(*events[stop])
            orc.set_routes(*events[stop])
    wd, wo = dev.link_weights(), orc.link_weights()
    assert np.array_equal(wd, wo, equal_nan=True)
    assert dev.counters()["agent_steps"] == orc.counters()["agent_steps"]


def test_device_reports_broken_routes():
    from spatial_queue_b200 import _lib
    from spatial_queue_b200.tick import DeviceSim
    z, flat = load_golden("grid8")
    # cut one agent's route short: it will queue at a node that is not its destination with no next link
    a = int(np.argmax(np.diff(flat.route_ptr) > 3))
    ptr = flat.route_ptr.copy()
    links = flat.route_links.copy()
    sim = DeviceSim(flat, seed=1)
    sim.set_routes([a], [0, 1], links[ptr[a]: ptr[a] + 1])
    with pytest.raises(_lib.B200LibraryError, match="no next link"):
        sim.run(0, 400)
        sim.sync()


@pytest.mark.parametrize("name", ["grid8", "grid7s", "bolinas", "grid5x", "grid6p"])
def test_device_router_replays_reference_golden(name):
    """The whole agent loop on the device: no routes and no route events from the golden file are used, only the
    network, the demand and the reference's parameters (reroute every 300 s, staggered by agent id; edge weights
    from queue lengths every 100 s).  sq_attach_router + sq_step must reproduce the reference run digest for
    digest -- i.e. the reverse-tree next-link tables pick the very paths the reference's per-agent Dijkstra did."""
    from spatial_queue_b200 import flatten as fl
    from spatial_queue_b200.tick import DeviceSim
    z, flat = load_golden(name)
    fl.router_arrays_flat(flat)
    g = fl.routing_graph_flat(flat)
    sim = DeviceSim(flat, seed=int(z["seed"]), routes_from_router=True)
    sim.attach_router(g, flat, reroute_freq=300, weight_freq=100)
    sim.set_shelter_after(3 * 3600)
    n = int(z["n_ticks"])
    for t in range(n):
        sim.step(t)
        if t % 7 == 0 or t < 50 or t == n - 1:
            assert state_digest(sim.state()) == int(z["digests"][t]), f"tick {t}"
    assert (sim.agent_state()["arrive_tick"] == z["arrive_tick"]).all()
    c = sim.counters()
    assert c["agent_steps"] == int(z["on_link"][:n].astype(np.int64).sum())
    sim.close()


def test_device_router_run_equals_steps():
    """sq_run (CUDA graphs between the routing refreshes) == sq_step tick by tick, router attached."""
    from spatial_queue_b200 import flatten as fl
    from spatial_queue_b200.tick import DeviceSim
    z, flat = load_golden("grid8")
    fl.router_arrays_flat(flat)
    sims = []
    for _ in range(2):
        sim = DeviceSim(flat, seed=3, routes_from_router=True)
        sim.attach_router(fl.routing_graph_flat(flat), flat, reroute_freq=60, weight_freq=100)
        sims.append(sim)
    a, b = sims
    for t in range(0, 437):
        a.step(t)
    b.run(0, 250)
    b.run(250, 187)
    sa, sb = a.state(), b.state()
    assert state_digest(sa) == state_digest(sb)
    assert a.counters()["agent_steps"] == b.counters()["agent_steps"]


def test_device_congestion_counter_matches_reference():
    """Row A6: link.congested on the device == the reference's, summed every 25 ticks and per link at the snapshots,
    over the 11.5k-tick run whose stuck vehicles wait half an hour and more."""
    from spatial_queue_b200.tick import DeviceSim
    from tickcheck import check_congestion_counter
    z, flat = load_golden("grid5x")
    sim = DeviceSim(flat, seed=int(z["seed"]))
    check_congestion_counter(sim, z)
    sim.close()


def test_config4_size_window_matches_oracle():
    """BASELINE config 4 at full network size (159 x 159 grid, 100,488 real + 25,281 source links) with 150k agents
    released over ten minutes: 900 ticks on the device (CUDA-graph replay) against the C oracle, bit-exact integer
    state and f64 capacities at five stops, equal run counters."""
    from oracle.tick import OracleSim
    from spatial_queue_b200 import synth
    from spatial_queue_b200.tick import DeviceSim
    flat = synth.grid_scenario(159, 150_000, horizon=600, seed=3)
    dev, orc = DeviceSim(flat, seed=9), OracleSim(flat, seed=9)
    for s in (dev, orc):
        s.set_shelter_after(3 * 3600)
        s.set_congest_after(120)        # low enough to trigger inside the window
    t = 0
    for stop in (1, 120, 400, 650, 900):
        dev.run(t, stop - t)
        orc.run(t, stop - t)
        t = stop
        sd, so = dev.state(), orc.state()
        assert state_digest(sd) == state_digest(so), f"tick {t - 1}"
        for k in ("in_c", "out_c", "storage_c"):
            assert sd[k].tobytes() == so[k].tobytes(), (k, t)
        assert (dev.link_congested() == orc.link_congested()).all()
    cd, co = dev.counters(), orc.counters()
    assert cd["agent_steps"] == co["agent_steps"] and cd["moves"] == co["moves"] and cd["arrivals"] == co["arrivals"]
    # row N4: the one-read-back summary equals the same reductions over the oracle's full state
    sm, ag, real = dev.summary(), orc.agent_state(), flat.link_is_virtual == 0
    for code, name in enumerate(("unloaded", "enroute", "arrive", "shelter_p", "shelter_a1")):
        assert sm[name] == int((ag["status"] == code).sum()), name
    assert sm["queued"] == int(so["q"].sum()) and sm["running"] == int(so["r"].sum())
    assert sm["congested"] == int(orc.link_congested().sum()) and sm["closed_links"] == 0
    assert sm["node_moves"] == int(so["node_move"].sum()) and sm["vehicles_left"] == int(so["left"][real].sum())
    assert sm["links_in_use"] == int(((so["q"] + so["r"])[real] > 0).sum())
    assert cd["moves"] > 500_000 and int(dev.link_congested().sum()) > 0
    dev.close(); orc.close()


def _town(device_routes=False):
    from spatial_queue_b200 import synth
    return synth.town_scenario(48, 40_000, horizon=300, seed=5, device_routes=device_routes)


def test_town_network_matches_oracle():
    """Irregular topology (stand-in for configs 2 / 3): streets removed, one-ways, diagonals, two hubs with more
    out-links than the node kernel caches, 1-4 lanes, signal and priority nodes, mixed vehicle lengths, 40k agents
    in five minutes (spill-back everywhere), links closed at tick 100 and half of them reopened at tick 250.
    CUDA tick against the C oracle: bit-exact state and capacities at every stop."""
    from oracle.tick import OracleSim
    from spatial_queue_b200.tick import DeviceSim
    flat = _town()
    N = len(flat.node_key)
    assert np.diff(flat.node_out_ptr).max() > 8 and (flat.node_type == 1).sum() > 20 and (flat.node_type == 2).sum() > 20
    assert (flat.link_reverse[N:] < 0).sum() > 100
    dev, orc = DeviceSim(flat, seed=11), OracleSim(flat, seed=11)
    for s in (dev, orc):
        s.set_shelter_after(400)
        s.set_congest_after(90)
    rng = np.random.default_rng(0)
    shut = (N + rng.choice(len(flat.edge_start), 60, replace=False)).astype(np.int32)
    t = 0
    for stop in (1, 50, 100, 180, 250, 420, 700):
        for s in (dev, orc):
            if t == 100:
                s.set_link_closed(shut, np.ones(len(shut), np.uint8))
            if t == 250:
                s.set_link_closed(shut[::2], np.zeros(len(shut[::2]), np.uint8))
            s.run(t, stop - t)
        t = stop
        sd, so = dev.state(), orc.state()
        assert state_digest(sd) == state_digest(so), f"tick {t - 1}"
        for k in ("in_c", "out_c", "storage_c"):
            assert sd[k].tobytes() == so[k].tobytes(), (k, t)
        assert (dev.link_congested() == orc.link_congested()).all()
    cd, co = dev.counters(), orc.counters()
    assert cd["agent_steps"] == co["agent_steps"] and cd["moves"] == co["moves"] and cd["arrivals"] == co["arrivals"]
    assert cd["moves"] > 100_000 and cd["arrivals"] > 1000
    assert (dev.agent_state()["status"] == 4).sum() > 0          # some gave up (shelter_a1)
    dev.close(); orc.close()


def test_town_device_router_on_free_flow_weights_equals_fixed_routes():
    """Rerouting every agent each 60 s on weights that never change must not change anything: the device router
    (reverse trees on the irregular network -> next-link tables -> staggered reroute blocks) against the C oracle
    walking the routes fixed at t=0."""
    from oracle.tick import OracleSim
    from spatial_queue_b200 import synth
    from spatial_queue_b200.tick import DeviceSim
    fixed, routed = _town(), _town(device_routes=True)
    orc = OracleSim(fixed, seed=4)
    dev = DeviceSim(routed, seed=4, routes_from_router=True)
    dev.attach_router(synth.routing_graph(routed), routed, reroute_freq=60, weight_freq=10 ** 6)
    t = 0
    for stop in (1, 61, 200, 500):
        dev.run(t, stop - t)
        orc.run(t, stop - t)
        t = stop
        assert state_digest(dev.state()) == state_digest(orc.state()), f"tick {t - 1}"
    dev.close(); orc.close()


def test_routing_graph_outlives_the_simulation_with_its_refreshed_weights():
    """The router leaves the newest edge weights on the device (no host round trip inside a refresh); the host's edge log
    catches up when somebody looks -- here after the simulation and its stream are gone: update_edge must start from the
    weights of the last refresh, not from the ones the graph was created with."""
    from oracle import sssp as osssp
    from spatial_queue_b200 import synth
    from spatial_queue_b200.tick import DeviceSim
    routed = _town(device_routes=True)
    g = synth.routing_graph(routed)
    dev = DeviceSim(routed, seed=4, routes_from_router=True)
    dev.attach_router(g, routed, reroute_freq=60, weight_freq=100)
    dev.run(0, 300); dev.sync()                       # refreshes at 0, 100, 200
    N = len(routed.node_key)
    w300 = np.asarray(dev.link_weights())[N:]         # what the refresh of tick 300 computes (state after tick 299)
    assert not np.isnan(w300).any()
    dev.run(300, 50); dev.sync()                      # refresh at 300, then ticks that change the queues again
    s, e = np.asarray(routed.edge_start), np.asarray(routed.edge_end)
    w_ref = np.asarray(routed.edge_fft, np.float64)
    assert (w300 != w_ref).any(), "the run never changed a weight: the test would prove nothing"
    origins = np.array([0, N // 3, N - 1], np.int32)
    dev.close()                                       # the stream the refreshes ran on is gone
    g.update_edge(int(s[0]), int(e[0]), 123.5)        # first host edit after the device-side refreshes
    d_new, _ = g.sssp_batch(origins)
    w_now = w300.copy()
    w_now[0] = 123.5
    d_orc = osssp.OracleGraph(s, e, w_now, max_vertex_id=N - 1).sssp_batch(origins)[0]
    assert d_new.tobytes() == d_orc.tobytes()
    g.free()


def test_stop_agents_matches_oracle():
    """sq_stop_agents (a driver rule shelters enroute agents between ticks: they leave their link's queue / run list
    at once, model/dta_meso_butte.py:256-263) against the oracle's list filtering: random agents anywhere in the
    rings (queue part, run part, source links, several on one link), every 40 ticks, on the irregular network."""
    from oracle.tick import OracleSim
    from spatial_queue_b200.tick import DeviceSim
    flat = _town()
    dev, orc = DeviceSim(flat, seed=2), OracleSim(flat, seed=2)
    for s in (dev, orc):
        s.set_congest_after(60)
    rng = np.random.default_rng(7)
    t, stopped = 0, 0
    for stop in range(40, 601, 40):
        dev.run(t, stop - t)
        orc.run(t, stop - t)
        t = stop
        ag = orc.agent_state()
        on = np.flatnonzero(ag["status"] == 1)
        pick = rng.choice(on, min(len(on), 400), replace=False).astype(np.int32)
        # several victims on the same links: everyone on the 3 fullest links as well
        ptr, order = orc.link_vehicles()
        full = np.argsort(np.diff(ptr))[-3:]
        pick = np.unique(np.concatenate([pick] + [order[ptr[l]:ptr[l + 1]] for l in full])).astype(np.int32)
        for s in (dev, orc):
            s.stop_agents(pick, 5, t)
        stopped += len(pick)
        sd, so = dev.state(), orc.state()
        assert state_digest(sd) == state_digest(so), f"after stops at {t}"
    dev.run(t, 100); orc.run(t, 100)
    sd, so = dev.state(), orc.state()
    assert state_digest(sd) == state_digest(so)
    for k in ("in_c", "out_c", "storage_c"):
        assert sd[k].tobytes() == so[k].tobytes(), k
    assert (dev.link_congested() == orc.link_congested()).all()
    cd, co = dev.counters(), orc.counters()
    assert cd["agent_steps"] == co["agent_steps"] and cd["moves"] == co["moves"]
    assert int((dev.agent_state()["status"] == 5).sum()) == stopped > 3000
    with pytest.raises(Exception):
        dev.stop_agents(pick[:1], 5, t)          # already stopped: not enroute
    dev.close(); orc.close()


@pytest.mark.parametrize("seed", range(24))
def test_random_small_towns_match_oracle(seed):
    """Differential fuzz: small irregular networks with random shape parameters (dead ends, isolated nodes, hubs,
    dense signals / priority nodes), short horizons (everything jams), random closures, reopenings and driver stops,
    a random give-up limit.  CUDA tick == C oracle at every stop, digest and capacities."""
    from oracle.tick import OracleSim
    from spatial_queue_b200 import synth
    from spatial_queue_b200.tick import DeviceSim
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.integers(6, 15))
    lon, lat = synth.grid_coords(n, seed)
    s, e, lanes, maxmph = synth.town_edges(n, seed, drop=float(rng.uniform(0.05, 0.4)), oneway=float(rng.uniform(0, 0.3)),
                                           diag=float(rng.uniform(0, 0.3)), hubs=int(rng.integers(0, 3)),
                                           hub_links=int(rng.integers(3, min(14, n * n - 2))))
    flat = synth.town_scenario(n, int(rng.integers(300, 4000)), horizon=int(rng.integers(20, 200)), seed=seed,
                               n_exits=int(rng.integers(1, 6)), signals=float(rng.uniform(0, 0.3)),
                               priority=float(rng.uniform(0, 0.3)))
    N, E = len(flat.node_key), len(flat.edge_start)
    dev, orc = DeviceSim(flat, seed=seed), OracleSim(flat, seed=seed)
    give_up = int(rng.integers(50, 400))
    for sim in (dev, orc):
        sim.set_shelter_after(give_up)
        sim.set_congest_after(int(give_up // 3))
    t = 0
    for _ in range(8):
        stop = t + int(rng.integers(1, 90))
        dev.run(t, stop - t)
        orc.run(t, stop - t)
        t = stop
        sd, so = dev.state(), orc.state()
        assert state_digest(sd) == state_digest(so), f"seed {seed} tick {t - 1}"
        for k in ("in_c", "out_c", "storage_c"):
            assert sd[k].tobytes() == so[k].tobytes(), (seed, k, t)
        assert (dev.link_congested() == orc.link_congested()).all()
        links = (N + rng.choice(E, min(E, 5), replace=False)).astype(np.int32)
        flags = rng.integers(0, 2, len(links)).astype(np.uint8)
        on = np.flatnonzero(so["status"] == 1)
        pick = rng.choice(on, min(len(on), int(rng.integers(0, 30))), replace=False).astype(np.int32)
        for sim in (dev, orc):
            sim.set_link_closed(links, flags)
            if len(pick):
                sim.stop_agents(pick, 5, t)
    cd, co = dev.counters(), orc.counters()
    assert cd["agent_steps"] == co["agent_steps"] and cd["moves"] == co["moves"] and cd["arrivals"] == co["arrivals"]
    dev.close(); orc.close()


def test_link_pair_flow_counters_match_oracle():
    """Row A6: run_node_model's second return value, (in_link, out_link) per moved vehicle (model/queue_class.py:
    377-383), as per-pair counters on the device; equal to the oracle's pair list after 600 ticks on the irregular
    network, and consistent with total_vehicles_left (row sum + arrivals)."""
    from oracle.tick import OracleSim
    from spatial_queue_b200.tick import DeviceSim
    flat = _town()
    dev, orc = DeviceSim(flat, seed=4), OracleSim(flat, seed=4)
    dev.enable_pair_moves()
    dev.run(0, 600)
    orc.run(0, 600)
    di, do_, dc = dev.pair_moves(flat)
    oi, oo, oc = orc.pair_moves(flat)
    assert len(dc) > 2000 and (di == oi).all() and (do_ == oo).all() and (dc == oc).all()
    cd = dev.counters()
    assert int(dc.sum()) == cd["moves"] - cd["arrivals"]
    left = dev.link_state()["left"].astype(np.int64)
    per_in = np.bincount(di, weights=dc, minlength=len(left)).astype(np.int64)
    assert (per_in <= left).all() and int((left - per_in).sum()) == cd["arrivals"]
    dev.close(); orc.close()
