"""The benchmarked tick workload, pinned at size (parity hole 1b): config-4 network, 150k agents, the device
router at bench.py's cadences (every agent rerouted each 60 s, weights refreshed every 100 s) against an independent
host path -- the C tick oracle driven by routes from a FORWARD Dijkstra per due agent on the oracle's own weights
(tests/hostrouter.py: the reference's algorithm, queue_class.py:491-512), which shares nothing with the device
router's reverse trees per destination."""
import numpy as np
import pytest

from tickcheck import state_digest


def _oracle_router(starts, ends, w, roots):
    from oracle import sssp as osssp
    og = osssp.OracleGraph(starts, ends, w, max_vertex_id=int(max(starts.max(), ends.max())))
    return og.sssp_batch(np.asarray(roots, np.int32), want_dist=False)[1]


def test_forward_router_equals_reverse_tables_on_free_flow_weights():
    """CPU check of the host path itself: at t = 0 (free-flow weights) the forward-Dijkstra routes of every agent are the
    routes synth.grid_scenario derives from the oracle's reverse trees (the grid's jittered weights leave no ties)."""
    from hostrouter import ForwardRouter
    from spatial_queue_b200 import synth
    flat = synth.grid_scenario(20, 3000, horizon=300, seed=3, router=_oracle_router)
    fr = ForwardRouter(flat, 60, 100, workers=2)
    try:
        idx = np.arange(len(flat.agent_key))
        ptr, links = fr.routes(flat.agent_origin, flat.agent_destin)
    finally:
        fr.close()
    assert (ptr == flat.route_ptr).all() and (links == flat.route_links).all()
    assert fr.trees_computed == len(np.unique(flat.agent_origin)) and len(idx) == 3000


@pytest.mark.gpu
def test_config4_device_router_matches_forward_dijkstra_host_path():
    """159 x 159 grid (100,488 real + 25,281 source links), 150,000 agents released over ten minutes (heavy spill-back),
    1,200 ticks.  Device: sq_attach_router(reroute 60 s, weights 100 s), sq_run.  Host: OracleSim + ForwardRouter.
    Integer state digest and f64 capacities must agree at every weight-refresh boundary and at the end; any
    disagreement between reverse-tree and forward-tree next links would show up here as a digest mismatch."""
    from hostrouter import ForwardRouter
    from oracle.tick import OracleSim
    from spatial_queue_b200 import synth
    from spatial_queue_b200.tick import DeviceSim
    REROUTE, WEIGHT, TICKS = 60, 100, 1200
    flat = synth.grid_scenario(159, 150_000, horizon=600, seed=3, device_routes=True)
    dev = DeviceSim(flat, seed=9, routes_from_router=True)
    dev.attach_router(synth.routing_graph(flat), flat, reroute_freq=REROUTE, weight_freq=WEIGHT)
    orc = OracleSim(flat, seed=9)
    for s in (dev, orc):
        s.set_shelter_after(3 * 3600)
    fr = ForwardRouter(flat, REROUTE, WEIGHT)
    rerouted = 0
    try:
        for t0 in range(0, TICKS, WEIGHT):
            dev.run(t0, WEIGHT)
            for t in range(t0, t0 + WEIGHT):
                rerouted += fr.tick(orc, t)
                orc.step(t)
            sd, so = dev.state(), orc.state()
            assert state_digest(sd) == state_digest(so), f"tick {t0 + WEIGHT - 1}"
            for k in ("in_c", "out_c", "storage_c"):
                assert sd[k].tobytes() == so[k].tobytes(), (k, t0 + WEIGHT)
            assert np.array_equal(dev.link_weights(), orc.link_weights(), equal_nan=True)
    finally:
        fr.close()
    cd, co = dev.counters(), orc.counters()
    assert cd["agent_steps"] == co["agent_steps"] and cd["moves"] == co["moves"] and cd["arrivals"] == co["arrivals"]
    assert (dev.agent_state()["arrive_tick"] == orc.agent_state()["arrive_tick"]).all()
    # the case is what it claims to be: everyone rerouted about 20 times, queues long enough to change the weights
    assert rerouted > 150_000 * (TICKS // REROUTE) * 0.8 and fr.trees_computed > 100_000
    # (free-flow weights are 9 - 24 s on this grid; queued links weigh more than twice that)
    assert cd["moves"] > 1_000_000 and float(np.nanmax(orc.link_weights())) > 50.0
    dev.close(); orc.close()
