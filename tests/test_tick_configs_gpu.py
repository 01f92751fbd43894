"""BASELINE configs 2 and 3 at their named sizes (missing item 1): the stand-in networks of SURVEY.md 8(d)
(the reference ships neither the Berkeley nor the Butte inputs), full agent counts, device against independent host paths."""
import numpy as np
import pytest

from tickcheck import state_digest

pytestmark = pytest.mark.gpu


def test_config2_berkeley_standin_full_run_matches_oracle():
    """Config 2 (benchmark/profile_berkeley.py:144: ticks 0..1400, no rerouting): ~38k links on ~15k nodes, signal and
    priority nodes, 50,000 agents on routes fixed at t = 0.  CUDA tick (sq_run) against the C oracle over the whole
    1,401 ticks: integer state digest and f64 capacities at seven stops, counters and arrival ticks at the end."""
    from oracle.tick import OracleSim
    from spatial_queue_b200 import synth
    from spatial_queue_b200.tick import DeviceSim
    flat = synth.berkeley_standin()
    N = len(flat.node_key)
    assert 37_000 < len(flat.link_end_node) - N < 40_000 and 14_000 < N < 16_000 and len(flat.agent_key) == 50_000
    dev, orc = DeviceSim(flat, seed=2), OracleSim(flat, seed=2)
    for s in (dev, orc):
        s.set_shelter_after(3 * 3600)
        s.set_congest_after(300)
    t = 0
    for stop in (1, 100, 300, 600, 900, 1200, 1401):
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
    assert (dev.agent_state()["arrive_tick"] == orc.agent_state()["arrive_tick"]).all()
    assert cd["moves"] > 1_000_000 and cd["arrivals"] > 5_000
    dev.close(); orc.close()


def test_config3_butte_standin_with_rules_and_closures_matches_host_path():
    """Config 3 (model/dta_meso_butte.py): ~31k links, 100,000 evacuees, rerouting every 60 s on weights refreshed every
    100 s, the shelter_a2 detour rule (factor 2), reroute upon closure (stagger 10) and the closure schedule of
    :56-115 (fire closures every 300 s, p = 0.15, reopening after 3,600 s; timed arterial closures) as a device event
    table -- everything inside sq_run, no host call per tick.  Against the C oracle driven by the independent host
    path of tests/hostrouter.py (forward Dijkstra per due agent, the same rules in Python), 1,000 ticks: digest and
    capacities every 100 ticks, statuses and stop ticks at the end."""
    from hostrouter import ForwardRouter
    from oracle.tick import OracleSim
    from spatial_queue_b200 import synth
    from spatial_queue_b200.tick import DeviceSim
    REROUTE, WEIGHT, TICKS = 60, 100, 1000
    flat = synth.butte_standin(device_routes=True)
    N = len(flat.node_key)
    assert 30_000 < len(flat.link_end_node) - N < 32_500 and len(flat.agent_key) == 100_000
    ev = synth.butte_closure_schedule(flat)
    in_window = [e for e in ev if e[0] < TICKS]
    assert len(in_window) > 40
    dev = DeviceSim(flat, seed=3, routes_from_router=True)
    dev.attach_router(synth.routing_graph(flat), flat, reroute_freq=REROUTE, weight_freq=WEIGHT)
    dev.set_reroute_rules(2.0, 10)
    dev.set_closure_schedule([e[0] for e in ev], [e[1] for e in ev], [e[2] for e in ev])
    orc = OracleSim(flat, seed=3)
    for s in (dev, orc):
        s.set_shelter_after(9000)
    fr = ForwardRouter(flat, REROUTE, WEIGHT, detour_factor=2.0, closure_reroute_freq=10, closures=ev)
    try:
        for t0 in range(0, TICKS, 100):
            dev.run(t0, 100)
            for t in range(t0, t0 + 100):
                fr.tick(orc, t)
                orc.step(t)
            sd, so = dev.state(), orc.state()
            assert state_digest(sd) == state_digest(so), f"tick {t0 + 99}"
            for k in ("in_c", "out_c", "storage_c"):
                assert sd[k].tobytes() == so[k].tobytes(), (k, t0 + 100)
    finally:
        fr.close()
    ad, ao = dev.agent_state(), orc.agent_state()
    assert (ad["status"] == ao["status"]).all() and (ad["arrive_tick"] == ao["arrive_tick"]).all()
    shd, sho = dev.shelter_state(), orc.shelter_state()
    assert (shd["stop_tick"] == sho["stop_tick"]).all() and (shd["link_shelter"] == sho["link_shelter"]).all()
    cd, co = dev.counters(), orc.counters()
    assert cd["agent_steps"] == co["agent_steps"] and cd["moves"] == co["moves"] and cd["arrivals"] == co["arrivals"]
    sm = dev.summary()
    assert sm["closed_links"] > 20 and cd["moves"] > 200_000
    print("config 3 window: shelter_a2", int((ad["status"] == 5).sum()), "host path stopped", fr.stopped_a2,
          "closed links", sm["closed_links"], "trees", fr.trees_computed)
    dev.close(); orc.close()
