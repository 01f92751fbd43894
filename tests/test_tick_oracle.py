"""CPU: the C tick oracle replays every golden scenario produced by the unmodified reference, tick by tick."""
import numpy as np
import pytest

from oracle.tick import OracleSim
from tickcheck import load_golden, replay


@pytest.mark.parametrize("name,every", [("grid8", 1), ("grid7s", 1), ("bolinas", 1), ("grid5x", 5), ("grid6p", 1)])
def test_oracle_replays_reference_golden(name, every):
    """grid5x runs for 11,500 ticks behind a signal approach that is never served: the vehicles stuck there give
    up after three hours on the link (status shelter_a1, run_traffic_simulation.py:100-104)."""
    z, flat = load_golden(name)
    sim = OracleSim(flat, seed=int(z["seed"]))
    sim.set_shelter_after(3 * 3600)
    n = replay(sim, z, every=every)
    ag = sim.agent_state()
    assert (ag["arrive_tick"] == z["arrive_tick"]).all()
    c = sim.counters()
    assert c["agent_steps"] == int(z["on_link"][:n].astype(np.int64).sum())
    assert c["moves"] == int(z["moves"][:n].astype(np.int64).sum())


def test_oracle_congestion_counter_matches_reference():
    from oracle.tick import OracleSim
    from tickcheck import check_congestion_counter
    z, flat = load_golden("grid5x")
    sim = OracleSim(flat, seed=int(z["seed"]))
    check_congestion_counter(sim, z)
    sim.close()


def test_oracle_on_irregular_network_conserves_vehicles():
    """Host logic of the irregular-network generator (synth.flat_from_edges / town_scenario, routed here by the C
    SSSP oracle) + the oracle's stop_agents: CSR consistency of the flat network, and vehicle conservation through
    closures, give-ups and driver-ordered stops."""
    from oracle import sssp as osssp
    from spatial_queue_b200 import synth

    def router(starts, ends, w, roots):
        og = osssp.OracleGraph(starts, ends, w, max_vertex_id=int(max(starts.max(), ends.max())))
        return og.sssp_batch(np.asarray(roots, np.int32), want_dist=False)[1]

    flat = synth.town_scenario(24, 6000, horizon=200, seed=5, router=router)
    N, L = len(flat.node_key), len(flat.link_start_node)
    assert (flat.link_end_node[flat.node_in_links] == np.repeat(np.arange(N), np.diff(flat.node_in_ptr))).all()
    assert (flat.link_start_node[flat.node_out_links] == np.repeat(np.arange(N), np.diff(flat.node_out_ptr))).all()
    rev = flat.link_reverse
    has = rev >= 0
    assert (flat.link_start_node[rev[has]] == flat.link_end_node[has]).all() and (flat.link_end_node[rev[has]] == flat.link_start_node[has]).all()
    assert (~has[N:]).sum() > 10 and np.diff(flat.node_out_ptr).max() > 8
    sim = OracleSim(flat, seed=1)
    sim.set_shelter_after(150)
    rng = np.random.default_rng(0)
    t, stopped = 0, 0
    for stop in (50, 100, 150, 200, 300, 400):
        sim.run(t, stop - t)
        t = stop
        on = np.flatnonzero(sim.agent_state()["status"] == 1)
        pick = rng.choice(on, min(50, len(on)), replace=False).astype(np.int32)
        sim.stop_agents(pick, 5, t)
        stopped += len(pick)
        if stop == 100:
            sim.set_link_closed(N + rng.choice(L - N, 40, replace=False).astype(np.int32), np.ones(40, np.uint8))
        st, ls = sim.agent_state()["status"], sim.link_state()
        assert int((st == 1).sum()) == int(ls["q"].sum() + ls["r"].sum())          # enroute <=> on exactly one link list
        ptr, order = sim.link_vehicles()
        assert len(np.unique(order)) == len(order) and (st[order] == 1).all()
    assert int((st == 5).sum()) == stopped and (st == 4).sum() > 0 and (st == 2).sum() > 500
    # node_move_link_pairs (model/queue_class.py:377-383) as counters: one pair per move that is not an arrival
    il, ol, cnt = sim.pair_moves(flat)
    c = sim.counters()
    assert int(cnt.sum()) == c["moves"] - c["arrivals"] and (flat.link_end_node[il] == flat.link_start_node[ol]).all()
    sim.close()
