"""CPU: the C tick oracle replays every golden scenario produced by the unmodified reference, tick by tick."""
import numpy as np
import pytest

from oracle.tick import OracleSim
from tickcheck import load_golden, replay


@pytest.mark.parametrize("name,every", [("grid8", 1), ("grid7s", 1), ("bolinas", 1), ("grid5x", 5)])
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
