"""CPU: the C tick oracle replays every golden scenario produced by the unmodified reference, tick by tick."""
import numpy as np
import pytest

from oracle.tick import OracleSim
from tickcheck import load_golden, replay


@pytest.mark.parametrize("name,every", [("grid8", 1), ("grid7s", 1), ("bolinas", 1)])
def test_oracle_replays_reference_golden(name, every):
    z, flat = load_golden(name)
    sim = OracleSim(flat, seed=int(z["seed"]))
    n = replay(sim, z, every=every)
    ag = sim.agent_state()
    assert (ag["arrive_tick"] == z["arrive_tick"]).all()
    c = sim.counters()
    assert c["agent_steps"] == int(z["on_link"][:n].astype(np.int64).sum())
    assert c["moves"] == int(z["moves"][:n].astype(np.int64).sum())
