"""Shared replay logic: drive a simulator (oracle or device) through a golden scenario and compare per tick."""
import os

import numpy as np

from oracle.digest import digest
from spatial_queue_b200.flatten import FlatNetwork

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    z = np.load(os.path.join(GOLD, f"tick_{name}.npz"))
    flat = FlatNetwork.from_dict({k[5:]: z[k] for k in z.files if k.startswith("flat_")})
    return z, flat


def state_digest(s, active=None):
    m = slice(None) if active is None else active
    return digest(s["q"], s["r"], s["left"], s["order"], s["status"][m], s["cur"][m], s["nxt"][m], s["enter"][m],
                  s["node_move"])


def route_events(z):
    """-> {tick: (agent_idx[], ptr[], links[])} from the golden's ev_* arrays."""
    ev_t, ev_a, ptr, links = z["ev_t"], z["ev_agent"], z["ev_ptr"], z["ev_links"]
    out = {}
    order = np.argsort(ev_t, kind="stable")
    for t in np.unique(ev_t):
        idx = order[np.searchsorted(ev_t[order], t): np.searchsorted(ev_t[order], t, side="right")]
        p = [0]
        ll = []
        for i in idx:
            seg = links[ptr[i]: ptr[i + 1]]
            ll.append(seg)
            p.append(p[-1] + len(seg))
        out[int(t)] = (ev_a[idx].astype(np.int32), np.array(p, np.int64),
                       np.concatenate(ll).astype(np.int32) if ll else np.zeros(0, np.int32))
    return out


def replay(sim, z, n_ticks=None, every=1, on_mismatch=None):
    """Step `sim` through the golden run; assert the digest at every `every`-th tick and at every snapshot."""
    events = route_events(z)
    n = int(z["n_ticks"]) if n_ticks is None else min(int(n_ticks), int(z["n_ticks"]))
    snaps = set(int(t) for t in z["snapshot_ticks"])
    digests = z["digests"]
    for t in range(0, n):
        if t in events:
            sim.set_routes(*events[t])
        sim.step(t)
        if t % every == 0 or t in snaps or t == n - 1:
            s = sim.state()
            d = state_digest(s)
            if d != int(digests[t]):
                detail = ""
                if t in snaps:
                    for k in ("q", "r", "left", "order", "status", "cur", "nxt", "enter", "node_move"):
                        g = z[f"snap{t}_{k}"]
                        if len(g) != len(s[k]) or (np.asarray(g) != np.asarray(s[k])).any():
                            bad = np.nonzero(np.asarray(g[: len(s[k])]) != np.asarray(s[k][: len(g)]))[0][:8]
                            detail += f" [{k}: first diffs at {bad.tolist()} gold {np.asarray(g)[bad].tolist()} got {np.asarray(s[k])[bad].tolist()}]"
                if on_mismatch:
                    on_mismatch(t, s)
                raise AssertionError(f"tick {t}: digest mismatch{detail}")
    return n


def check_congestion_counter(sim, z, every=25):
    """link.congested (reset every tick, run_traffic_simulation.py:80-82, 100-102) against the reference run: the sum
    over links every `every` ticks, the per-link counts at the snapshot ticks.  Needs the shelter rule of the golden
    (3 h) because sheltered vehicles leave the links."""
    sim.set_shelter_after(3 * 3600)
    sim.set_congest_after(1800)
    n = int(z["n_ticks"])
    snaps = set(int(t) for t in z["snapshot_ticks"])
    events = route_events(z)
    checked = 0
    for t in range(n):
        if t in events:
            sim.set_routes(*events[t])
        sim.step(t)
        if t % every == 0 or t in snaps or t == n - 1:
            c = sim.link_congested()
            assert int(c.sum()) == int(z["congested_total"][t]), f"tick {t}"
            if t in snaps:
                assert (c == z[f"snap{t}_congested"]).all(), f"tick {t}"
            checked += int(c.sum() > 0)
    assert checked > 50      # the scenario really has vehicles stuck for more than half an hour
