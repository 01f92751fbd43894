"""Runs the REFERENCE tick (unmodified /root/reference code) on stubs with a counter-based RNG shim and records
golden vectors.  TEST INFRASTRUCTURE; only usable in the build container (/root/reference must exist).

Recipe: SURVEY.md appendix A.  The reference's queue_class.py / run_traffic_simulation.py are imported as they
lie under /root/reference; nothing is copied into this repository.  The only rebinding is the two module-level
names `queue_class.shuffle` and `queue_class.random` plus a context-setting wrapper around
`Node.run_node_model`; no reference file is edited.
"""
import importlib
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference"
REF_PROJECT = REF + "/projects/bolinas_civic"

from spatial_queue_b200 import rng as sqrng  # noqa: E402
from spatial_queue_b200.flatten import STATUS_CODES  # noqa: E402


class _Ctx:
    seed = 0
    t = 0
    node = 0
    n_shuffle = 0
    n_toss = 0


CTX = _Ctx()


def shim_shuffle(lst):
    """Replacement for random.shuffle at queue_class.py:231/236 (slots 0 and 1 of the RNG contract)."""
    k = CTX.n_shuffle
    CTX.n_shuffle += 1
    n = len(lst)
    if n >= 2:
        u = sqrng.uniform(CTX.seed, CTX.t, CTX.node, k)
        r = int(np.floor(u * n))
        lst.insert(0, lst.pop(r))


class ShimRandom:
    """Replacement for the `random` module name inside queue_class (only .choices is used, :360)."""

    @staticmethod
    def choices(population, weights=None, k=1):
        assert k == 1 and len(population) == 2
        j = CTX.n_toss
        CTX.n_toss += 1
        u = sqrng.uniform(CTX.seed, CTX.t, CTX.node, 2 + j)
        w0, w1 = weights
        total = w0 + w1
        return [population[0] if u * total < w0 else population[1]]


_IMPORTED = {}


def import_reference(seed=0):
    """-> (queue_class module, run_traffic_simulation module) of the Bolinas project, shim installed."""
    if _IMPORTED:
        CTX.seed = seed
        return _IMPORTED["qc"], _IMPORTED["rts"]
    if not os.path.isdir(REF_PROJECT):
        raise RuntimeError("/root/reference is not present: the reference harness only runs in the build container")
    lsp = os.path.join(HERE, "_ref", "liblsp.so")
    if not os.path.exists(lsp):
        import subprocess
        subprocess.run(["make", "-s", "-C", HERE, "ref"], check=True)
    tmp = tempfile.mkdtemp(prefix="sq_ref_sp_")
    os.makedirs(os.path.join(tmp, "sp", "build"))
    shutil.copy(REF_PROJECT + "/sp/interface.py", os.path.join(tmp, "sp", "interface.py"))  # /tmp only, never the repo
    open(os.path.join(tmp, "sp", "__init__.py"), "w").close()
    os.symlink(lsp, os.path.join(tmp, "sp", "build", "liblsp.so"))
    sys.path.insert(0, os.path.join(HERE, "refstubs"))
    sys.path.insert(0, tmp)
    import sp.interface  # noqa: F401  (must win over the read-only submodule directory of the project)
    sys.path.insert(0, REF_PROJECT)
    qc = importlib.import_module("queue_class")
    rts = importlib.import_module("run_traffic_simulation")
    qc.shuffle = shim_shuffle
    qc.random = ShimRandom()
    orig = qc.Node.run_node_model

    def run_node_model(self, t_now, *a, **k):
        CTX.t, CTX.node, CTX.n_shuffle, CTX.n_toss = int(t_now), int(self.node_id), 0, 0
        return orig(self, t_now, *a, **k)

    qc.Node.run_node_model = run_node_model
    CTX.seed = seed
    _IMPORTED.update(qc=qc, rts=rts, tmp=tmp)
    return qc, rts


def make_write_path():
    wp = tempfile.mkdtemp(prefix="sq_ref_out_")
    for d in ("demand_inputs/od_csv", "simulation_outputs/log", "simulation_outputs/network",
              "simulation_outputs/link_weights"):
        os.makedirs(os.path.join(wp, d))
    return wp


class Recorder:
    """Captures per-tick state of a reference Network in the flat index space of spatial_queue_b200.flatten."""

    def __init__(self, network, flat, snapshot_ticks=()):
        from .digest import digest
        self._digest = digest
        self.net = network
        self.flat = flat
        self.agents = dict(network.agents)  # keep arrived agents reachable after the driver deletes them
        self.links = list(network.links.values())
        self.nodes = [network.nodes[nid] for nid in flat.node_ids]
        self.active = np.array([STATUS_CODES[a.status] != STATUS_CODES["shelter_p"] for a in self.agents.values()])
        self.snapshot_ticks = set(snapshot_ticks)
        self.digests, self.on_link, self.moves, self.active_nodes = [], [], [], []
        self.congested_total = []
        self.snapshots = {}
        self.arrive_tick = np.full(len(self.agents), -1, np.int32)
        self._prev_move = 0

    def state(self):
        li, ai = self.flat.link_index, self.flat.agent_index
        q = np.array([len(l.queue_vehicles) for l in self.links], np.int32)
        r = np.array([len(l.run_vehicles) for l in self.links], np.int32)
        left = np.array([l.total_vehicles_left for l in self.links], np.int32)
        order = np.array([ai[v] for l in self.links for v in (l.queue_vehicles + l.run_vehicles)], np.int32)
        A = len(self.agents)
        st = np.zeros(A, np.int32); cur = np.full(A, -1, np.int32); nxt = np.full(A, -1, np.int32)
        ent = np.full(A, -1, np.int32)
        for i, a in enumerate(self.agents.values()):
            st[i] = STATUS_CODES[a.status]
            tl = getattr(a, "this_link", None)
            cur[i] = li[tl] if tl is not None else -1
            nxt[i] = li[a.next_link] if a.next_link is not None else -1
            ent[i] = a.current_link_enter_time if a.current_link_enter_time is not None else -1
        nm = np.array([n.node_move for n in self.nodes], np.int64)
        return dict(q=q, r=r, left=left, order=order, status=st, cur=cur, nxt=nxt, enter=ent, node_move=nm)

    def record(self, t):
        s = self.state()
        m = self.active
        self.digests.append(self._digest(s["q"], s["r"], s["left"], s["order"], s["status"][m], s["cur"][m],
                                         s["nxt"][m], s["enter"][m], s["node_move"]))
        self.on_link.append(int(s["q"].sum() + s["r"].sum()))
        tot = int(s["node_move"].sum())
        self.moves.append(tot - self._prev_move)
        self._prev_move = tot
        newly = (s["status"] == STATUS_CODES["arrive"]) & (self.arrive_tick < 0)
        self.arrive_tick[newly] = t
        cong = np.array([l.congested for l in self.links], np.int32)
        self.congested_total.append(int(cong.sum()))
        if t in self.snapshot_ticks:
            s["congested"] = cong
            self.snapshots[t] = s

    def to_npz_dict(self):
        out = dict(digests=np.array(self.digests, np.uint64), on_link=np.array(self.on_link, np.int32),
                   moves=np.array(self.moves, np.int32), arrive_tick=self.arrive_tick,
                   active=self.active.astype(np.uint8), snapshot_ticks=np.array(sorted(self.snapshots), np.int32),
                   # sum over links of link.congested after each tick (run_traffic_simulation.py:80-82, 100-102: the
                   # counter is reset every tick); per link in the snapshots
                   congested_total=np.array(self.congested_total, np.int32))
        for t, s in self.snapshots.items():
            for k, v in s.items():
                out[f"snap{t}_{k}"] = v
        return out


class RouteLog:
    """Wraps Agent.get_path to log every route CHANGE after t=0 as (tick, agent index, link sequence)."""

    def __init__(self, qc, network, flat):
        self.events = []  # (t, agent_idx, [link idx ...])
        self.flat = flat
        self.net = network
        self._orig = qc.Agent.get_path
        self._qc = qc
        log = self

        def get_path(agent, t, g=None):
            from spatial_queue_b200.flatten import route_to_links
            before = None
            if t > 0:
                before = route_to_links(agent.route, agent.current_link_end_nid, log.net.node2link_dict,
                                        log.flat.link_index, len(log.flat.link_ids) + 1)
            res = log._orig(agent, t, g=g)
            if t > 0:
                after = route_to_links(agent.route, agent.current_link_end_nid, log.net.node2link_dict,
                                       log.flat.link_index, len(log.flat.link_ids) + 1)
                if after != before:
                    log.events.append((t, log.flat.agent_index[agent.agent_id], after))
            return res

        qc.Agent.get_path = get_path

    def close(self):
        self._qc.Agent.get_path = self._orig

    def to_npz_dict(self):
        ev_t = np.array([e[0] for e in self.events], np.int32)
        ev_a = np.array([e[1] for e in self.events], np.int32)
        ptr = np.cumsum([0] + [len(e[2]) for e in self.events]).astype(np.int64)
        links = np.array([l for e in self.events for l in e[2]], np.int32)
        return dict(ev_t=ev_t, ev_agent=ev_a, ev_ptr=ptr, ev_links=links)
