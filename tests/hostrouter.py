"""An independent host routing path for the tick parity tests (test infrastructure).

It restates what the reference's agent loop does (run_traffic_simulation.py:76-111, queue_class.py:452-455, 491-512):
every `weight_freq` ticks the edge weights are recomputed from the queue lengths; an agent that is due
(t == 0 or t % reroute_freq == agent_id % reroute_freq) gets the shortest path from the end node of its current link to
its destination by a FORWARD Dijkstra from that node (the C oracle's restatement of graph.cc:153-218, raw
first-found predecessors), and its route is replaced by that path.  The product's device router answers the same
question with one reverse tree per destination; this path shares nothing with it.

Trees are memoised per (weight epoch, origin node) and computed on all host cores."""
import multiprocessing as mp
import os

import numpy as np

_G = {}


def _init(s, e, V):
    _G["s"], _G["e"], _G["V"] = s, e, V


def _trees(args):
    w, nodes = args
    from oracle import sssp as osssp
    g = osssp.OracleGraph(_G["s"], _G["e"], w, max_vertex_id=_G["V"] - 1)
    out = np.empty((len(nodes), _G["V"]), np.int32)
    for i, o in enumerate(nodes):
        _, p = g.dijkstra(int(o), -1)
        out[i] = p
    return out


class ForwardRouter:
    def __init__(self, flat, reroute_freq, weight_freq, workers=None):
        self.f = flat
        self.N = len(flat.node_key)
        self.s = np.ascontiguousarray(flat.edge_start, np.int32)
        self.e = np.ascontiguousarray(flat.edge_end, np.int32)
        self.w = np.ascontiguousarray(flat.edge_fft, np.float64).copy()
        self.rf, self.wf = int(reroute_freq), int(weight_freq)
        self.workers = workers or min(os.cpu_count() or 1, 32)
        self.pool = mp.get_context("fork").Pool(self.workers, initializer=_init, initargs=(self.s, self.e, self.N)) \
            if self.workers > 1 else None
        _init(self.s, self.e, self.N)
        key = self.s.astype(np.int64) * self.N + self.e
        self.korder = np.argsort(key, kind="stable")
        self.ksorted = key[self.korder]
        self.row_of = np.full(self.N, -1, np.int64)
        self.par = None
        self.n_rows = 0
        self.trees_computed = 0

    def close(self):
        if self.pool:
            self.pool.terminate()
            self.pool = None

    def refresh_weights(self, link_weights):
        """`link_weights`: the simulator's per-link weights (NaN = keep: closed or source link)."""
        w = np.asarray(link_weights)[self.N:]
        m = ~np.isnan(w)
        self.w[m] = w[m]
        self.row_of[:] = -1
        self.n_rows = 0

    def _ensure(self, nodes):
        need = np.unique(nodes[self.row_of[nodes] < 0])
        if len(need) == 0:
            return
        if self.par is None or self.n_rows + len(need) > len(self.par):
            cap = min(self.N, max(2 * (self.n_rows + len(need)), 1024))
            new = np.empty((cap, self.N), np.int16 if self.N < 32768 else np.int32)
            if self.par is not None and self.n_rows:
                new[: self.n_rows] = self.par[: self.n_rows]
            self.par = new
        if self.pool and len(need) >= 4 * self.workers:
            chunks = np.array_split(need, min(len(need) // 4, 4 * self.workers))
            res = self.pool.map(_trees, [(self.w, c) for c in chunks])
        else:
            chunks, res = [need], [_trees((self.w, need))]
        for c, r in zip(chunks, res):
            self.par[self.n_rows: self.n_rows + len(c)] = r
            self.row_of[c] = np.arange(self.n_rows, self.n_rows + len(c))
            self.n_rows += len(c)
        self.trees_computed += len(need)

    def routes(self, origin_nodes, dest_nodes):
        """-> (ptr i64[n+1], links i32[...]): for every pair the links of the path origin -> destination."""
        o = np.asarray(origin_nodes, np.int64)
        d = np.asarray(dest_nodes, np.int64)
        n = len(o)
        self._ensure(o)
        rows = self.row_of[o]
        cur = d.copy()
        alive = np.flatnonzero(cur != o)
        who, hop, link = [], [], []
        h = 0
        while len(alive):
            p = self.par[rows[alive], cur[alive]].astype(np.int64)
            if (p < 0).any():
                raise RuntimeError("destination unreachable from an agent's node")
            pos = np.searchsorted(self.ksorted, p * self.N + cur[alive])
            who.append(alive)
            hop.append(np.full(len(alive), h))
            link.append(self.N + self.korder[pos])
            cur[alive] = p
            alive = alive[p != o[alive]]
            h += 1
            if h > self.N:
                raise RuntimeError("predecessor chain does not end")
        ptr = np.zeros(n + 1, np.int64)
        if not who:
            return ptr, np.zeros(0, np.int32)
        who, hop, link = np.concatenate(who), np.concatenate(hop), np.concatenate(link)
        length = np.bincount(who, minlength=n)
        ptr[1:] = np.cumsum(length)
        out = np.empty(ptr[-1], np.int32)
        out[ptr[who] + length[who] - 1 - hop] = link
        return ptr, out

    def due(self, t, status, agent_key):
        live = (status == 0) | (status == 1)
        if t == 0:
            return np.flatnonzero(live)
        return np.flatnonzero(live & (agent_key % self.rf == t % self.rf))

    def tick(self, sim, t):
        """What the reference does before the link model of tick t, applied to `sim` (an OracleSim)."""
        if t % self.wf == 0:
            self.refresh_weights(sim.link_weights())
        ag = sim.agent_state()
        idx = self.due(t, ag["status"], np.asarray(self.f.agent_key))
        if len(idx):
            node = self.f.link_end_node[ag["cur"][idx]]
            ptr, links = self.routes(node, np.asarray(self.f.agent_destin)[idx])
            sim.set_routes(idx.astype(np.int32), ptr, links)
        return len(idx)
