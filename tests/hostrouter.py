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
    w, nodes, want_dist = args
    from oracle import sssp as osssp
    g = osssp.OracleGraph(_G["s"], _G["e"], w, max_vertex_id=_G["V"] - 1)
    out = np.empty((len(nodes), _G["V"]), np.int32)
    dist = np.empty((len(nodes), _G["V"]), np.float64) if want_dist else None
    for i, o in enumerate(nodes):
        d, p = g.dijkstra(int(o), -1)
        out[i] = p
        if want_dist:
            dist[i] = d
    return out, dist


class ForwardRouter:
    """detour_factor / closure_reroute_freq: Butte's two extra rules (model/dta_meso_butte.py:193-209); `closures`:
    [(tick, link, closed)] applied at the start of their tick (Link.close_link_to_newcomers / open_link_to_newcomers,
    queue_class.py:457-464: the edge weighs 1e8 / its free-flow time)."""

    def __init__(self, flat, reroute_freq, weight_freq, workers=None, detour_factor=0.0, closure_reroute_freq=0, closures=()):
        self.f = flat
        self.detour, self.cf = float(detour_factor or 0.0), int(closure_reroute_freq or 0)
        self.keep_dist = bool(self.detour) or len(closures) > 0
        self.events = sorted(((int(t), int(l), int(c)) for t, l, c in closures), key=lambda e: e[0])
        self.ev_cursor = 0
        self.closed = np.zeros(len(flat.link_end_node), bool)
        self.route_dist = np.full(len(flat.agent_key), 5e7)
        self.stopped_a2 = 0
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
        self.dist = None
        self.n_rows = 0
        self.trees_computed = 0
        self.fft = np.ascontiguousarray(flat.edge_fft, np.float64)

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

    def set_closed(self, links, closed):
        links, closed = np.asarray(links, np.int64), np.asarray(closed, bool)
        self.closed[links] = closed
        e = links - self.N
        self.w[e] = np.where(closed, 1e8, self.fft[e])
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
            if self.keep_dist:
                nd = np.empty((cap, self.N), np.float64)
                if self.dist is not None and self.n_rows:
                    nd[: self.n_rows] = self.dist[: self.n_rows]
                self.dist = nd
        if self.pool and len(need) >= 4 * self.workers:
            chunks = np.array_split(need, min(len(need) // 4, 4 * self.workers))
            res = self.pool.map(_trees, [(self.w, c, self.keep_dist) for c in chunks])
        else:
            chunks, res = [need], [_trees((self.w, need, self.keep_dist))]
        for c, (r, dd) in zip(chunks, res):
            self.par[self.n_rows: self.n_rows + len(c)] = r
            if self.keep_dist:
                self.dist[self.n_rows: self.n_rows + len(c)] = dd
            self.row_of[c] = np.arange(self.n_rows, self.n_rows + len(c))
            self.n_rows += len(c)
        self.trees_computed += len(need)

    def distances(self, origin_nodes, dest_nodes):
        o = np.asarray(origin_nodes, np.int64)
        self._ensure(o)
        return self.dist[self.row_of[o], np.asarray(dest_nodes, np.int64)]

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

    def tick(self, sim, t):
        """What the agent loop does before the link model of tick t, applied to `sim` (an OracleSim): closures due,
        weight refresh, reroutes of the regular stagger slot and (Butte) of agents whose next link is closed, the
        detour test, and the removal of sheltering vehicles.  Returns the number of agents rerouted."""
        if t == 0:
            # the routes every agent starts with are computed on the open, empty network (the reference's agent loop
            # runs before the link model that applies the closures of tick 0; the device routes at attach time)
            n0 = self._reroute(sim, 0)
            self._apply_events(sim, 0)
            return n0
        self._apply_events(sim, t)
        if t % self.wf == 0:
            self.refresh_weights(sim.link_weights())
        return self._reroute(sim, t)

    def _apply_events(self, sim, t):
        ev = []
        while self.ev_cursor < len(self.events) and self.events[self.ev_cursor][0] <= t:
            ev.append(self.events[self.ev_cursor])
            self.ev_cursor += 1
        if ev:
            last = {}
            for _, l, c in ev:
                last[l] = c
            links, closed = np.array(list(last), np.int32), np.array(list(last.values()), np.uint8)
            sim.set_link_closed(links, closed)
            self.set_closed(links, closed)

    def _reroute(self, sim, t):
        ag = sim.agent_state()
        key = np.asarray(self.f.agent_key)
        live = (ag["status"] == 0) | (ag["status"] == 1)
        if t == 0:
            due = live
        else:
            due = live & (key % self.rf == t % self.rf)
            if self.cf:
                nxt = ag["nxt"]
                due = due | (live & (key % self.cf == t % self.cf) & (nxt >= 0) & self.closed[np.maximum(nxt, 0)])
        idx = np.flatnonzero(due)
        if len(idx) == 0:
            return 0
        node = self.f.link_end_node[ag["cur"][idx]]
        dest = np.asarray(self.f.agent_destin)[idx]
        if self.keep_dist:
            d = self.distances(node, dest)
            ok = ~(d > 1e8)                 # out of reach: the old route is kept (queue_class.py:501-503)
            idx, node, dest, d = idx[ok], node[ok], dest[ok], d[ok]
            if len(idx) == 0:
                return 0
        ptr, links = self.routes(node, dest)
        sim.set_routes(idx.astype(np.int32), ptr, links)
        if self.keep_dist:
            prev = self.route_dist[idx]
            self.route_dist[idx] = d
            if self.detour and t > 0:
                a2 = idx[(ag["status"][idx] == 1) & (d > prev * self.detour)]
                if len(a2):
                    sim.stop_agents(a2.astype(np.int32), 5, t)      # shelter_a2, off the link before the link model
                    self.stopped_a2 += len(a2)
        return len(idx)
