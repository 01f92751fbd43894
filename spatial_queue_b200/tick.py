"""ctypes front-end of the device tick (include/sq_tick.h): FlatNetwork -> SqNetDesc -> SqSim."""
import ctypes as C

import numpy as np

from . import _lib

_FIELDS = [
    ("n_nodes", C.c_int32), ("n_links", C.c_int32), ("n_agents", C.c_int32), ("reserved0", C.c_int32),
    ("seed", C.c_uint64),
    ("link_start_node", C.c_void_p), ("link_end_node", C.c_void_p), ("link_is_virtual", C.c_void_p),
    ("link_lanes_floor", C.c_void_p), ("link_cap_per_s", C.c_void_p), ("link_storage_cap", C.c_void_p),
    ("link_fft", C.c_void_p), ("link_closed", C.c_void_p), ("link_reverse", C.c_void_p),
    ("link_capacity", C.c_void_p), ("link_maxmph", C.c_void_p),
    ("node_key", C.c_void_p), ("node_type", C.c_void_p), ("node_in_ptr", C.c_void_p), ("node_in_links", C.c_void_p),
    ("node_in_phase", C.c_void_p), ("node_in_prio", C.c_void_p), ("node_out_ptr", C.c_void_p),
    ("node_out_links", C.c_void_p), ("node_turn_ptr", C.c_void_p), ("turn_angle", C.c_void_p),
    ("agent_key", C.c_void_p), ("agent_origin", C.c_void_p), ("agent_destin", C.c_void_p), ("agent_dept", C.c_void_p),
    ("agent_vlen", C.c_void_p), ("agent_status", C.c_void_p), ("route_ptr", C.c_void_p), ("route_links", C.c_void_p),
]
_DTYPES = {
    "link_start_node": np.int32, "link_end_node": np.int32, "link_is_virtual": np.uint8, "link_lanes_floor": np.int32,
    "link_cap_per_s": np.float64, "link_storage_cap": np.float64, "link_fft": np.float64, "link_closed": np.uint8,
    "link_reverse": np.int32, "link_capacity": np.float64, "link_maxmph": np.float64, "node_key": np.int64,
    "node_type": np.uint8, "node_in_ptr": np.int32, "node_in_links": np.int32, "node_in_phase": np.uint8,
    "node_in_prio": np.int32, "node_out_ptr": np.int32, "node_out_links": np.int32, "node_turn_ptr": np.int64,
    "turn_angle": np.float64, "agent_key": np.int64, "agent_origin": np.int32, "agent_destin": np.int32,
    "agent_dept": np.int32, "agent_vlen": np.int32, "agent_status": np.uint8, "route_ptr": np.int64,
    "route_links": np.int32,
}


class SqNetDesc(C.Structure):
    _fields_ = _FIELDS


class SqRouterDesc(C.Structure):
    """include/sq_tick.h SqRouterDesc."""
    _fields_ = [("graph", C.c_void_p), ("n_graph_edges", C.c_int32), ("n_dest", C.c_int32),
                ("dest_nodes", C.c_void_p), ("agent_dslot", C.c_void_p), ("agent_key", C.c_void_p),
                ("node_vertex", C.c_void_p), ("link_edge", C.c_void_p), ("n_pairs", C.c_int32),
                ("pair_start", C.c_void_p), ("pair_end", C.c_void_p), ("pair_link", C.c_void_p),
                ("reroute_freq", C.c_int32), ("weight_freq", C.c_int32)]


def make_desc(flat, seed, flags=0):
    """-> (SqNetDesc, keepalive list).  `flat` is a flatten.FlatNetwork (or anything with the same arrays)."""
    d = SqNetDesc()
    d.reserved0 = int(flags)
    keep = []
    d.n_nodes = len(flat.node_key)
    d.n_links = len(flat.link_end_node)
    d.n_agents = len(flat.agent_key)
    d.seed = int(seed)
    for name, dt in _DTYPES.items():
        a = np.ascontiguousarray(getattr(flat, name), dtype=dt)
        if a.size == 0:
            a = np.zeros(1, dt)
        keep.append(a)
        setattr(d, name, a.ctypes.data)
    return d, keep


class _SimBase:
    """Shared read-back / control surface; subclasses bind the function prefix of a library."""
    _prefix = None

    def _fn(self, name):
        raise NotImplementedError

    def _check(self, rc, what):
        raise NotImplementedError

    def step(self, t):
        self._check(self._fn("step")(self._h, int(t)), "step")

    def set_congest_after(self, seconds):
        self._check(self._fn("set_congest_after")(self._h, int(seconds)), "set_congest_after")

    def stop_agents(self, agent_idx, status, t):
        """A driver rule takes enroute agents off the road between ticks (status code >= 3, e.g. 5 = shelter_a2):
        they leave their link's lists at once (model/dta_meso_butte.py:256-263)."""
        a = np.ascontiguousarray(agent_idx, np.int32)
        self._check(self._fn("stop_agents")(self._h, len(a), a.ctypes.data, int(status), int(t)), "stop_agents")

    def link_congested(self):
        out = np.empty(max(self.n_links, 1), np.int64)
        self._check(self._fn("get_link_congested")(self._h, out.ctypes.data), "get_link_congested")
        return out[: self.n_links]

    def run(self, t0, n):
        self._check(self._fn("run")(self._h, int(t0), int(n)), "run")

    def set_routes(self, agent_idx, ptr, links):
        a = np.ascontiguousarray(agent_idx, np.int32)
        p = np.ascontiguousarray(ptr, np.int64)
        l = np.ascontiguousarray(links, np.int32)
        if l.size == 0:
            l = np.zeros(1, np.int32)
        self._check(self._fn("set_routes")(self._h, len(a), a.ctypes.data, p.ctypes.data, l.ctypes.data), "set_routes")

    def set_link_closed(self, link_idx, closed):
        a = np.ascontiguousarray(link_idx, np.int32)
        c = np.ascontiguousarray(closed, np.uint8)
        self._check(self._fn("set_link_closed")(self._h, len(a), a.ctypes.data, c.ctypes.data), "set_link_closed")

    def link_state(self):
        L = self.n_links
        q, r, left = np.empty(L, np.int32), np.empty(L, np.int32), np.empty(L, np.int32)
        ic, oc, sc = np.empty(L), np.empty(L), np.empty(L)
        self._check(self._fn("get_link_state")(self._h, q.ctypes.data, r.ctypes.data, left.ctypes.data, ic.ctypes.data,
                                               oc.ctypes.data, sc.ctypes.data), "get_link_state")
        return dict(q=q, r=r, left=left, in_c=ic, out_c=oc, storage_c=sc)

    def link_vehicles(self):
        ptr = np.empty(self.n_links + 1, np.int64)
        ag = np.empty(max(self.n_agents, 1), np.int32)
        self._check(self._fn("get_link_vehicles")(self._h, ptr.ctypes.data, ag.ctypes.data), "get_link_vehicles")
        return ptr, ag[: ptr[-1]]

    def agent_state(self):
        A = self.n_agents
        st = np.empty(A, np.uint8)
        cur, nxt, ent, arr = (np.empty(A, np.int32) for _ in range(4))
        self._check(self._fn("get_agent_state")(self._h, st.ctypes.data, cur.ctypes.data, nxt.ctypes.data,
                                                ent.ctypes.data, arr.ctypes.data), "get_agent_state")
        return dict(status=st, cur=cur, nxt=nxt, enter=ent, arrive_tick=arr)

    def node_state(self):
        nm = np.empty(self.n_nodes, np.int64)
        self._check(self._fn("get_node_state")(self._h, nm.ctypes.data), "get_node_state")
        return dict(node_move=nm)

    def link_weights(self):
        w = np.empty(self.n_links, np.float64)
        self._check(self._fn("get_link_weights")(self._h, w.ctypes.data), "get_link_weights")
        return w

    def pair_moves(self, flat):
        """Vehicles moved per (in_link, out_link) pair -- what Node.run_node_model returns as node_move_link_pairs
        (model/queue_class.py:377-383), counted: -> (in_link i32[], out_link i32[], count i64[]) over the pairs with at
        least one move.  `flat` supplies the node CSR the table is shaped after.  DeviceSim: enable_pair_moves first."""
        n_turn = int(flat.node_turn_ptr[-1])
        tbl = np.zeros(max(n_turn, 1), np.int64)
        self._check(self._fn("get_pair_moves")(self._h, tbl.ctypes.data), "get_pair_moves")
        tbl = tbl[:n_turn]
        idx = np.flatnonzero(tbl)
        node = np.searchsorted(np.asarray(flat.node_turn_ptr), idx, side="right") - 1
        outdeg = (np.asarray(flat.node_out_ptr)[node + 1] - np.asarray(flat.node_out_ptr)[node]).astype(np.int64)
        rel = idx - np.asarray(flat.node_turn_ptr)[node]
        il = np.asarray(flat.node_in_links)[np.asarray(flat.node_in_ptr)[node] + rel // outdeg]
        ol = np.asarray(flat.node_out_links)[np.asarray(flat.node_out_ptr)[node] + rel % outdeg]
        return il.astype(np.int32), ol.astype(np.int32), tbl[idx]

    def counters(self):
        out = np.zeros(6, np.int64)
        self._check(self._fn("get_counters")(self._h, out.ctypes.data), "get_counters")
        return dict(zip(("ticks", "agent_steps", "moves", "active_nodes", "arrivals", "launches"), out.tolist()))

    def state(self):
        """Everything the parity digest covers, in the layout of oracle.ref_harness.Recorder.state()."""
        ls, ag, nd = self.link_state(), self.agent_state(), self.node_state()
        _, order = self.link_vehicles()
        return dict(q=ls["q"], r=ls["r"], left=ls["left"], order=order, status=ag["status"].astype(np.int32),
                    cur=ag["cur"], nxt=ag["nxt"], enter=ag["enter"], node_move=nd["node_move"],
                    arrive_tick=ag["arrive_tick"], in_c=ls["in_c"], out_c=ls["out_c"], storage_c=ls["storage_c"])


class DeviceSim(_SimBase):
    """The GPU tick.  Raises B200LibraryError when the library or the device is missing."""

    def __init__(self, flat, seed=0, routes_from_router=False):
        """`routes_from_router=True`: agents carry no routes; `attach_router` must follow before the first tick."""
        L = _lib.lib()
        self._L = L
        self.n_nodes, self.n_links, self.n_agents = len(flat.node_key), len(flat.link_end_node), len(flat.agent_key)
        desc, keep = make_desc(flat, seed, flags=1 if routes_from_router else 0)
        self._h = L.sq_create(C.byref(desc))
        if not self._h:
            raise _lib.B200LibraryError("sq_create failed: " + L.sq_last_error().decode())

    def _fn(self, name):
        return getattr(self._L, "sq_" + name)

    def _check(self, rc, what):
        if rc != 0:
            raise _lib.B200LibraryError(f"sq_{what} failed ({rc}): {self._L.sq_last_error().decode()}")

    def summary(self):
        """include/sq_tick.h sq_get_summary: the drivers' periodic statistics from one 128-byte read-back."""
        o = np.zeros(16, np.int64)
        self._check(self._L.sq_get_summary(self._h, o.ctypes.data), "get_summary")
        names = ("unloaded", "enroute", "arrive", "shelter_p", "shelter_a1", "shelter_a2")
        d = {k: int(o[i]) for i, k in enumerate(names)}
        d.update(queued=int(o[8]), running=int(o[9]), congested=int(o[10]), closed_links=int(o[11]), node_moves=int(o[12]),
                 vehicles_left=int(o[13]), links_in_use=int(o[14]), on_closed_links=int(o[15]))
        return d

    def set_closure_schedule(self, ticks, link_idx, closed):
        """Timed closures / reopenings as a device event table (include/sq_tick.h sq_set_closure_schedule)."""
        t = np.ascontiguousarray(ticks, np.int32)
        a = np.ascontiguousarray(link_idx, np.int32)
        c = np.ascontiguousarray(closed, np.uint8)
        assert len(t) == len(a) == len(c)
        self._check(self._L.sq_set_closure_schedule(self._h, len(t), t.ctypes.data, a.ctypes.data, c.ctypes.data),
                    "set_closure_schedule")

    def attach_router(self, graph, flat, reroute_freq=300, weight_freq=100):
        """Device-side rerouting on `graph` (an sp_interface.Graph whose edges are the real links).  `flat` needs
        the router arrays of flatten.router_arrays / synth.flat_grid: dest_nodes, agent_dslot, node_vertex, link_edge
        and the (start, end) -> link pairs."""
        r = SqRouterDesc()
        keep = []

        def arr(a, dt):
            a = np.ascontiguousarray(a, dtype=dt)
            if a.size == 0:
                a = np.zeros(1, dt)
            keep.append(a)
            return a.ctypes.data
        r.graph = graph._h
        r.n_graph_edges = int(graph.nedges_logged)
        r.n_dest = len(flat.dest_nodes)
        r.dest_nodes = arr(flat.dest_nodes, np.int32)
        r.agent_dslot = arr(flat.agent_dslot, np.int32)
        r.agent_key = arr(flat.agent_key, np.int64)
        r.node_vertex = arr(flat.node_vertex, np.int32)
        r.link_edge = arr(flat.link_edge, np.int32)
        r.n_pairs = len(flat.pair_link)
        r.pair_start = arr(flat.pair_start, np.int32)
        r.pair_end = arr(flat.pair_end, np.int32)
        r.pair_link = arr(flat.pair_link, np.int32)
        r.reroute_freq, r.weight_freq = int(reroute_freq), int(weight_freq)
        self._check(self._L.sq_attach_router(self._h, C.byref(r)), "attach_router")
        self._router_graph = graph   # keep the graph alive as long as the sim uses it

    def set_reroute_rules(self, detour_factor=0.0, closure_reroute_freq=0):
        """Butte's shelter_a2 detour rule and reroute-upon-closure inside the device router (include/sq_tick.h)."""
        self._check(self._L.sq_set_reroute_rules(self._h, float(detour_factor or 0.0), int(closure_reroute_freq or 0)),
                    "set_reroute_rules")

    def enable_pair_moves(self):
        """Start counting moves per (in_link, out_link) pair (include/sq_tick.h sq_enable_pair_moves)."""
        self._check(self._L.sq_enable_pair_moves(self._h), "enable_pair_moves")

    def enable_timeline(self, t0, n_ticks):
        """Measurement aid (include/sq_tick.h sq_enable_timeline): warp entry / exit times of both phase kernels."""
        self._check(self._L.sq_enable_timeline(self._h, int(t0), int(n_ticks)), "enable_timeline")

    def timeline(self):
        """-> u64[n_ticks, 2 (link, node), warps, 2 (entry, exit)] in ns of %globaltimer; unused slots are 0."""
        import ctypes as C
        n, st = C.c_int32(0), C.c_int32(0)
        self._check(self._L.sq_get_timeline(self._h, C.addressof(n), C.addressof(st), None), "get_timeline")
        out = np.zeros((n.value, 2, st.value, 2), np.uint64)
        self._check(self._L.sq_get_timeline(self._h, None, None, out.ctypes.data), "get_timeline")
        return out

    def set_shelter_after(self, seconds):
        self._check(self._L.sq_set_shelter_after(self._h, int(seconds)), "set_shelter_after")

    def shelter_state(self):
        ls, st = np.empty(self.n_links, np.int32), np.empty(max(self.n_agents, 1), np.int32)
        self._check(self._L.sq_get_shelter_state(self._h, ls.ctypes.data, st.ctypes.data), "get_shelter_state")
        return dict(link_shelter=ls, stop_tick=st[: self.n_agents])

    def sync(self):
        self._check(self._L.sq_sync(self._h), "sync")

    def set_stream(self, cuda_stream_ptr):
        self._check(self._L.sq_set_stream(self._h, C.c_void_p(cuda_stream_ptr)), "set_stream")

    def close(self):
        if self._h:
            self._L.sq_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
