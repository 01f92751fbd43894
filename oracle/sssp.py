"""ctypes bindings of the SSSP oracles (test infrastructure only; see oracle/__init__.py)."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = None


def build(ref=True):
    """Compile liboracle.so (and _ref/liblsp.so when /root/reference is present)."""
    subprocess.run(["make", "-s", "-C", HERE, "all" if ref else os.path.join(HERE, "liboracle.so")],
                   check=True, capture_output=True)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.oracle_graph_create.restype = C.c_void_p
        L.oracle_graph_create.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.oracle_graph_free.argtypes = [C.c_void_p]
        L.oracle_graph_nvertices.argtypes = [C.c_void_p]
        L.oracle_dijkstra.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.oracle_canonical_parents.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        L.oracle_sssp_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        _LIB = L
    return _LIB


class OracleGraph:
    """sp Graph semantics (graph.cc:4-28, 138-150) + Dijkstra (graph.cc:153-218), restated in C."""

    def __init__(self, starts, ends, weights, max_vertex_id=None, directed=True):
        self.s = np.ascontiguousarray(starts, dtype=np.int32)
        self.e = np.ascontiguousarray(ends, dtype=np.int32)
        self.w = np.ascontiguousarray(weights, dtype=np.float64)
        if max_vertex_id is None:
            max_vertex_id = int(max(self.s.max(), self.e.max())) if len(self.s) else 0
        self.h = lib().oracle_graph_create(self.s.ctypes.data, self.e.ctypes.data, self.w.ctypes.data,
                                           len(self.s), int(max_vertex_id), int(bool(directed)))
        self.V = lib().oracle_graph_nvertices(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().oracle_graph_free(self.h)
            self.h = None

    def dijkstra(self, origin, destination=-1):
        """-> (dist f64[V], parent_raw i32[V]); early exit at `destination` like the reference."""
        d = np.empty(self.V, dtype=np.float64)
        p = np.empty(self.V, dtype=np.int32)
        lib().oracle_dijkstra(self.h, int(origin), int(destination), d.ctypes.data, p.ctypes.data)
        return d, p

    def canonical_parents(self, origin, dist):
        p = np.empty(self.V, dtype=np.int32)
        dist = np.ascontiguousarray(dist, dtype=np.float64)
        lib().oracle_canonical_parents(self.h, int(origin), dist.ctypes.data, p.ctypes.data)
        return p

    def sssp_batch(self, origins, want_parent=True, want_dist=True):
        o = np.ascontiguousarray(origins, dtype=np.int32)
        d = np.empty((len(o), self.V), dtype=np.float64) if want_dist else None
        p = np.empty((len(o), self.V), dtype=np.int32) if want_parent else None
        lib().oracle_sssp_batch(self.h, o.ctypes.data, len(o), d.ctypes.data if want_dist else None,
                                p.ctypes.data if want_parent else None)
        return d, p


# ---- the real reference library (oracle/_ref/liblsp.so), when it was built ---------------------------

def ref_available():
    return os.path.exists(os.path.join(HERE, "_ref", "liblsp.so"))


def ref_lib():
    """The reference's own sp library, bound exactly like sp/interface.py:12-51 binds it."""
    global _REF
    if _REF is None:
        L = C.CDLL(os.path.join(HERE, "_ref", "liblsp.so"))
        L.distance.restype = C.c_double
        L.distance.argtypes = [C.c_void_p, C.c_int]
        L.parent.argtypes = [C.c_void_p, C.c_int]
        L.origin.argtypes = [C.c_void_p]
        L.clear.argtypes = [C.c_void_p]
        L.dijkstra.restype = C.c_void_p
        L.dijkstra.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.update_edge.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double]
        L.simplegraph.restype = C.c_void_p
        L.simplegraph.argtypes = [C.c_bool]
        L.readgraph.restype = C.c_void_p
        L.readgraph.argtypes = [C.c_char_p, C.c_bool]
        L.creategraph.restype = C.c_void_p
        L.creategraph.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_bool]
        L.writegraph.argtypes = [C.c_void_p, C.c_char_p]
        _REF = L
    return _REF


class RefGraph:
    """Thin handle on the reference library for differential tests and the CPU baseline."""

    def __init__(self, starts=None, ends=None, weights=None, max_vertex_id=None, directed=True, handle=None):
        L = ref_lib()
        if handle is not None:
            self.h = handle
            self.V = None
            return
        s = np.ascontiguousarray(starts, dtype=np.int32)
        e = np.ascontiguousarray(ends, dtype=np.int32)
        w = np.ascontiguousarray(weights, dtype=np.float64)
        if max_vertex_id is None:
            max_vertex_id = int(max(s.max(), e.max()))
        self.h = L.creategraph(s.ctypes.data, e.ctypes.data, w.ctypes.data, len(s), int(max_vertex_id), bool(directed))
        self.V = int(max_vertex_id) + 1

    @classmethod
    def simple(cls, directed=True):
        g = cls(handle=ref_lib().simplegraph(bool(directed)))
        g.V = 7
        return g

    def update_edge(self, a, b, w):
        ref_lib().update_edge(self.h, int(a), int(b), float(w))

    def dijkstra(self, origin, destination=-1, V=None):
        """-> (dist, parent) read back through the reference's accessors (py.cc:69-87)."""
        L = ref_lib()
        V = V or self.V
        sp = L.dijkstra(self.h, int(origin), int(destination))
        d = np.array([L.distance(sp, v) for v in range(V)], dtype=np.float64)
        p = np.array([L.parent(sp, v) for v in range(V)], dtype=np.int32)
        L.clear(sp)
        return d, p

    def one_to_one(self, origin, destination):
        L = ref_lib()
        sp = L.dijkstra(self.h, int(origin), int(destination))
        d = L.distance(sp, int(destination))
        L.clear(sp)
        return d

    def time_trees(self, origins):
        """Run full trees (destination -1) and free them; returns nothing (used by the timed CPU leg)."""
        L = ref_lib()
        for o in origins:
            L.clear(L.dijkstra(self.h, int(o), -1))
