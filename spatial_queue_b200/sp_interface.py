"""Host-side mirror of the reference's `sp/interface.py` on top of the B200 library.

Same names, argument meaning and return conventions as projects/bolinas_civic/sp/interface.py:15-67
(`ShortestPath.origin/distance/parent/route/clear`, `Graph.dijkstra/update_edge/writegraph`,
`simplegraph`, `readgraph`, `from_dataframe`), so `queue_class.py` can bind to it unchanged
(`from sp import interface` -> `from spatial_queue_b200 import sp_interface as interface`).

Additive: `Graph.sssp_batch` (one tree per origin, host arrays), `Graph.sssp_batch_device` (torch CUDA
tensors handed over as raw device pointers), `Graph.set_weights`, `Graph.remove_edge`.
"""
import ctypes as C

import numpy as np

from ._lib import lib, check

DBL_MAX = np.finfo(np.float64).max


def pinned_empty(shape, dtype):
    """numpy array over page-locked host memory from the library (sp_alloc_host); freed with the array."""
    import weakref
    dtype = np.dtype(dtype)
    nbytes = int(np.prod(shape)) * dtype.itemsize
    ptr = lib().sp_alloc_host(nbytes)
    if not ptr:
        raise MemoryError(f"sp_alloc_host({nbytes}) failed: {lib().sp_last_error().decode()}")
    buf = (C.c_char * max(nbytes, 1)).from_address(ptr)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    weakref.finalize(buf, lib().sp_free_host, ptr)
    return arr


class ShortestPath:
    """One shortest-path tree (reference: interface.py:15-32, struct ShortestPath in sp/include/graph.h:20-53)."""

    def __init__(self, handle):
        self._h = handle

    @property
    def origin(self):
        return lib().origin(self._h)

    def distance(self, destination):
        return lib().distance(self._h, int(destination))

    def parent(self, destination):
        return lib().parent(self._h, int(destination))

    def route(self, destination):
        """Yield (parent, node) hops from the origin to `destination` (interface.py:26-30).

        The reference recurses once per hop (and dies past ~990 hops); this walks iteratively and is bounded by
        the vertex count, so a broken predecessor chain cannot loop forever."""
        origin = self.origin
        hops = []
        node = int(destination)
        limit = 1 << 31
        while node != origin:
            par = self.parent(node)
            if par < 0:
                raise KeyError(f"vertex {node} has no predecessor towards origin {origin}")
            hops.append((par, node))
            node = par
            limit -= 1
            if limit <= 0 or len(hops) > 50_000_000:
                raise RuntimeError("predecessor chain does not terminate")
        return iter(reversed(hops))

    def clear(self):
        if self._h:
            lib().clear(self._h)
            self._h = None


class Graph:
    """Reference: interface.py:36-42 (class Graph) over sp/src/py.cc."""

    def __init__(self, handle):
        self._h = handle

    # -- reference surface --
    def dijkstra(self, origin, destination):
        return ShortestPath(lib().dijkstra(self._h, int(origin), int(destination)))

    def update_edge(self, origin, destination, weight):
        w = weight.value if isinstance(weight, C.c_double) else float(weight)
        return lib().update_edge(self._h, int(origin), int(destination), w)

    def writegraph(self, filename):
        if isinstance(filename, str):
            filename = filename.encode()
        return lib().writegraph(self._h, filename)

    # -- additive --
    @property
    def nvertices(self):
        return lib().sp_nvertices(self._h)

    @property
    def nedges(self):
        return lib().sp_nedges(self._h)

    @property
    def nedges_logged(self):
        """Edges ever inserted (creategraph order; what sp_set_weights / sp_set_weights_device index)."""
        return lib().sp_nedges_logged(self._h)

    def remove_edge(self, v1, v2):
        check(lib().sp_remove_edge(self._h, int(v1), int(v2)), "remove_edge")

    def set_weights(self, weights):
        w = np.ascontiguousarray(weights, dtype=np.float64)
        check(lib().sp_set_weights(self._h, w.ctypes.data, len(w)), "sp_set_weights")

    def set_tuning(self, threads=0, ctas_per_sm=-1, delta_factor=0.0, smem_queue=-1):
        check(lib().sp_set_tuning(self._h, threads, ctas_per_sm, float(delta_factor), smem_queue), "sp_set_tuning")

    def set_parent_format(self, fmt):
        """0 = int32 vertex ids, 1 = one-byte in-edge codes (device batches; include/sq_sp.h)."""
        check(lib().sp_set_parent_format(self._h, int(fmt)), "sp_set_parent_format")

    def decode_parents_device(self, codes_dev, parent_dev, stream=None, reverse=False):
        import torch
        sp = C.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)
        check(lib().sp_decode_parents_device(self._h, C.c_void_p(codes_dev.data_ptr()), int(codes_dev.shape[0]),
                                             C.c_void_p(parent_dev.data_ptr()), sp, 1 if reverse else 0),
              "sp_decode_parents_device")

    def set_algo(self, algo=3, queue_factor=0.0):
        """3 = auto (default), 4 = group kernel, 2 = warp kernel, 0 = ELL kernel, 1 = bitmap kernel; see include/sq_sp.h."""
        check(lib().sp_set_algo(self._h, int(algo), float(queue_factor)), "sp_set_algo")

    def set_warp_tuning(self, ring_cap=0, ilp=0, ctas_per_sm=-1):
        check(lib().sp_set_warp_tuning(self._h, int(ring_cap), int(ilp), int(ctas_per_sm)), "sp_set_warp_tuning")

    def set_group_tuning(self, warps_per_sm=0, batch=0, prefetch=-1, min_trees=0, min_density=0):
        """Group kernel (algo 4 / auto on dense batches): see include/sq_sp.h."""
        check(lib().sp_set_group_tuning(self._h, int(warps_per_sm), int(batch), int(prefetch), int(min_trees),
                                        int(min_density)), "sp_set_group_tuning")

    def set_smem_rows(self, on=True):
        """Lone trees keep their distance row in shared memory when the graph fits (include/sq_sp.h)."""
        check(lib().sp_set_smem_rows(self._h, 1 if on else 0), "sp_set_smem_rows")

    def set_group_mode(self, mode=0, ring_cap=0):
        """0 = grouped warp kernel (default), 1 = sssp_group_kernel; see include/sq_sp.h."""
        check(lib().sp_set_group_mode(self._h, int(mode), int(ring_cap)), "sp_set_group_mode")

    def sync(self):
        check(lib().sp_sync(self._h), "sp_sync")

    def sssp_batch(self, origins, want_dist=True, want_parent=True, pinned=True, out=None, reverse=False):
        """Full trees for `origins` -> (dist f64[n,V] or None, parent i32[n,V] or None), host arrays.

        The results land in page-locked memory by default so that the device-to-host copies overlap the
        computation (`pinned=False` gives ordinary numpy arrays; `out=(dist, parent)` reuses caller arrays)."""
        o = np.ascontiguousarray(origins, dtype=np.int32)
        V = self.nvertices
        if out is not None:
            d, p = out
        else:
            alloc = pinned_empty if (pinned and len(o) * V >= (1 << 16)) else np.empty
            d = alloc((len(o), V), np.float64) if want_dist else None
            p = alloc((len(o), V), np.int32) if want_parent else None
        want_dist, want_parent = d is not None, p is not None
        check(lib().sp_sssp_batch_ex(self._h, o.ctypes.data, len(o),
                                     d.ctypes.data if want_dist else None,
                                     p.ctypes.data if want_parent else None, 1 if reverse else 0), "sp_sssp_batch")
        return d, p

    def sssp_batch_codes(self, origins, want_dist=False, out=None, reverse=False):
        """Full trees for `origins` with the predecessors as one-byte in-edge codes (sp_sssp_batch_codes):
        -> (dist f64[n,V] or None, codes u8[n,V]) in page-locked host memory; `out=(dist or None, codes)` reuses arrays."""
        o = np.ascontiguousarray(origins, dtype=np.int32)
        V = self.nvertices
        if out is not None:
            d, c = out
        else:
            d = pinned_empty((len(o), V), np.float64) if want_dist else None
            c = pinned_empty((len(o), V), np.uint8)
        check(lib().sp_sssp_batch_codes(self._h, o.ctypes.data, len(o), d.ctypes.data if d is not None else None,
                                        c.ctypes.data, 1 if reverse else 0), "sp_sssp_batch_codes")
        return d, c

    def sssp_batch_codes4(self, origins, want_dist=False, out=None, reverse=False):
        """sssp_batch_codes with two vertices per byte (sp_sssp_batch_codes4): codes u8[n, (V + 1) // 2], low nibble = even
        vertex, 15 = none; graphs with in-degree <= 15 only."""
        o = np.ascontiguousarray(origins, dtype=np.int32)
        V = self.nvertices
        if out is not None:
            d, c = out
        else:
            d = pinned_empty((len(o), V), np.float64) if want_dist else None
            c = pinned_empty((len(o), (V + 1) // 2), np.uint8)
        check(lib().sp_sssp_batch_codes4(self._h, o.ctypes.data, len(o), d.ctypes.data if d is not None else None,
                                         c.ctypes.data, 1 if reverse else 0), "sp_sssp_batch_codes4")
        return d, c

    def decode_parents_host4(self, codes, reverse=False):
        """codes u8[n, (V + 1) // 2] -> parent i32[n, V] on the host (sp_decode_parents_host4)."""
        c = np.ascontiguousarray(codes, dtype=np.uint8)
        p = np.empty((c.shape[0], self.nvertices), np.int32)
        check(lib().sp_decode_parents_host4(self._h, c.ctypes.data, int(c.shape[0]), p.ctypes.data, 1 if reverse else 0),
              "sp_decode_parents_host4")
        return p

    def decode_parents_host(self, codes, reverse=False):
        """codes u8[n,V] -> parent i32[n,V] on the host (sp_decode_parents_host)."""
        c = np.ascontiguousarray(codes, dtype=np.uint8)
        p = np.empty(c.shape, np.int32)
        check(lib().sp_decode_parents_host(self._h, c.ctypes.data, int(c.shape[0]), p.ctypes.data, 1 if reverse else 0),
              "sp_decode_parents_host")
        return p

    def route_batch(self, origins, dests):
        """`Agent.get_path` for many agents in one call (queue_class.py:491-512): for every pair the distance
        origin -> destination (DBL_MAX if unreachable) and the path as a vertex sequence.  Returns
        (dist f64[n], off i64[n], length i32[n], pool i32[total]): path k = pool[off[k] : off[k] + length[k]] =
        origin, ..., destination (empty when unreachable or origin == destination).  One tree per distinct origin
        is computed and walked on the GPU; only distances and paths come back."""
        o = np.ascontiguousarray(origins, dtype=np.int32)
        d = np.ascontiguousarray(dests, dtype=np.int32)
        if o.shape != d.shape:
            raise ValueError("origins and dests must have the same length")
        n = len(o)
        dist, off, ln = np.empty(n, np.float64), np.empty(n, np.int64), np.empty(n, np.int32)
        check(lib().sp_route_batch(self._h, o.ctypes.data, d.ctypes.data, n, dist.ctypes.data, off.ctypes.data,
                                   ln.ctypes.data), "sp_route_batch")
        total = int((off + ln).max()) if n else 0
        pool = np.empty(total, np.int32)
        if total:
            C.memmove(pool.ctypes.data, lib().sp_route_pool(self._h), 4 * total)
        return dist, off, ln, pool

    def set_weights_device(self, weights_dev, stream=None):
        """Per-edge weights from a torch CUDA float64 tensor (creategraph order, NaN = keep)."""
        import torch
        sp = C.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)
        check(lib().sp_set_weights_device(self._h, C.c_void_p(weights_dev.data_ptr()), int(weights_dev.numel()), sp),
              "sp_set_weights_device")

    def sssp_batch_device(self, origins_dev, dist_dev=None, parent_dev=None, stream=None, reverse=False):
        """Device-resident batch.  Arguments are torch CUDA tensors (int32 origins [n]; float64 dist [n,V];
        int32 parent [n,V]); only their data pointers cross the C ABI.  Asynchronous on `stream`."""
        n = int(origins_dev.numel())
        sp = C.c_void_p(stream.cuda_stream) if stream is not None else None
        if stream is None:
            import torch
            sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        check(lib().sp_sssp_batch_device_ex(self._h, C.c_void_p(origins_dev.data_ptr()), n,
                                            C.c_void_p(dist_dev.data_ptr()) if dist_dev is not None else None,
                                            C.c_void_p(parent_dev.data_ptr()) if parent_dev is not None else None,
                                            sp, 1 if reverse else 0), "sp_sssp_batch_device")

    def sssp_batch_device_noalloc(self, origins_dev, dist_dev, parent_dev=None, stream=None):
        """sssp_batch_device for callers that must not allocate (include/sq_sp.h sp_sssp_batch_device_noalloc): static
        scratch, at most 65,536 vertices, graph already synced; calls that overlap on one GPU race."""
        import torch
        sp = C.c_void_p((stream or torch.cuda.current_stream()).cuda_stream)
        check(lib().sp_sssp_batch_device_noalloc(self._h, C.c_void_p(origins_dev.data_ptr()), int(origins_dev.numel()),
                                                 C.c_void_p(dist_dev.data_ptr()),
                                                 C.c_void_p(parent_dev.data_ptr()) if parent_dev is not None else None, sp),
              "sp_sssp_batch_device_noalloc")

    def last_stats(self):
        out = (C.c_longlong * 5)()
        lib().sp_last_stats(self._h, out)
        return dict(zip(("trees", "rounds", "relaxations", "transitions", "overflows"), list(out)))

    def free(self):
        if self._h:
            lib().sp_free_graph(self._h)
            self._h = None


def simplegraph(directed=True):
    return Graph(lib().simplegraph(bool(directed)))


def readgraph(filename, directed=True):
    if isinstance(filename, str):
        filename = filename.encode()
    return Graph(lib().readgraph(filename, bool(directed)))


def from_arrays(starts, ends, weights, max_vertex_id=None, directed=True):
    s = np.ascontiguousarray(starts, dtype=np.int32)
    e = np.ascontiguousarray(ends, dtype=np.int32)
    w = np.ascontiguousarray(weights, dtype=np.float64)
    if max_vertex_id is None:
        max_vertex_id = int(max(s.max(), e.max()))
    return Graph(lib().creategraph(s.ctypes.data, e.ctypes.data, w.ctypes.data, len(s), int(max_vertex_id), bool(directed)))


def from_dataframe(edges=None, start_node_col=None, end_node_col=None, weight_col=None, directed=True):
    """interface.py:59-67: edge count from the frame, vertex count = largest id seen in either column."""
    return from_arrays(edges[start_node_col].values, edges[end_node_col].values,
                       np.asarray(edges[weight_col].values, dtype=np.float64),
                       int(np.max(edges[[start_node_col, end_node_col]].values)), directed)
