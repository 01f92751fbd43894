"""Row B1, literally: libsqb200.so under the reference's own ctypes binding (projects/bolinas_civic/sp/interface.py).

In the build container the reference file itself is staged in a temporary `sp/` package whose build/liblsp.so is the
product library (the way oracle/ref_harness.py stages it for the reference library; nothing is copied into the
repository): it must import, bind every symbol with its own restype / argtypes, and build graphs through
simplegraph / creategraph / from_dataframe / writegraph -- calls that need no GPU.  On the GPU box, where the reference
tree does not exist, the same binding pattern (empty ctypes Structures, POINTER restypes, byref(self), .contents;
interface.py:12-67) is restated here and the known answers of sp/tests/graph_test.cc are asked through it."""
import os
import shutil
import sys
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_IF = "/root/reference/projects/bolinas_civic/sp/interface.py"
LIB = os.path.join(ROOT, "spatial_queue_b200", "libsqb200.so")


@pytest.mark.skipif(not os.path.exists(REF_IF), reason="the reference tree only exists in the build container")
def test_reference_interface_module_binds_the_product_library():
    tmp = tempfile.mkdtemp(prefix="sq_b1_")
    try:
        os.makedirs(os.path.join(tmp, "spb1", "build"))
        shutil.copy(REF_IF, os.path.join(tmp, "spb1", "interface.py"))      # /tmp only
        open(os.path.join(tmp, "spb1", "__init__.py"), "w").close()
        os.symlink(LIB, os.path.join(tmp, "spb1", "build", "liblsp.so"))
        sys.path.insert(0, tmp)
        try:
            import importlib
            iface = importlib.import_module("spb1.interface")
        finally:
            sys.path.remove(tmp)
        import pandas as pd
        g = iface.simplegraph()
        g.update_edge(1, 3, iface.c_double(0.5))
        out = os.path.join(tmp, "g.mtx")
        assert g.writegraph(out.encode())
        rows = [l.split() for l in open(out) if not l.startswith("%")]
        assert rows[0] == ["7", "7", "11"] and ["1", "3", "0.5"] in rows[1:]
        df = pd.DataFrame({"start": [0, 1, 2, 3], "end": [1, 2, 3, 0], "wgh": [0.1, 0.5, 1.9, 1.1]})
        g2 = iface.from_dataframe(df, "start", "end", "wgh")
        assert g2.writegraph(os.path.join(tmp, "g2.mtx").encode())
        assert sum(1 for l in open(os.path.join(tmp, "g2.mtx")) if not l.startswith("%")) == 5
        for name in ("simplegraph", "readgraph", "creategraph", "writegraph", "dijkstra", "update_edge", "origin",
                     "distance", "clear", "parent"):
            assert hasattr(iface.libsp, name)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def _bind_like_the_reference():
    """interface.py:12-51, restated: what a maintainer's unmodified binding does with whatever liblsp.so it finds."""
    from ctypes import POINTER, Structure, byref, c_bool, c_double, c_int, c_int32, cdll
    from numpy.ctypeslib import ndpointer
    libsp = cdll.LoadLibrary(LIB)
    libsp.distance.restype = c_double

    class ShortestPath(Structure):
        @property
        def origin(self):
            return libsp.origin(byref(self))

        def distance(self, destination):
            return libsp.distance(byref(self), destination)

        def parent(self, destination):
            return libsp.parent(byref(self), destination)

        def route(self, destination):
            if destination != self.origin:
                parent = self.parent(destination)
                yield from self.route(parent)
                yield parent, destination

        def clear(self):
            return libsp.clear(byref(self))

    libsp.dijkstra.restype = POINTER(ShortestPath)

    class Graph(Structure):
        def dijkstra(self, origin, destination):
            return libsp.dijkstra(byref(self), origin, destination).contents

        def update_edge(self, origin, destination, weight):
            return libsp.update_edge(byref(self), origin, destination, weight)

    libsp.simplegraph.restype = POINTER(Graph)
    libsp.creategraph.restype = POINTER(Graph)
    libsp.creategraph.argtypes = [ndpointer(c_int32, flags="C_CONTIGUOUS"), ndpointer(c_int32, flags="C_CONTIGUOUS"),
                                  ndpointer(c_double, flags="C_CONTIGUOUS"), c_int, c_int, c_bool]
    return libsp, c_double


@pytest.mark.gpu
def test_known_answers_through_the_reference_style_binding():
    libsp, c_double = _bind_like_the_reference()
    g = libsp.simplegraph(True).contents
    sp = g.dijkstra(1, 3)                                   # graph_test.cc:31
    assert sp.origin == 1 and sp.distance(3) == pytest.approx(7.2, rel=1e-7)
    assert [v for _, v in sp.route(3)] == [2, 4, 3]
    sp.clear()
    g.update_edge(1, 3, c_double(3.7))                      # graph_test.cc:42
    sp = g.dijkstra(1, 3)
    assert sp.distance(3) == pytest.approx(3.7, rel=1e-7)
    sp.clear()
    g = libsp.simplegraph(False).contents                   # graph_test.cc:91
    sp = g.dijkstra(1, 3)
    assert sp.distance(3) == pytest.approx(5.6, rel=1e-7)
    sp.clear()
    z = np.load(os.path.join(ROOT, "tests", "golden", "sp_network.npz"))
    g = libsp.creategraph(z["starts"].astype(np.int32), z["ends"].astype(np.int32), z["weights"].astype(np.float64),
                          len(z["starts"]), int(z["header_n"]), True).contents
    sp = g.dijkstra(1020, 20)                               # graph_test.cc:127
    assert sp.distance(20) == 12409.660000000002
    sp.clear()
