"""sp_sssp_batch_device_noalloc (include/sq_sp.h): the allocation-free entry point against sp_sssp_batch_device, through
the C ABI.  Collected last on purpose (pytest -x at the end of a round: a failure here must not hide other tests)."""
import ctypes as C

import numpy as np
import pytest

from conftest import grid_graph

pytestmark = pytest.mark.gpu


def _graph(spi, n, seed):
    s, e, w = grid_graph(n, seed=seed)
    V = n * n
    return spi.from_arrays(s, e, w, max_vertex_id=V - 1), V


def test_noalloc_entry_point_matches_the_existing_one_twice_in_a_row():
    import torch
    from spatial_queue_b200 import sp_interface as spi
    from spatial_queue_b200._lib import lib
    rng = np.random.default_rng(9)
    V, E = 30000, 150000                              # an expander: wavefronts of many thousand vertices, so the queues
    s = rng.integers(0, V, E).astype(np.int32)        # outgrow their 2,048 shared-memory slots and use the static scratch
    e = rng.integers(0, V, E).astype(np.int32)
    w = rng.random(E) * 30.0 + 0.5
    g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
    g.sync()
    st = torch.cuda.current_stream().cuda_stream
    free_before = None
    for call in range(2):                             # the second call finds the first one's leftovers in the scratch
        n = 40 if call == 0 else 23
        origins = torch.from_numpy(rng.integers(0, V, n).astype(np.int32)).cuda()
        d_ref = torch.empty((n, V), dtype=torch.float64, device="cuda")
        p_ref = torch.empty((n, V), dtype=torch.int32, device="cuda")
        d_new = torch.full((n, V), -1.0, dtype=torch.float64, device="cuda")
        p_new = torch.full((n, V), -7, dtype=torch.int32, device="cuda")
        g.sssp_batch_device(origins, d_ref, p_ref)
        torch.cuda.synchronize()
        if call == 1:
            free_before = torch.cuda.mem_get_info()[0]
        rc = lib().sp_sssp_batch_device_noalloc(g._h, C.c_void_p(origins.data_ptr()), n, C.c_void_p(d_new.data_ptr()),
                                                C.c_void_p(p_new.data_ptr()), C.c_void_p(st))
        assert rc == 0, lib().sp_last_error().decode()
        torch.cuda.synchronize()
        if call == 1:                                 # the device's free memory did not move across the call
            assert torch.cuda.mem_get_info()[0] == free_before
        # the tolerance of the SSSP parity tests: bit-exact distances, identical canonical predecessors
        assert torch.equal(d_new.view(torch.int64), d_ref.view(torch.int64))
        assert torch.equal(p_new, p_ref)
    # distances only
    d_only = torch.empty((n, V), dtype=torch.float64, device="cuda")
    g.sssp_batch_device_noalloc(origins, d_only)
    torch.cuda.synchronize()
    assert torch.equal(d_only.view(torch.int64), d_ref.view(torch.int64))
    g.free()


def test_noalloc_entry_point_states_its_limits():
    import torch
    from spatial_queue_b200 import sp_interface as spi
    from spatial_queue_b200._lib import lib
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    o = torch.zeros(1, dtype=torch.int32, device="cuda")
    # beyond 65,536 vertices: a clear error, nothing launched
    g, V = _graph(spi, 260, seed=1)                   # 67,600 vertices
    g.sync()
    d = torch.empty((1, V), dtype=torch.float64, device="cuda")
    rc = lib().sp_sssp_batch_device_noalloc(g._h, C.c_void_p(o.data_ptr()), 1, C.c_void_p(d.data_ptr()), None, st)
    assert rc != 0 and b"65536" in lib().sp_last_error()
    g.free()
    # edits not yet on the device: the call refuses instead of allocating
    g, V = _graph(spi, 20, seed=2)
    d = torch.empty((1, V), dtype=torch.float64, device="cuda")
    rc = lib().sp_sssp_batch_device_noalloc(g._h, C.c_void_p(o.data_ptr()), 1, C.c_void_p(d.data_ptr()), None, st)
    assert rc != 0 and b"sp_sync" in lib().sp_last_error()
    g.sync()
    rc = lib().sp_sssp_batch_device_noalloc(g._h, C.c_void_p(o.data_ptr()), 1, C.c_void_p(d.data_ptr()), None, st)
    assert rc == 0, lib().sp_last_error().decode()
    rc = lib().sp_sssp_batch_device_noalloc(g._h, C.c_void_p(o.data_ptr()), 1, None, None, st)
    assert rc != 0
    torch.cuda.synchronize()
    g.free()
