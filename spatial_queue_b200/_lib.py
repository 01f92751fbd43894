"""ctypes loader of the product library.  No fallback: a missing library or GPU is an error."""
import ctypes as C
import os

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libsqb200.so")
_lib = None


class B200LibraryError(RuntimeError):
    pass


def lib():
    """Load libsqb200.so (built by spatial_queue_b200.build / __graft_entry__.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200LibraryError(
                f"{LIB_PATH} is missing: run `python -m spatial_queue_b200.build` (needs nvcc). "
                "There is no CPU fallback for the tick or the SSSP path.")
        L = C.CDLL(LIB_PATH)
        _declare(L)
        _lib = L
    return _lib


def _declare(L):
    vp, ci, cd, cb = C.c_void_p, C.c_int, C.c_double, C.c_bool
    # ---- sp surface (include/sq_sp.h part 1) ----
    L.simplegraph.restype = vp; L.simplegraph.argtypes = [cb]
    L.readgraph.restype = vp; L.readgraph.argtypes = [C.c_char_p, cb]
    L.creategraph.restype = vp; L.creategraph.argtypes = [vp, vp, vp, ci, ci, cb]
    L.writegraph.restype = cb; L.writegraph.argtypes = [vp, C.c_char_p]
    L.dijkstra.restype = vp; L.dijkstra.argtypes = [vp, ci, ci]
    L.update_edge.restype = None; L.update_edge.argtypes = [vp, ci, ci, cd]
    L.origin.restype = ci; L.origin.argtypes = [vp]
    L.distance.restype = cd; L.distance.argtypes = [vp, ci]
    L.clear.restype = None; L.clear.argtypes = [vp]
    L.parent.restype = ci; L.parent.argtypes = [vp, ci]
    # ---- additive ----
    L.sp_remove_edge.restype = ci; L.sp_remove_edge.argtypes = [vp, ci, ci]
    L.sp_free_graph.restype = None; L.sp_free_graph.argtypes = [vp]
    L.sp_nvertices.restype = ci; L.sp_nvertices.argtypes = [vp]
    L.sp_nedges.restype = ci; L.sp_nedges.argtypes = [vp]
    L.sp_nedges_logged.restype = ci; L.sp_nedges_logged.argtypes = [vp]
    L.sp_sssp_batch.restype = ci; L.sp_sssp_batch.argtypes = [vp, vp, ci, vp, vp]
    L.sp_sssp_batch_ex.restype = ci; L.sp_sssp_batch_ex.argtypes = [vp, vp, ci, vp, vp, ci]
    L.sp_sssp_batch_codes.restype = ci; L.sp_sssp_batch_codes.argtypes = [vp, vp, ci, vp, vp, ci]
    L.sp_decode_parents_host.restype = ci; L.sp_decode_parents_host.argtypes = [vp, vp, ci, vp, ci]
    L.sp_sssp_batch_codes4.restype = ci; L.sp_sssp_batch_codes4.argtypes = [vp, vp, ci, vp, vp, ci]
    L.sp_decode_parents_host4.restype = ci; L.sp_decode_parents_host4.argtypes = [vp, vp, ci, vp, ci]
    L.sp_route_batch.restype = ci; L.sp_route_batch.argtypes = [vp, vp, vp, ci, vp, vp, vp]
    L.sp_route_pool.restype = vp; L.sp_route_pool.argtypes = [vp]
    L.sp_sssp_batch_device.restype = ci; L.sp_sssp_batch_device.argtypes = [vp, vp, ci, vp, vp, vp]
    L.sp_sssp_batch_device_noalloc.restype = ci; L.sp_sssp_batch_device_noalloc.argtypes = [vp, vp, ci, vp, vp, vp]
    L.sp_sssp_batch_device_ex.restype = ci; L.sp_sssp_batch_device_ex.argtypes = [vp, vp, ci, vp, vp, vp, ci]
    L.sp_set_weights_device.restype = ci; L.sp_set_weights_device.argtypes = [vp, vp, ci, vp]
    L.sp_set_weights.restype = ci; L.sp_set_weights.argtypes = [vp, vp, ci]
    L.sp_sync.restype = ci; L.sp_sync.argtypes = [vp]
    L.sp_has_pending_edits.restype = ci; L.sp_has_pending_edits.argtypes = [vp]
    L.sp_set_tuning.restype = ci; L.sp_set_tuning.argtypes = [vp, ci, ci, cd, ci]
    L.sp_alloc_host.restype = vp; L.sp_alloc_host.argtypes = [C.c_size_t]
    L.sp_free_host.restype = None; L.sp_free_host.argtypes = [vp]
    L.sp_set_parent_format.restype = ci; L.sp_set_parent_format.argtypes = [vp, ci]
    L.sp_decode_parents_device.restype = ci; L.sp_decode_parents_device.argtypes = [vp, vp, ci, vp, vp, ci]
    L.sp_set_algo.restype = ci; L.sp_set_algo.argtypes = [vp, ci, cd]
    L.sp_set_warp_tuning.restype = ci; L.sp_set_warp_tuning.argtypes = [vp, ci, ci, ci]
    L.sp_set_group_tuning.restype = ci; L.sp_set_group_tuning.argtypes = [vp, ci, ci, ci, ci, ci]
    L.sp_set_group_mode.restype = ci; L.sp_set_group_mode.argtypes = [vp, ci, ci]
    L.sp_set_smem_rows.restype = ci; L.sp_set_smem_rows.argtypes = [vp, ci]
    L.sp_last_stats.restype = ci; L.sp_last_stats.argtypes = [vp, vp]
    L.sp_last_error.restype = C.c_char_p; L.sp_last_error.argtypes = []
    L.sp_kernel_launches.restype = C.c_longlong; L.sp_kernel_launches.argtypes = []


    # ---- tick surface (include/sq_tick.h) ----
    i32, i64 = C.c_int32, C.c_int64
    L.sq_create.restype = vp; L.sq_create.argtypes = [vp]
    L.sq_destroy.restype = None; L.sq_destroy.argtypes = [vp]
    L.sq_step.restype = ci; L.sq_step.argtypes = [vp, i32]
    L.sq_run.restype = ci; L.sq_run.argtypes = [vp, i32, i32]
    L.sq_sync.restype = ci; L.sq_sync.argtypes = [vp]
    L.sq_set_stream.restype = ci; L.sq_set_stream.argtypes = [vp, vp]
    L.sq_set_routes.restype = ci; L.sq_set_routes.argtypes = [vp, i32, vp, vp, vp]
    L.sq_attach_router.restype = ci; L.sq_attach_router.argtypes = [vp, vp]
    L.sq_set_shelter_after.restype = ci; L.sq_set_shelter_after.argtypes = [vp, i32]
    L.sq_set_congest_after.restype = ci; L.sq_set_congest_after.argtypes = [vp, i32]
    L.sq_stop_agents.restype = ci; L.sq_stop_agents.argtypes = [vp, i32, vp, i32, i32]
    L.sq_get_link_congested.restype = ci; L.sq_get_link_congested.argtypes = [vp, vp]
    L.sq_get_summary.restype = ci; L.sq_get_summary.argtypes = [vp, vp]
    L.sq_get_shelter_state.restype = ci; L.sq_get_shelter_state.argtypes = [vp, vp, vp]
    L.sq_set_link_closed.restype = ci; L.sq_set_link_closed.argtypes = [vp, i32, vp, vp]
    L.sq_set_closure_schedule.restype = ci; L.sq_set_closure_schedule.argtypes = [vp, i32, vp, vp, vp]
    L.sq_get_link_state.restype = ci; L.sq_get_link_state.argtypes = [vp] * 7
    L.sq_get_link_vehicles.restype = ci; L.sq_get_link_vehicles.argtypes = [vp, vp, vp]
    L.sq_get_agent_state.restype = ci; L.sq_get_agent_state.argtypes = [vp] * 6
    L.sq_get_node_state.restype = ci; L.sq_get_node_state.argtypes = [vp, vp]
    L.sq_get_link_weights.restype = ci; L.sq_get_link_weights.argtypes = [vp, vp]
    L.sq_set_reroute_rules.restype = ci; L.sq_set_reroute_rules.argtypes = [vp, cd, i32]
    L.sq_enable_pair_moves.restype = ci; L.sq_enable_pair_moves.argtypes = [vp]
    L.sq_get_pair_moves.restype = ci; L.sq_get_pair_moves.argtypes = [vp, vp]
    L.sq_enable_timeline.restype = ci; L.sq_enable_timeline.argtypes = [vp, i32, i32]
    L.sq_get_timeline.restype = ci; L.sq_get_timeline.argtypes = [vp, vp, vp, vp]
    L.sq_build_node_records.restype = ci; L.sq_build_node_records.argtypes = [vp, vp]
    L.sq_get_counters.restype = ci; L.sq_get_counters.argtypes = [vp, vp]
    L.sq_last_error.restype = C.c_char_p; L.sq_last_error.argtypes = []
    L.sq_kernel_launches.restype = C.c_longlong; L.sq_kernel_launches.argtypes = []


def check(rc, what="call"):
    if rc != 0:
        raise B200LibraryError(f"{what} failed ({rc}): {lib().sp_last_error().decode()}")
