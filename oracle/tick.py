"""ctypes binding of the C tick oracle (tick_oracle.c).  Test infrastructure only."""
import ctypes as C

from spatial_queue_b200.tick import _SimBase, make_desc
from . import sssp as _s


def _lib():
    L = _s.lib()
    if not getattr(L, "_tick_declared", False):
        vp, ci = C.c_void_p, C.c_int
        L.oracle_tick_create.restype = vp; L.oracle_tick_create.argtypes = [vp]
        L.oracle_tick_destroy.argtypes = [vp]
        L.oracle_tick_step.argtypes = [vp, ci]
        L.oracle_tick_run.argtypes = [vp, ci, ci]
        L.oracle_tick_set_routes.argtypes = [vp, ci, vp, vp, vp]
        L.oracle_tick_set_link_closed.argtypes = [vp, ci, vp, vp]
        L.oracle_tick_get_link_state.argtypes = [vp] * 7
        L.oracle_tick_get_link_vehicles.argtypes = [vp, vp, vp]
        L.oracle_tick_get_agent_state.argtypes = [vp] * 6
        L.oracle_tick_get_node_state.argtypes = [vp, vp]
        L.oracle_tick_get_link_weights.argtypes = [vp, vp]
        L.oracle_tick_get_counters.argtypes = [vp, vp]
        L.oracle_tick_get_pair_moves.argtypes = [vp, vp]
        L.oracle_tick_set_shelter_after.argtypes = [vp, ci]
        L.oracle_tick_set_congest_after.argtypes = [vp, ci]
        L.oracle_tick_stop_agents.restype = ci; L.oracle_tick_stop_agents.argtypes = [vp, ci, vp, ci, ci]
        L.oracle_tick_get_link_congested.argtypes = [vp, vp]
        L.oracle_tick_get_shelter_state.argtypes = [vp, vp, vp]
        L.oracle_tick_error.restype = C.c_char_p; L.oracle_tick_error.argtypes = [vp]
        L._tick_declared = True
    return L


class OracleSim(_SimBase):
    def __init__(self, flat, seed=0):
        self._L = _lib()
        self.n_nodes, self.n_links, self.n_agents = len(flat.node_key), len(flat.link_end_node), len(flat.agent_key)
        desc, keep = make_desc(flat, seed)
        self._h = self._L.oracle_tick_create(C.byref(desc))

    def _fn(self, name):
        return getattr(self._L, "oracle_tick_" + name)

    def _check(self, rc, what):
        if rc != 0:
            raise RuntimeError(f"oracle_tick_{what} failed ({rc}): {self._L.oracle_tick_error(self._h).decode()}")

    def set_shelter_after(self, seconds):
        self._check(self._L.oracle_tick_set_shelter_after(self._h, int(seconds)), "set_shelter_after")

    def shelter_state(self):
        import numpy as np
        ls, st = np.empty(self.n_links, np.int32), np.empty(max(self.n_agents, 1), np.int32)
        self._check(self._L.oracle_tick_get_shelter_state(self._h, ls.ctypes.data, st.ctypes.data), "get_shelter_state")
        return dict(link_shelter=ls, stop_tick=st[: self.n_agents])

    def sync(self):
        pass

    def close(self):
        if self._h:
            self._L.oracle_tick_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()
