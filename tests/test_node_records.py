"""The node kernel's static opposite-approach lists (NodeRec, tick.cu) against the walk they replace.

run_node_model picks the opposite approach among the in-links that still have a queue as the one whose turn onto the
primary link's reverse is the straightest (queue_class.py:267-293): a walk over the remaining candidates with a strict
`<`, so that ties go to whoever is visited first -- the element the last shuffle moved to the front, then list order; in
priority order at priority nodes.  The device reads a list built at set-up instead ("first of the list that has a
queue") and keeps the walk where its order matters.  Host-only check (no GPU): for every node, every primary in-link and
EVERY set of remaining candidates, list and walk agree."""
import ctypes as C
import itertools

import numpy as np

NR_NONE, NR_SLOW, NR_NODE_SLOW = 1 << 30, 1 << 31, 1 << 29


def _records(flat):
    from spatial_queue_b200._lib import lib
    from spatial_queue_b200.tick import make_desc
    desc, keep = make_desc(flat, seed=0)
    out = np.zeros((len(flat.node_key), 16), np.int32)
    assert lib().sq_build_node_records(C.byref(desc), out.ctypes.data) == 0
    return out[:, :8], out[:, 8:].view(np.uint32)


def _walk(angles, rest, first, node_type, prio):
    """pick_secondary of tick.cu (= the reference's loop): argmin |angle| over `rest` in visiting order, strict <."""
    if node_type == 2:
        order = sorted((j for j in rest), key=lambda j: prio[j])
        seen, visit = set(), []
        for j in order:                       # links that share a rank: only the first of a rank is ever visited
            if prio[j] not in seen:
                seen.add(prio[j]); visit.append(j)
    else:
        visit = ([first] if first in rest else []) + [j for j in sorted(rest) if j != first]
    best, arg = None, -1
    for j in visit:
        a = abs(angles[j])
        if best is None or a < best:
            best, arg = a, j
    return arg


def _check(flat, max_nodes=None):
    ins, sec = _records(flat)
    ip, op = np.asarray(flat.node_in_ptr), np.asarray(flat.node_out_ptr)
    il, ol = np.asarray(flat.node_in_links), np.asarray(flat.node_out_links)
    tp, turn, rev = np.asarray(flat.node_turn_ptr), np.asarray(flat.turn_angle), np.asarray(flat.link_reverse)
    ntype, prio_all = np.asarray(flat.node_type), np.asarray(flat.node_in_prio)
    out_local = {}
    for n in range(len(ip) - 1):
        for j, l in enumerate(ol[op[n]:op[n + 1]]):
            out_local[int(l)] = j
    fast = slow = none = 0
    nodes = range(len(ip) - 1) if max_nodes is None else np.random.default_rng(0).choice(len(ip) - 1, max_nodes, replace=False)
    for n in nodes:
        indeg, outdeg = int(ip[n + 1] - ip[n]), int(op[n + 1] - op[n])
        assert (ins[n, :min(indeg, 8)] == il[ip[n]:ip[n] + min(indeg, 8)]).all() and (ins[n, indeg:] == -1).all()
        if indeg > 8:
            assert sec[n, 0] & NR_NODE_SLOW
            continue
        tb = turn[tp[n]:tp[n + 1]].reshape(indeg, outdeg) if outdeg else np.zeros((indeg, 0))
        prio = prio_all[ip[n]:ip[n + 1]]
        for i in range(indeg):
            w = int(sec[n, i])
            r = int(rev[il[ip[n] + i]])
            if r < 0:
                assert w & NR_NONE; none += 1
                continue
            if w & NR_SLOW:
                slow += 1
                continue
            fast += 1
            cnt = w & 7
            lst = [(w >> (3 + 3 * a)) & 7 for a in range(cnt)]
            others = [j for j in range(indeg) if j != i]
            assert sorted(lst) == others
            angles = tb[:, out_local[r]]
            cands = [j for j in others if ntype[n] != 2 or prio[j] >= 0]     # priority nodes: only listed links are candidates
            for k in range(1, len(cands) + 1):
                for rest in itertools.combinations(cands, k):
                    got = next(j for j in lst if j in rest)
                    for first in (-1,) + rest:
                        assert got == _walk(angles, set(rest), first, int(ntype[n]), prio), (n, i, rest, first)
    return fast, slow, none


def test_lists_equal_the_walk_on_an_irregular_network():
    """Streets removed, one-ways, diagonals, hubs, signal and priority nodes (the stand-in generator of configs 2 / 3)."""
    from oracle import sssp as osssp
    from spatial_queue_b200 import synth

    def router(starts, ends, w, roots):           # the generator routes its demand; on a machine without a GPU, by the C oracle
        og = osssp.OracleGraph(starts, ends, w, max_vertex_id=int(max(starts.max(), ends.max())))
        return og.sssp_batch(np.asarray(roots, np.int32), want_dist=False)[1]

    flat = synth.town_scenario(24, 500, horizon=60, seed=3, router=router)
    fast, slow, none = _check(flat)
    assert fast > 1000 and none > 100


def test_exact_ties_keep_the_walk():
    """A lattice with its turn angles rounded to whole degrees: left and right turns have the same |angle|, the order of
    the walk decides, and the record says so (NR_SLOW) instead of freezing an order; lists without ties still agree."""
    from spatial_queue_b200 import synth
    lon, lat = synth.grid_coords(9, seed=1, jitter=0.0)
    s, e = synth.grid_edges(9)
    _, lanes, maxmph = synth.grid_attributes(9, s, e, lon, lat, 8)
    flat = synth.flat_from_edges(lon, lat, s, e, lanes, maxmph)
    flat.agent_key = np.zeros(0, np.int64)
    for name in ("agent_origin", "agent_destin", "agent_dept", "agent_vlen", "route_links"):
        setattr(flat, name, np.zeros(0, np.int32))
    flat.agent_status = np.zeros(0, np.uint8)
    flat.route_ptr = np.zeros(1, np.int64)
    flat.turn_angle = np.round(np.asarray(flat.turn_angle))       # whole degrees: +90 and -90 tie exactly
    fast, slow, none = _check(flat)
    assert slow > 100 and fast > 0
