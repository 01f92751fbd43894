"""GPU parity tests for HP-2: the CUDA tree kernel, called through the C ABI, against the oracle."""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import grid_graph

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
DBL_MAX = np.finfo(np.float64).max


@pytest.fixture(scope="module")
def spi():
    from spatial_queue_b200 import sp_interface
    return sp_interface


def test_known_answers_through_the_abi(spi):
    # sp/tests/graph_test.cc:31-64 (directed) and :91-103 (undirected), replayed through the C ABI
    g = spi.simplegraph(True)
    sp = g.dijkstra(1, 3)
    assert sp.origin == 1
    assert sp.distance(3) == pytest.approx(7.2, rel=1e-7)
    assert [h[1] for h in sp.route(3)] == [2, 4, 3]
    sp.clear()
    g.update_edge(1, 3, 3.7)
    sp = g.dijkstra(1, 3); assert sp.distance(3) == pytest.approx(3.7, rel=1e-7); sp.clear()
    g = spi.simplegraph(True)
    g.remove_edge(3, 1)
    sp = g.dijkstra(1, 3); assert sp.distance(3) == pytest.approx(7.2, rel=1e-7); sp.clear()
    g.remove_edge(2, 4)
    sp = g.dijkstra(1, 3); assert sp.distance(3) == pytest.approx(9.1, rel=1e-7); sp.clear()
    g = spi.simplegraph(False)
    sp = g.dijkstra(1, 3); assert sp.distance(3) == pytest.approx(5.6, rel=1e-7); sp.clear()
    g.remove_edge(3, 1)
    sp = g.dijkstra(1, 3); assert sp.distance(3) == pytest.approx(9.1, rel=1e-7); sp.clear()
    # bad ids: empty tree, distance() = DBL_MAX (py.cc:74-77)
    sp = g.dijkstra(99, -1); assert sp.distance(3) == DBL_MAX; sp.clear()
    # missing edge: ignored, not an abort
    g.update_edge(5, 1, 1.0)


def test_network_mtx_golden(spi):
    gold = json.load(open(os.path.join(GOLD, "sp_golden.json")))
    z = np.load(os.path.join(GOLD, "sp_network.npz"))
    g = spi.from_arrays(z["starts"], z["ends"], z["weights"], max_vertex_id=int(z["header_n"]))
    assert g.nvertices == gold["network_mtx"]["nvertices"]
    sp = g.dijkstra(1020, 20)
    assert sp.distance(20) == 12409.660000000002  # graph_test.cc:127
    assert float(sp.distance(20)).hex() == gold["one_to_one_1020_20_hex"]
    sp.clear()
    origins = [t["origin"] for t in gold["trees"]]
    d, p = g.sssp_batch(origins)
    for i, t in enumerate(gold["trees"]):
        assert hashlib.sha256(d[i].tobytes()).hexdigest() == t["dist_sha256"]
        assert hashlib.sha256(p[i].tobytes()).hexdigest() == t["canonical_parent_sha256"]
        assert int((d[i] < 1e300).sum()) == t["reached"]


@pytest.mark.parametrize("directed", [True, False])
@pytest.mark.parametrize("threads,smemq,algo,qfac", [(256, 0, 1, 0), (128, 32, 1, 0), (1024, 64, 1, 0),   # bitmap kernel
                                                      (128, 0, 0, 0), (64, 0, 0, 0), (32, 0, 0, 0), (1024, 0, 0, 0), (512, 0, 0, 0),   # ELL kernel
                                                      (256, 0, 0, 4.0), (64, 0, 0, 0.02),                 # roomy / tight queues
                                                      (1, 1024, 2, 0), (2, 64, 2, 0), (4, 256, 2, 0),    # warp kernel: (ilp, ring)
                                                      (8, 0, 4, 0), (16, 1, 4, 0), (32, 1, 4, 0),         # group kernel: (batch, prefetch)
                                                      (2, 64, 5, 0), (4, 1024, 5, 0),                     # grouped warp kernel: (ilp, ring)
                                                      (4, 0, 6, 0), (8, 0, 6, 0), (16, 0, 6, 0),          # union kernel: pops in flight
                                                      (256, 0, 3, 0)])    # auto: a handful of trees -> distance rows in shared memory
def test_random_graphs_bit_exact(spi, directed, threads, smemq, algo, qfac):
    from oracle import sssp as osssp
    rng = np.random.default_rng(11 + threads)
    for trial in range(5):
        V = int(rng.integers(3, 3000))
        E = int(rng.integers(1, 6 * V))
        s = rng.integers(0, V, E).astype(np.int32)
        e = rng.integers(0, V, E).astype(np.int32)
        w = np.where(rng.random(E) < 0.05, 1e8, rng.random(E) * 100.0)
        w = np.where(rng.random(E) < 0.3, np.round(w) + 1.0, w)  # ties, strictly positive
        og = osssp.OracleGraph(s, e, w, max_vertex_id=V - 1, directed=directed)
        g = spi.from_arrays(s, e, w, max_vertex_id=V - 1, directed=directed)
        g.set_algo(min(algo, 4), qfac)
        if algo == 2:
            g.set_warp_tuning(ring_cap=smemq, ilp=threads)
            g.set_tuning(delta_factor=float(rng.choice([0.25, 2.0, 50.0])))
        elif algo == 4:
            # parallel edges, self loops, out-degrees beyond four (sequential path), closed links (threshold jumps)
            g.set_group_mode(1)
            g.set_group_tuning(batch=threads, prefetch=smemq)
            g.set_tuning(delta_factor=float(rng.choice([0.25, 2.0, 50.0])))
        elif algo == 5:
            # the warp kernel on the group layout; the 64-entry ring overflows on these graphs (bitmap-kernel rerun)
            g.set_algo(4)
            g.set_group_mode(0, smemq)
            g.set_warp_tuning(ilp=threads)
            g.set_tuning(delta_factor=float(rng.choice([0.25, 2.0, 50.0])))
        elif algo == 6:
            # shared frontier + atomics: parallel edges, self loops, overflow edges, closed links
            g.set_algo(4)
            g.set_group_mode(2)
            g.set_group_tuning(batch=threads)
            g.set_tuning(delta_factor=float(rng.choice([0.25, 2.0, 50.0])))
        else:
            g.set_tuning(threads=threads, smem_queue=smemq, delta_factor=float(rng.choice([0.25, 2.0, 50.0])))
        origins = np.concatenate([rng.integers(0, V, 7 if algo < 4 else 29), [-1, V, V + 5]]).astype(np.int32)
        d, p = g.sssp_batch(origins)
        do, po = og.sssp_batch(origins)
        assert d.tobytes() == do.tobytes()
        assert (p == po).all()
        g.free()


def test_queue_overflow_falls_back_to_the_bitmap_kernel(spi):
    """A dense random graph with tiny queue capacities: the ELL kernel gives up on (some) trees and the bitmap
    kernel reruns them; results stay bit-exact."""
    from oracle import sssp as osssp
    rng = np.random.default_rng(5)
    V, E = 20000, 160000
    s = rng.integers(0, V, E).astype(np.int32)
    e = rng.integers(0, V, E).astype(np.int32)
    w = rng.random(E) * 10.0 + 0.01
    g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
    g.set_algo(0, 0.0016)      # 384 shared + 32 global entries per queue
    origins = rng.integers(0, V, 24).astype(np.int32)
    d, p = g.sssp_batch(origins)
    assert g.last_stats()["overflows"] > 0
    do, po = osssp.OracleGraph(s, e, w, max_vertex_id=V - 1).sssp_batch(origins)
    assert d.tobytes() == do.tobytes() and (p == po).all()
    g.free()


def test_warp_kernel_ring_overflow_falls_back(spi):
    """Same, for the warp kernel with a 64-entry ring: wavefronts of a dense random graph do not fit, the trees go
    to the fallback kernel; a grid with a roomy ring must not overflow at all."""
    from oracle import sssp as osssp
    rng = np.random.default_rng(6)
    V, E = 20000, 160000
    s = rng.integers(0, V, E).astype(np.int32)
    e = rng.integers(0, V, E).astype(np.int32)
    w = rng.random(E) * 10.0 + 0.01
    g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
    g.set_algo(2)
    g.set_warp_tuning(ring_cap=64, ilp=2)
    origins = rng.integers(0, V, 24).astype(np.int32)
    d, p = g.sssp_batch(origins)
    assert g.last_stats()["overflows"] > 0
    do, po = osssp.OracleGraph(s, e, w, max_vertex_id=V - 1).sssp_batch(origins)
    assert d.tobytes() == do.tobytes() and (p == po).all()
    g.free()
    s, e, w = grid_graph(100, seed=3)
    g = spi.from_arrays(s, e, w)
    g.set_algo(2)
    origins = rng.integers(0, 100 * 100, 64).astype(np.int32)
    d, p = g.sssp_batch(origins)
    assert g.last_stats()["overflows"] == 0
    do, po = osssp.OracleGraph(s, e, w).sssp_batch(origins)
    assert d.tobytes() == do.tobytes() and (p == po).all()
    g.free()


def test_weight_updates_reach_the_device(spi):
    from oracle import sssp as osssp
    s, e, w = grid_graph(24, seed=5)
    g = spi.from_arrays(s, e, w)
    rng = np.random.default_rng(0)
    w2 = w.copy()
    for _ in range(3):
        idx = rng.choice(len(w), 200, replace=False)
        w2[idx] = np.where(rng.random(200) < 0.2, 1e8, w[idx] * (1 + 5 * rng.random(200)))
        for i in idx[:50]:
            g.update_edge(int(s[i]), int(e[i]), float(w2[i]))
        if len(idx) > 50:
            g.set_weights(w2)
        og = osssp.OracleGraph(s, e, w2)
        origins = rng.integers(0, 24 * 24, 5).astype(np.int32)
        d, p = g.sssp_batch(origins)
        do, po = og.sssp_batch(origins)
        assert d.tobytes() == do.tobytes() and (p == po).all()


@pytest.mark.parametrize("algo", [0, 1, 2, 3, 4, 5, 6])
def test_grid_batch_and_device_api(spi, algo):
    import torch
    from oracle import sssp as osssp
    n = 64
    s, e, w = grid_graph(n, seed=1)
    g = spi.from_arrays(s, e, w)
    g.set_algo(min(algo, 4))
    g.set_group_mode({4: 1, 6: 2}.get(algo, 0))
    og = osssp.OracleGraph(s, e, w)
    # 2000 origins: above the auto switch-over (13 trees per SM), so algo 3 runs the warp kernel here
    origins = np.random.default_rng(2).permutation(n * n)[:2000].astype(np.int32)
    o_dev = torch.from_numpy(origins).cuda()
    d_dev = torch.empty((len(origins), n * n), dtype=torch.float64, device="cuda")
    p_dev = torch.empty((len(origins), n * n), dtype=torch.int32, device="cuda")
    g.sssp_batch_device(o_dev, d_dev, p_dev)
    torch.cuda.synchronize()
    do, po = og.sssp_batch(origins)
    assert d_dev.cpu().numpy().tobytes() == do.tobytes()
    assert (p_dev.cpu().numpy() == po).all()


def test_full_size_properties(spi):
    """BASELINE config 5 size (224x224 grid): fixed-point and tree properties that need no oracle."""
    n = 224
    s, e, w = grid_graph(n, seed=1)
    g = spi.from_arrays(s, e, w)
    origins = np.array([0, n * n - 1, (n // 2) * n + n // 2, 12345], dtype=np.int32)
    d, p = g.sssp_batch(origins)
    for i, o in enumerate(origins):
        di = d[i]
        assert di[o] == 0.0 and p[i][o] == -1
        assert (di < 1e300).all()                       # grid is strongly connected
        assert (di[e] <= di[s] + w).all()               # no edge can still relax
        pv = p[i]
        v = np.arange(n * n)
        m = v != o
        # every predecessor is an in-neighbour that attains the distance exactly
        key = {(int(a), int(b)): float(ww) for a, b, ww in zip(s, e, w)}
        idx = np.random.default_rng(1).choice(v[m], 2000, replace=False)
        for vv in idx:
            u = int(pv[vv])
            assert di[u] + key[(u, int(vv))] == di[vv]
        # sum of distances symmetric check: d(o -> x) on the reversed graph equals d(x -> o)
    # reversed graph trees from x give d(x->o): compare one pair exactly through a second graph
    gr = spi.from_arrays(e, s, w)
    dr, _ = gr.sssp_batch(origins[:1])
    # path costs are summed in opposite orders, so equality is to rounding only
    assert np.allclose(dr[0][origins[1]], d[1][origins[0]], rtol=1e-12)


def test_reverse_trees_and_device_weight_refresh(spi):
    """reverse=True == the oracle on the transposed edge list (bit-exact); sp_set_weights_device == sp_set_weights,
    with NaN entries leaving weights alone, on a graph with vertices of in/out degree > 4 (overflow records)."""
    import torch
    from oracle import sssp as osssp
    rng = np.random.default_rng(21)
    V, E = 3000, 14000
    s = rng.integers(0, V, E).astype(np.int32)
    e = rng.integers(0, V, E).astype(np.int32)
    w = rng.random(E) * 50.0 + 0.5
    g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
    dests = rng.integers(0, V, 9).astype(np.int32)
    o_dev = torch.from_numpy(dests).cuda()
    d_dev = torch.empty((len(dests), V), dtype=torch.float64, device="cuda")
    p_dev = torch.empty((len(dests), V), dtype=torch.int32, device="cuda")

    def check(weights):
        g.sssp_batch_device(o_dev, d_dev, p_dev, reverse=True)
        torch.cuda.synchronize()
        do, po = osssp.OracleGraph(e, s, weights, max_vertex_id=V - 1).sssp_batch(dests)
        assert d_dev.cpu().numpy().tobytes() == do.tobytes()
        assert (p_dev.cpu().numpy() == po).all()
        d, p = g.sssp_batch(dests[:3])
        do, po = osssp.OracleGraph(s, e, weights, max_vertex_id=V - 1).sssp_batch(dests[:3])
        assert d.tobytes() == do.tobytes() and (p == po).all()

    check(w)
    w2 = w.copy()
    upd = rng.random(E) < 0.4
    w2[upd] = np.where(rng.random(int(upd.sum())) < 0.1, 1e8, rng.random(int(upd.sum())) * 80.0 + 0.1)
    wd = torch.from_numpy(np.where(upd, w2, np.nan)).cuda()
    g.set_weights_device(wd)
    check(w2)
    # a later host-side edit starts from the refreshed weights, not from stale ones
    g.update_edge(int(s[0]), int(e[0]), 3.25)
    last = np.nonzero((s == s[0]) & (e == e[0]))[0][-1]
    w2[last] = 3.25
    check(w2)
    # two device refreshes in a row (the host log is brought up to date lazily), then a structural edit: the rebuild
    # starts from the newest weights
    for k in range(2):
        upd = rng.random(E) < 0.3
        w2[upd] = rng.random(int(upd.sum())) * 60.0 + 0.2
        g.set_weights_device(torch.from_numpy(np.where(upd, w2, np.nan)).cuda())
    check(w2)
    a, b = int(s[5]), int(e[5])
    g.remove_edge(a, b)
    keep = ~((s == a) & (e == b))
    keep[np.nonzero(~keep)[0][:-1]] = True      # the last edge logged under the key is the one removed
    g.sssp_batch_device(o_dev, d_dev, p_dev, reverse=True)
    torch.cuda.synchronize()
    do, po = osssp.OracleGraph(e[keep], s[keep], w2[keep], max_vertex_id=V - 1).sssp_batch(dests)
    assert d_dev.cpu().numpy().tobytes() == do.tobytes()
    g.free()


@pytest.mark.parametrize("algo", [3, 4, 5, 6])
@pytest.mark.parametrize("reverse", [False, True])
def test_one_byte_predecessor_codes_decode_to_the_same_trees(spi, reverse, algo):
    """The compact table the multi-GPU path all-gathers (one byte per vertex) expands to exactly the int32 tree."""
    import torch
    rng = np.random.default_rng(31)
    V, E = 2500, 13000                      # in-degrees beyond 4: overflow codes are exercised
    s = rng.integers(0, V, E).astype(np.int32)
    e = rng.integers(0, V, E).astype(np.int32)
    w = np.round(rng.random(E) * 20.0) + 1.0   # many exact ties
    g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
    g.set_algo(min(algo, 4))
    g.set_group_mode({4: 1, 6: 2}.get(algo, 0), 64 if algo == 5 else 0)   # 5: tiny ring, some trees rerun by the bitmap kernel
    origins = torch.from_numpy(rng.integers(0, V, 40).astype(np.int32)).cuda()
    d = torch.empty((40, V), dtype=torch.float64, device="cuda")
    p32 = torch.empty((40, V), dtype=torch.int32, device="cuda")
    g.sssp_batch_device(origins, d, p32, reverse=reverse)
    g.set_parent_format(1)
    codes = torch.empty((40, V), dtype=torch.uint8, device="cuda")
    g.sssp_batch_device(origins, d, codes, reverse=reverse)
    out = torch.empty((40, V), dtype=torch.int32, device="cuda")
    g.decode_parents_device(codes, out, reverse=reverse)
    torch.cuda.synchronize()
    assert torch.equal(out, p32)
    g.set_parent_format(0)
    dh, ph = g.sssp_batch(origins.cpu().numpy()[:5], reverse=reverse)
    assert (ph == p32[:5].cpu().numpy()).all()
    if algo >= 4:
        # and the group path's tables are the other kernels' tables
        g.set_algo(0)
        d0 = torch.empty_like(d)
        p0 = torch.empty_like(p32)
        g.sssp_batch_device(origins, d0, p0, reverse=reverse)
        torch.cuda.synchronize()
        assert torch.equal(d0.view(torch.int64), d.view(torch.int64)) and torch.equal(p0, p32)
    g.free()


def test_auto_mode_learns_from_ring_overflows(spi):
    """Auto kernel choice on a graph whose wavefronts outgrow the warp kernel's ring (a dense random graph, 2000
    origins): the first batch hands trees to the fallback kernel, the following ones run with a larger ring or on the
    ELL kernel; every batch is bit-exact."""
    from oracle import sssp as osssp
    rng = np.random.default_rng(8)
    V, E = 6000, 60000
    s = rng.integers(0, V, E).astype(np.int32)
    e = rng.integers(0, V, E).astype(np.int32)
    w = rng.random(E) * 10.0 + 0.01
    g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
    g.set_warp_tuning(ring_cap=64)
    og = osssp.OracleGraph(s, e, w, max_vertex_id=V - 1)
    origins = rng.integers(0, V, 2000).astype(np.int32)
    do, po = og.sssp_batch(origins)
    overflows = []
    for _ in range(4):
        d, p = g.sssp_batch(origins)
        overflows.append(g.last_stats()["overflows"])
        assert d.tobytes() == do.tobytes() and (p == po).all()
    assert overflows[0] > 50 and overflows[-1] < overflows[0]
    g.free()


def _walk(par, origin, dest):
    path = [dest]
    while path[-1] != origin:
        path.append(int(par[path[-1]]))
    return path[::-1]


def test_route_batch_equals_get_path_of_the_oracle(spi):
    """sp_route_batch (Agent.get_path for a batch: distance(destin) + route(destin), queue_class.py:491-512) against
    the oracle's Dijkstra with the reference's early exit at the destination: distances bit for bit, paths equal to
    the walk over the canonical predecessors, path cost re-summed forward == distance exactly; repeated origins,
    origin == destination, unreachable pairs, an origin outside the graph; a call in several chunks of trees."""
    from oracle import sssp as osssp
    rng = np.random.default_rng(5)
    for V, E, n in ((40, 60, 200), (2500, 9000, 3000)):
        s = rng.integers(0, V, E).astype(np.int32)
        e = rng.integers(0, V, E).astype(np.int32)
        w = np.where(rng.random(E) < 0.3, np.round(rng.random(E) * 20) + 1.0, rng.random(E) * 100.0)
        og = osssp.OracleGraph(s, e, w, max_vertex_id=V - 1)
        g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
        o = rng.integers(0, V, n).astype(np.int32)
        o[: n // 4] = o[0]                                # many agents at the same node
        d = rng.integers(0, V, n).astype(np.int32)
        d[5] = o[5]
        o[7] = V + 3
        dist, off, ln, pool = g.route_batch(o, d)
        wmin = {}
        for a, b, c in zip(s, e, w):
            wmin[(int(a), int(b))] = min(c, wmin.get((int(a), int(b)), np.inf))
        trees = {}
        reached = 0
        for k in range(n):
            if k == 7:
                assert dist[k] > 1e300 and ln[k] == 0
                continue
            do, _ = og.dijkstra(int(o[k]), int(d[k]))      # early exit at the destination, like the reference
            assert dist[k].tobytes() == do[d[k]].tobytes(), k
            path = pool[off[k]: off[k] + ln[k]].tolist()
            if do[d[k]] > 1e300 or o[k] == d[k]:
                assert path == []
                continue
            reached += 1
            if int(o[k]) not in trees:
                full, _ = og.dijkstra(int(o[k]))
                trees[int(o[k])] = og.canonical_parents(int(o[k]), full)
            assert path == _walk(trees[int(o[k])], int(o[k]), int(d[k])), k
            acc = 0.0
            for a, b in zip(path[:-1], path[1:]):
                acc = acc + wmin[(a, b)]
            assert acc == dist[k]
        assert reached > n // 5
        g.free()


@pytest.mark.parametrize("chunk", [None, 97])
def test_route_batch_on_the_grid_matches_full_tables(spi, chunk, monkeypatch):
    if chunk:
        monkeypatch.setenv("SQB200_ROUTE_CHUNK", str(chunk))      # 1500 trees in 16 chunks
    s, e, w = grid_graph(64, seed=3)
    V = 64 * 64
    g = spi.from_arrays(s, e, w)
    rng = np.random.default_rng(1)
    o = rng.permutation(V)[:1500].astype(np.int32)
    d = rng.integers(0, V, 1500).astype(np.int32)
    dist, off, ln, pool = g.route_batch(o, d)
    dd, pp = g.sssp_batch(o)
    assert dist.tobytes() == dd[np.arange(1500), d].tobytes()
    for k in range(0, 1500, 37):
        assert pool[off[k]: off[k] + ln[k]].tolist() == (_walk(pp[k], int(o[k]), int(d[k])) if o[k] != d[k] else [])
    assert int(ln.sum()) == len(pool) and ln.max() > 60
    g.free()


# ---- the benchmarked configuration, at its size (VERDICT r1 parity hole 1a) -------------------------------------------

def _config5():
    import bench
    s, e, w, V = bench.sssp_workload(224)
    return np.asarray(s, np.int32), np.asarray(e, np.int32), np.asarray(w, np.float64), V, bench


def _congested_weights(s, e, w, seed=2):
    """second weight set, w = q / (cap + 0.1) * 3600 * 25 / maxmph + fft with q ~ Poisson(3)
    (queue_class.py:463-466 on one-lane 1900 veh/h links at 25 mph)."""
    q = np.random.default_rng(seed).poisson(3.0, len(w)).astype(np.float64)
    return q / (1900.0 + 0.1) * 3600 * 25 / 25.0 + w


@pytest.mark.parametrize("algo,n", [(3, 2368), (4, 2368), (5, 2368), (6, 2368), (3, 18944), (6, 18944)])
def test_config5_batches_bit_exact_against_the_oracle(spi, algo, n):
    """bench.py's HP-2 workload itself: 224x224 grid, bench.origins_for origins, through sp_sssp_batch_device in auto
    mode (sssp_warp_kernel, at 2,368 origins and at the benchmarked 18,944) and with each kernel of the group path
    forced (4: sssp_group_kernel, 5: the warp kernel on the group layout, 6: sssp_union_kernel); distances bit for bit and canonical predecessors equal to the C oracle for every tree of the small
    batch and for a 1,024-tree sample of the large one; eight trees also against the reference library itself."""
    import torch
    from oracle import sssp as osssp
    s, e, w, V, bench = _config5()
    g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
    g.set_algo(min(algo, 4))
    g.set_group_mode({4: 1, 6: 2}.get(algo, 0))
    origins = bench.origins_for(0, n, V, step=0)
    o_dev = torch.from_numpy(origins).cuda()
    d_dev = torch.empty((n, V), dtype=torch.float64, device="cuda")
    p_dev = torch.empty((n, V), dtype=torch.int32, device="cuda")
    g.sssp_batch_device(o_dev, d_dev, p_dev)
    torch.cuda.synchronize()
    st = g.last_stats()
    sample = np.arange(n) if n <= 4096 else np.random.default_rng(0).choice(n, 1024, replace=False)
    og = osssp.OracleGraph(s, e, w, max_vertex_id=V - 1)
    do, po = og.sssp_batch(origins[sample])
    idx = torch.from_numpy(sample).cuda()
    assert d_dev[idx].cpu().numpy().tobytes() == do.tobytes()
    assert (p_dev[idx].cpu().numpy() == po).all()
    # size-independent properties over the whole batch: origin rows, reachability, no edge can still relax
    assert bool((d_dev[torch.arange(n, device="cuda"), o_dev.long()] == 0).all())
    assert bool((d_dev < 1e300).all())
    sd, ed, wd = (torch.from_numpy(x).cuda() for x in (s.astype(np.int64), e.astype(np.int64), w))
    for lo in range(0, n, 1024):
        blk = d_dev[lo:lo + 1024]
        assert bool((blk[:, ed] <= blk[:, sd] + wd).all())
    if osssp.ref_available():
        rg = osssp.RefGraph(s, e, w, max_vertex_id=V - 1)
        for i in sample[:8]:
            dr, _ = rg.dijkstra(int(origins[i]), -1)
            assert d_dev[int(i)].cpu().numpy().tobytes() == dr.tobytes()
    # the one-byte code path of the same batch decodes to the same tables
    g.set_parent_format(1)
    codes = torch.empty((n, V), dtype=torch.uint8, device="cuda")
    g.sssp_batch_device(o_dev, None, codes)
    out = torch.empty((min(n, 2368), V), dtype=torch.int32, device="cuda")
    g.decode_parents_device(codes[:len(out)], out)
    torch.cuda.synchronize()
    assert torch.equal(out, p_dev[:len(out)])
    g.set_parent_format(0)
    del st
    g.free()


@pytest.mark.parametrize("algo", [3, 4, 5, 6])
def test_config5_congested_weight_set(spi, algo):
    """Same graph with the second, congested weight set of SURVEY.md 8(d) (queues ~ Poisson(3)) pushed through
    sp_set_weights_device, 1,024 origins (auto mode: the warp kernel; forced: the group kernel)."""
    import torch
    from oracle import sssp as osssp
    s, e, w, V, bench = _config5()
    w2 = _congested_weights(s, e, w)
    g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
    g.set_algo(min(algo, 4))
    g.set_group_mode({4: 1, 6: 2}.get(algo, 0))
    g.set_weights_device(torch.from_numpy(w2).cuda())
    origins = bench.origins_for(0, 1024, V, step=3)
    o_dev = torch.from_numpy(origins).cuda()
    d_dev = torch.empty((1024, V), dtype=torch.float64, device="cuda")
    p_dev = torch.empty((1024, V), dtype=torch.int32, device="cuda")
    g.sssp_batch_device(o_dev, d_dev, p_dev)
    torch.cuda.synchronize()
    do, po = osssp.OracleGraph(s, e, w2, max_vertex_id=V - 1).sssp_batch(origins)
    assert d_dev.cpu().numpy().tobytes() == do.tobytes()
    assert (p_dev.cpu().numpy() == po).all()
    g.free()


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_group_kernel_on_an_irregular_network_with_closures(spi, mode):
    """The group kernel where its assumptions are weakest: a town-like network (streets removed, one-ways, hubs with
    more than four out-links), 3 % of the links closed at 1e8 (threshold jumps, pending-list rescans), origins
    clustered AND scattered, duplicates and out-of-range origins in the batch; forward and reverse trees."""
    import torch
    from oracle import sssp as osssp
    from spatial_queue_b200 import synth
    rng = np.random.default_rng(4)
    s, e, _, _ = synth.town_edges(40, seed=3)
    s, e = np.asarray(s, np.int32), np.asarray(e, np.int32)
    V = 40 * 40
    w = rng.random(len(s)) * 30.0 + 2.0
    w[rng.random(len(s)) < 0.03] = 1e8
    g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
    g.set_algo(4)
    g.set_group_mode(mode)
    origins = np.concatenate([rng.integers(0, V, 300), np.arange(100, 140), [5, 5, 5, -1, V + 7]]).astype(np.int32)
    n = len(origins)
    for reverse in (False, True):
        o_dev = torch.from_numpy(origins).cuda()
        d_dev = torch.empty((n, V), dtype=torch.float64, device="cuda")
        p_dev = torch.empty((n, V), dtype=torch.int32, device="cuda")
        g.sssp_batch_device(o_dev, d_dev, p_dev, reverse=reverse)
        torch.cuda.synchronize()
        og = osssp.OracleGraph(e, s, w, max_vertex_id=V - 1) if reverse else osssp.OracleGraph(s, e, w, max_vertex_id=V - 1)
        do, po = og.sssp_batch(origins)
        assert d_dev.cpu().numpy().tobytes() == do.tobytes()
        assert (p_dev.cpu().numpy() == po).all()
    g.free()


def test_large_graph_overflow_is_rerun_not_dropped(spi):
    """beyond ~436k vertices the bitmap kernel cannot be the fallback.  A 700 x 700 grid (490k vertices), the
    warp kernel forced with a 64-entry ring: every tree overflows, is listed, and must come back complete (rerun on
    sssp_group_kernel, whose queues are bounded by V)."""
    from oracle import sssp as osssp
    n = 700
    s, e, w = grid_graph(n, seed=2)
    g = spi.from_arrays(s, e, w)
    g.set_algo(2)
    g.set_warp_tuning(ring_cap=64, ilp=2)
    origins = np.array([0, n * n // 2 + 17, n * n - 1, 123456, 7], np.int32)
    d, p = g.sssp_batch(origins)
    assert g.last_stats()["overflows"] == len(origins)
    do, po = osssp.OracleGraph(s, e, w).sssp_batch(origins)
    assert d.tobytes() == do.tobytes() and (p == po).all()
    g.free()


@pytest.mark.parametrize("reverse", [False, True])
def test_host_batch_with_one_byte_codes_decodes_to_the_int32_tables(spi, reverse):
    """sp_sssp_batch_codes (host buffers, 1 byte per vertex over PCIe) + sp_decode_parents_host == sp_sssp_batch's int32
    predecessor table; distances optional; in-degrees beyond four and exact ties; several chunks."""
    rng = np.random.default_rng(41)
    V, E = 2500, 13000
    s = rng.integers(0, V, E).astype(np.int32)
    e = rng.integers(0, V, E).astype(np.int32)
    w = np.round(rng.random(E) * 20.0) + 1.0
    g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
    origins = np.concatenate([rng.integers(0, V, 5000), [-1, V + 3]]).astype(np.int32)
    d0, p0 = g.sssp_batch(origins, reverse=reverse)
    d1, c1 = g.sssp_batch_codes(origins, want_dist=True, reverse=reverse)
    assert d1.tobytes() == d0.tobytes()
    assert (g.decode_parents_host(c1, reverse=reverse) == p0).all()
    _, c2 = g.sssp_batch_codes(origins, reverse=reverse)
    assert (c2 == c1).all()
    g.free()


@pytest.mark.parametrize("reverse", [False, True])
def test_host_batch_with_four_bit_codes(spi, reverse):
    """sp_sssp_batch_codes4 (two vertices per byte) + sp_decode_parents_host4 == sp_sssp_batch's int32 table on a road-like
    graph (odd vertex count, several chunks, invalid origins); a graph with an in-degree above 15 is refused."""
    s, e, w = grid_graph(53, seed=11)
    V = 53 * 53
    g = spi.from_arrays(s, e, np.round(w) + 1.0, max_vertex_id=V - 1)      # many exact ties
    rng = np.random.default_rng(43)
    origins = np.concatenate([rng.integers(0, V, 5000), [-1, V + 3]]).astype(np.int32)
    d0, p0 = g.sssp_batch(origins, reverse=reverse)
    d1, c4 = g.sssp_batch_codes4(origins, want_dist=True, reverse=reverse)
    assert c4.shape == (len(origins), (V + 1) // 2)
    assert d1.tobytes() == d0.tobytes()
    assert (g.decode_parents_host4(c4, reverse=reverse) == p0).all()
    _, c1 = g.sssp_batch_codes(origins, reverse=reverse)
    lo = np.minimum(c1[:, 0::2], 15)
    hi = np.minimum(np.concatenate([c1[:, 1::2], np.full((len(origins), (V + 1) // 2 - V // 2), 255, np.uint8)], axis=1), 15)
    assert (c4 == (lo | (hi << 4))).all()
    g.free()
    hub_s = np.arange(1, 18, dtype=np.int32)
    g = spi.from_arrays(np.concatenate([hub_s, [0]]).astype(np.int32), np.concatenate([np.zeros(17, np.int32), [1]]).astype(np.int32),
                        np.ones(18), max_vertex_id=17)
    with pytest.raises(Exception, match="in-degree <= 15"):
        g.sssp_batch_codes4(np.array([1], np.int32), reverse=False)
    g.free()


def test_lone_trees_with_shared_memory_rows_equal_global_rows(spi):
    """Auto mode on a handful of trees (the device router's reverse tree per destination): sssp_ell_kernel keeps the
    distance row in shared memory when the graph fits (config 4's 25,281 vertices do).  Same tables with the switch off,
    forward and reverse, closed links included; and a graph one vertex too many for the row falls back by itself."""
    import torch
    from oracle import sssp as osssp
    s, e, w = grid_graph(159, seed=1)
    V = 159 * 159
    rng = np.random.default_rng(3)
    w = w.copy()
    w[rng.random(len(w)) < 0.01] = 1e8
    g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
    dests = np.array([0, 158, V - 1, V - 159, 79, 79 * 159, V - 80, 80 * 159 - 1], np.int32)
    for reverse in (False, True):
        d1, p1 = g.sssp_batch(dests, reverse=reverse)
        g.set_smem_rows(False)
        d0, p0 = g.sssp_batch(dests, reverse=reverse)
        g.set_smem_rows(True)
        assert d1.tobytes() == d0.tobytes() and (p1 == p0).all()
        og = osssp.OracleGraph(e, s, w, max_vertex_id=V - 1) if reverse else osssp.OracleGraph(s, e, w, max_vertex_id=V - 1)
        do, po = og.sssp_batch(dests)
        assert d1.tobytes() == do.tobytes() and (p1 == po).all()
    g.free()
    s, e, w = grid_graph(162, seed=1)          # 26,244 vertices: the row does not fit any more
    g = spi.from_arrays(s, e, w)
    d, p = g.sssp_batch(dests)
    do, po = osssp.OracleGraph(s, e, w).sssp_batch(dests)
    assert d.tobytes() == do.tobytes() and (p == po).all()
    g.free()
