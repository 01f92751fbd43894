"""Synthetic road networks and demand in the reference's input formats (SURVEY.md 8d recipe).

`grid_frames` returns the edge / node DataFrames with exactly the columns the loader reads
(`eid,nid_s,nid_e,length,lanes,maxmph,type,geometry` and `nid,osmid,lon,lat`; bolinas_edges_sim.csv:1,
bolinas_nodes_sim.csv:1) and `grid_od` an OD frame (`origin_osmid,destin_osmid,dept_time,type`)."""
import numpy as np
import pandas as pd

M_PER_DEG_LAT = 111_000.0
M_PER_DEG_LON = 88_000.0   # ~cos(37.6 deg) * 111 km


def grid_coords(n, seed=1, jitter=1e-4):
    rng = np.random.default_rng(seed)
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    lon = -122.9 + 0.0022 * ii + rng.normal(0, jitter, (n, n))
    lat = 37.6 + 0.0018 * jj + rng.normal(0, jitter, (n, n))
    return lon.ravel(), lat.ravel()


def grid_edges(n):
    """Directed 4-neighbour links both ways; node id = i*n + j."""
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    nid = ii * n + jj
    s, e = [], []
    for di, dj in ((1, 0), (0, 1)):
        a = nid[: n - di, : n - dj].ravel()
        b = nid[di:, dj:].ravel()
        s += [a, b]
        e += [b, a]
    return np.concatenate(s).astype(np.int64), np.concatenate(e).astype(np.int64)


def grid_attributes(n, s, e, lon, lat, arterial_every=8):
    x, y = lon * M_PER_DEG_LON, lat * M_PER_DEG_LAT
    length = np.hypot(x[s] - x[e], y[s] - y[e])
    si, sj, ei, ej = s // n, s % n, e // n, e % n
    arterial = ((si == ei) & (si % arterial_every == 0)) | ((sj == ej) & (sj % arterial_every == 0))
    lanes = np.where(arterial, 2, 1)
    maxmph = np.where(arterial, 35, 25)
    return length, lanes, maxmph


def grid_frames(n, seed=1, arterial_every=8):
    lon, lat = grid_coords(n, seed)
    s, e = grid_edges(n)
    length, lanes, maxmph = grid_attributes(n, s, e, lon, lat, arterial_every)
    geom = [f"LINESTRING ({lon[a]!r} {lat[a]!r}, {lon[b]!r} {lat[b]!r})" for a, b in zip(s, e)]
    edges = pd.DataFrame({"eid": np.arange(len(s)), "nid_s": s, "nid_e": e, "length": length, "lanes": lanes,
                          "maxmph": maxmph, "type": "residential", "geometry": geom})
    nodes = pd.DataFrame({"nid": np.arange(n * n), "osmid": 10_000_000 + np.arange(n * n), "lon": lon, "lat": lat})
    return edges, nodes


def grid_exits(n):
    m = n // 2
    ij = [(0, 0), (0, n - 1), (n - 1, 0), (n - 1, n - 1), (0, m), (m, 0), (n - 1, m), (m, n - 1)]
    return np.array([i * n + j for i, j in ij], np.int64)


def grid_od(n, n_agents, seed=2, horizon=7200, exits=None):
    rng = np.random.default_rng(seed)
    exits = grid_exits(n) if exits is None else np.asarray(exits)
    origin = rng.integers(0, n * n, n_agents)
    destin = exits[rng.integers(0, len(exits), n_agents)]
    clash = origin == destin
    origin[clash] = (origin[clash] + 1) % (n * n)   # the reference crashes on origin == destination (load_vehicle KeyError)
    dept = 1 + (np.arange(n_agents) * horizon) // max(n_agents, 1)
    return pd.DataFrame({"origin_osmid": 10_000_000 + origin, "destin_osmid": 10_000_000 + destin,
                         "dept_time": dept, "type": "local"})


# ---- direct flat construction for large synthetic cases (no Python objects) ----------------------------------

def _utm10n_np(lon, lat):
    """Vectorised twin of geo.utm10n."""
    from . import geo
    phi = np.radians(lat)
    lam = np.radians(lon) - geo._LON0
    e = 2.0 * np.sqrt(geo._N) / (1.0 + geo._N)
    t = np.sinh(np.arctanh(np.sin(phi)) - e * np.arctanh(e * np.sin(phi)))
    xi = np.arctan2(t, np.cos(lam))
    eta = np.arctanh(np.sin(lam) / np.sqrt(1.0 + t * t))
    x, y = eta.copy(), xi.copy()
    for j, a in enumerate(geo._ALPHA, start=1):
        x += a * np.cos(2 * j * xi) * np.sinh(2 * j * eta)
        y += a * np.sin(2 * j * xi) * np.cosh(2 * j * eta)
    return geo._FE + geo._K0 * geo._AHAT * x, geo._K0 * geo._AHAT * y


def flat_from_edges(lon, lat, s, e, lanes, maxmph):
    """FlatNetwork (without agents) of an arbitrary directed network given as arrays, built with array operations
    only but following the loader's rules (queue_class.py:54-62, 137-176, 416): virtual source links first in node
    order, then the real links in edge order; incoming order = that order filtered by end node; turn angles from the
    projected two-point geometry with the reference's wrap rule.  Links need not have a reverse link and nodes need
    not have out-links."""
    from .flatten import FlatNetwork
    s, e = np.asarray(s, np.int64), np.asarray(e, np.int64)
    N, E = len(lon), len(s)
    L = N + E
    x0, y0 = lon * M_PER_DEG_LON, lat * M_PER_DEG_LAT
    length = np.hypot(x0[s] - x0[e], y0[s] - y0[e])
    fft = length / maxmph * 2.237
    capacity = (1900 * lanes).astype(np.float64)
    x, y = _utm10n_np(lon, lat)
    ang = np.arctan2(y[e] - y[s], x[e] - x[s])   # in_angle == out_angle for a straight two-point link

    f = FlatNetwork()
    f.node_key = np.arange(N, dtype=np.int64)
    f.link_start_node = np.concatenate([np.full(N, -1), s]).astype(np.int32)
    f.link_end_node = np.concatenate([np.arange(N), e]).astype(np.int32)
    f.link_is_virtual = np.concatenate([np.ones(N), np.zeros(E)]).astype(np.uint8)
    f.link_lanes_floor = np.concatenate([np.full(N, 100), lanes]).astype(np.int32)
    cap_all = np.concatenate([np.full(N, 100000.0), capacity])
    f.link_capacity = cap_all
    f.link_cap_per_s = cap_all / 3600
    f.link_storage_cap = np.concatenate([np.full(N, 18.0), np.maximum(18, length * lanes)])
    f.link_fft = np.concatenate([np.zeros(N), fft])
    f.link_closed = np.zeros(L, np.uint8)
    f.link_maxmph = np.concatenate([np.full(N, 1e5), length / fft * 2.23694])
    # reverse link (e -> s), where there is one
    key = s * N + e
    order = np.argsort(key, kind="stable")
    skey = key[order]
    pos = np.minimum(np.searchsorted(skey, e * N + s), E - 1)
    rev = np.where(skey[pos] == e * N + s, N + order[pos], -1)
    f.link_reverse = np.concatenate([np.full(N, -1), rev]).astype(np.int32)

    # CSR of incoming (virtual first, then real in edge order) and outgoing links
    in_order = np.argsort(e, kind="stable")
    in_cnt = np.bincount(e, minlength=N) + 1
    f.node_in_ptr = np.concatenate([[0], np.cumsum(in_cnt)]).astype(np.int32)
    in_links = np.empty(f.node_in_ptr[-1], np.int32)
    in_links[f.node_in_ptr[:-1]] = np.arange(N)
    real_pos = np.arange(E) - np.repeat(np.concatenate([[0], np.cumsum(in_cnt - 1)])[:-1], in_cnt - 1)
    in_links[np.repeat(f.node_in_ptr[:-1] + 1, in_cnt - 1) + real_pos] = N + in_order
    f.node_in_links = in_links
    out_order = np.argsort(s, kind="stable")
    out_cnt = np.bincount(s, minlength=N)
    f.node_out_ptr = np.concatenate([[0], np.cumsum(out_cnt)]).astype(np.int32)
    f.node_out_links = (N + out_order).astype(np.int32)
    f.node_in_phase = np.zeros(len(in_links), np.uint8)
    f.node_in_prio = np.full(len(in_links), -1, np.int32)
    f.node_type = np.zeros(N, np.uint8)

    # turn table (queue_class.py:166-176)
    tcnt = in_cnt * out_cnt
    f.node_turn_ptr = np.concatenate([[0], np.cumsum(tcnt)]).astype(np.int64)
    node_of_entry = np.repeat(np.arange(N), tcnt)
    k = np.arange(f.node_turn_ptr[-1]) - f.node_turn_ptr[node_of_entry]
    i_loc, j_loc = k // out_cnt[node_of_entry], k % out_cnt[node_of_entry]
    il = in_links[f.node_in_ptr[node_of_entry] + i_loc]
    ol = f.node_out_links[f.node_out_ptr[node_of_entry] + j_loc]
    ang_all = np.concatenate([np.zeros(N), ang])
    turn = -ang_all[il] + ang_all[ol]
    turn = np.where(turn >= np.pi, 2 * np.pi - turn, turn)
    turn = np.where(turn <= -np.pi, 2 * np.pi + turn, turn)
    turn = turn / np.pi * 180
    turn = np.where(il < N, 0.0, turn)
    f.turn_angle = turn
    # (start, end) -> link table
    f.pair_start, f.pair_end, f.pair_link = s.astype(np.int32), e.astype(np.int32), (N + np.arange(E)).astype(np.int32)
    f.edge_start, f.edge_end, f.edge_fft = s.astype(np.int32), e.astype(np.int32), fft
    # router arrays (flatten.router_arrays): node i is graph vertex i, real link N + k is edge k
    f.node_vertex = np.arange(N, dtype=np.int32)
    f.link_edge = np.concatenate([np.full(N, -1), np.arange(E)]).astype(np.int32)
    return f


def flat_grid(n, seed=1, arterial_every=8):
    """FlatNetwork (without agents) of the n x n grid."""
    lon, lat = grid_coords(n, seed)
    s, e = grid_edges(n)
    _, lanes, maxmph = grid_attributes(n, s, e, lon, lat, arterial_every)
    f = flat_from_edges(lon, lat, s, e, lanes, maxmph)
    f.n_grid = n
    return f


def routes_from_next_hop(next_link, link_end, origin, dest_idx, max_hops):
    """Materialise link sequences: next_link[d, v] is the link to take at node v towards destination slot d
    (-1 at the destination).  Returns (route_ptr, route_links)."""
    A = len(origin)
    cur = origin.astype(np.int64).copy()
    alive = np.arange(A)
    cols = []
    for _ in range(max_hops):
        l = next_link[dest_idx[alive], cur[alive]]
        keep = l >= 0
        alive = alive[keep]
        if len(alive) == 0:
            break
        l = l[keep]
        cols.append((alive.copy(), l))
        cur[alive] = link_end[l]
    lens = np.zeros(A, np.int64)
    for a, _ in cols:
        lens[a] += 1
    ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    links = np.empty(ptr[-1], np.int32)
    for h, (a, l) in enumerate(cols):
        links[ptr[a] + h] = l
    return ptr, links


def add_agents(f, origin, destin, dept, vlen, next_link, dest_nodes):
    """Attach agents (already in network.agents order) and their routes to a flat network."""
    A = len(origin)
    f.agent_key = np.arange(A, dtype=np.int64)
    f.agent_origin = origin.astype(np.int32)
    f.agent_destin = destin.astype(np.int32)
    f.agent_dept = dept.astype(np.int32)
    f.agent_vlen = vlen.astype(np.int32)
    f.agent_status = np.zeros(A, np.uint8)
    slot = {int(d): i for i, d in enumerate(dest_nodes)}
    dest_idx = np.array([slot[int(d)] for d in destin], np.int64) if A < 100000 else \
        np.searchsorted(np.sort(dest_nodes), destin)
    if A >= 100000:
        srt = np.argsort(dest_nodes)
        dest_idx = srt[dest_idx]
    f.dest_nodes = np.asarray(dest_nodes, np.int32)
    f.agent_dslot = np.asarray(dest_idx, np.int32)
    if next_link is None:      # routes come from the device router (sq_attach_router)
        f.route_ptr, f.route_links = np.zeros(A + 1, np.int64), np.zeros(0, np.int32)
        return f
    f.route_ptr, f.route_links = routes_from_next_hop(next_link, f.link_end_node, origin, dest_idx,
                                                      max_hops=4 * int(np.sqrt(len(f.node_key))) + 8)
    return f


def next_link_tables(f, exits, router=None):
    """next_link[d, v] = link to take at node v towards exit d (-1 at the exit and where there is no path), from one
    tree per exit on the reversed graph.  Routing runs on the GPU through the product library unless `router`
    supplies a callable (starts, ends, w, roots) -> parent[len(roots), V] (how the CPU tests inject the oracle)."""
    s, e, w = f.edge_start, f.edge_end, f.edge_fft
    N = len(f.node_key)
    if router is None:
        from . import sp_interface
        g = sp_interface.from_arrays(e, s, w, max_vertex_id=N - 1)      # reversed graph
        _, parent = g.sssp_batch(np.asarray(exits, np.int32), want_dist=False)
        g.free()
    else:
        parent = router(e, s, w, exits)
    # next link at node v towards exit d: the link (v -> parent[d, v])
    key = s.astype(np.int64) * N + e
    order = np.argsort(key)
    v = np.arange(N, dtype=np.int64)
    next_link = np.full((len(exits), N), -1, np.int32)
    for d in range(len(exits)):
        p = parent[d].astype(np.int64)
        ok = p >= 0
        pos = np.searchsorted(key[order], v[ok] * N + p[ok])
        next_link[d, v[ok]] = N + order[pos]
    return next_link


def grid_scenario(n, n_agents, horizon=7200, seed=1, arterial_every=8, exits=None, router=None, device_routes=False):
    """The C4-style workload: n x n grid, agents from uniform origins to the exit nodes with
    staggered departures, routed at t=0 on free-flow weights.

    Routing runs on the GPU through the product library: one tree per exit on the reversed graph gives the
    next hop of every node towards that exit (`router` may supply a callable (starts, ends, w, roots) ->
    parent[len(roots), V] instead, which is how the CPU tests inject the oracle)."""
    f = flat_grid(n, seed, arterial_every)
    N = n * n
    exits = grid_exits(n) if exits is None else np.asarray(exits, np.int64)
    od = grid_od(n, n_agents, seed + 1, horizon, exits)
    origin = (od.origin_osmid.values - 10_000_000).astype(np.int64)
    destin = (od.destin_osmid.values - 10_000_000).astype(np.int64)
    rng = np.random.default_rng(seed + 2)
    perm = rng.permutation(n_agents)            # add_demand shuffles the OD rows (queue_class.py:198)
    origin, destin, dept = origin[perm], destin[perm], od.dept_time.values[perm]
    if device_routes:
        # no host routing at all: DeviceSim(flat, routes_from_router=True) + attach_router(graph(flat), flat)
        add_agents(f, origin, destin, dept, np.full(n_agents, 8), None, exits)
        return f
    next_link = next_link_tables(f, exits, router)
    f.next_link_table = next_link
    add_agents(f, origin, destin, dept, np.full(n_agents, 8), next_link, exits)
    return f


def town_edges(n, seed=1, drop=0.12, oneway=0.08, diag=0.06, hubs=2, hub_links=10):
    """An irregular street network on the jittered n x n lattice (the stand-in for the Berkeley / Butte inputs the
    reference does not ship): a share of the streets removed, a share one-way, diagonal shortcuts (node degrees from
    0 to 8), a few hubs with `hub_links` extra two-way links to random nodes (more out-links than the node kernel
    caches), links in shuffled order, 1-4 lanes, three speed classes."""
    rng = np.random.default_rng(seed)
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    nid = ii * n + jj
    und = [np.stack([nid[:-1, :].ravel(), nid[1:, :].ravel()], 1), np.stack([nid[:, :-1].ravel(), nid[:, 1:].ravel()], 1)]
    und = np.concatenate(und)
    und = und[rng.random(len(und)) >= drop]
    d1 = np.stack([nid[:-1, :-1].ravel(), nid[1:, 1:].ravel()], 1)
    d2 = np.stack([nid[1:, :-1].ravel(), nid[:-1, 1:].ravel()], 1)
    dg = np.concatenate([d1[rng.random(len(d1)) < diag / 2], d2[rng.random(len(d2)) < diag / 2]])
    und = np.concatenate([und, dg])
    for h in rng.choice(n * n, hubs, replace=False):
        far = rng.choice(np.setdiff1d(np.arange(n * n), [h]), hub_links, replace=False)
        und = np.concatenate([und, np.stack([np.full(hub_links, h), far], 1)])
    und = np.unique(np.sort(und, 1), axis=0)                         # no duplicate streets
    und = und[rng.permutation(len(und))]
    u = rng.random(len(und))
    fwd, bwd = u >= oneway / 2, (u < oneway / 2) | (u >= oneway)     # one direction only below `oneway`
    s = np.concatenate([und[fwd, 0], und[bwd, 1]])
    e = np.concatenate([und[fwd, 1], und[bwd, 0]])
    perm = rng.permutation(len(s))
    s, e = s[perm], e[perm]
    lanes = rng.choice([1, 1, 1, 2, 2, 3, 4], len(s))
    maxmph = rng.choice([25, 25, 35, 45], len(s))
    return s.astype(np.int64), e.astype(np.int64), lanes, maxmph


def town_frames(n, seed=1, **shape):
    """The irregular network of `town_edges` as the loader's CSV frames (edges, nodes), like `grid_frames`."""
    lon, lat = grid_coords(n, seed)
    s, e, lanes, maxmph = town_edges(n, seed, **shape)
    x, y = lon * M_PER_DEG_LON, lat * M_PER_DEG_LAT
    length = np.hypot(x[s] - x[e], y[s] - y[e])
    geom = [f"LINESTRING ({lon[a]!r} {lat[a]!r}, {lon[b]!r} {lat[b]!r})" for a, b in zip(s, e)]
    edges = pd.DataFrame({"eid": np.arange(len(s)), "nid_s": s, "nid_e": e, "length": length, "lanes": lanes,
                          "maxmph": maxmph, "type": "residential", "geometry": geom})
    nodes = pd.DataFrame({"nid": np.arange(n * n), "osmid": 10_000_000 + np.arange(n * n), "lon": lon, "lat": lat})
    return edges, nodes


def town_scenario(n, n_agents, horizon=600, seed=1, n_exits=8, signals=0.04, priority=0.03, router=None,
                  device_routes=False, exits_on_rim=False, exits_in_core=False, **shape):
    """Irregular network + signal and priority nodes + demand from every node that can reach its exit.  Signal nodes
    alternate between the roughly north-south and the roughly east-west approaches (the source link is always
    green), priority nodes rank all their approaches in a random order."""
    rng = np.random.default_rng(seed + 10)
    lon, lat = grid_coords(n, seed)
    s, e, lanes, maxmph = town_edges(n, seed, **shape)
    f = flat_from_edges(lon, lat, s, e, lanes, maxmph)
    f.edge_lanes, f.edge_maxmph = np.asarray(lanes), np.asarray(maxmph)
    N = n * n
    indeg = np.diff(f.node_in_ptr)
    pick = rng.random(N)
    f.node_type = np.where((pick < signals) & (indeg >= 3), 1, np.where((pick > 1 - priority) & (indeg >= 3), 2, 0)).astype(np.uint8)
    node_of = np.repeat(np.arange(N), indeg)
    il = f.node_in_links
    real = il >= N
    dx = np.abs(lon[f.link_end_node[il]] - lon[np.maximum(f.link_start_node[il], 0)]) * M_PER_DEG_LON
    dy = np.abs(lat[f.link_end_node[il]] - lat[np.maximum(f.link_start_node[il], 0)]) * M_PER_DEG_LAT
    phase = np.where(real, np.where(dy >= dx, 1, 2), 3)
    f.node_in_phase = np.where(f.node_type[node_of] == 1, phase, 0).astype(np.uint8)
    # priority ranks: a random order of each node's approaches (rank = position among the node's sorted draws)
    draw = rng.random(len(il))
    o = np.lexsort((draw, node_of))
    rank = np.empty(len(il), np.int64)
    rank[o] = np.arange(len(il)) - f.node_in_ptr[node_of[o]]
    f.node_in_prio = np.where(f.node_type[node_of] == 2, rank, -1).astype(np.int32)

    candidates = np.arange(N)
    if exits_in_core:
        # exits the network can actually reach: nodes of the largest strongly connected component (a removed street can
        # leave a rim node with a handful of upstream nodes; every agent bound for it would then start at one node)
        from scipy.sparse import coo_matrix
        from scipy.sparse.csgraph import connected_components
        _, lab = connected_components(coo_matrix((np.ones(len(s)), (s, e)), shape=(N, N)), directed=True, connection="strong")
        candidates = np.flatnonzero(lab == np.bincount(lab).argmax())
    if exits_on_rim:
        # an evacuation: the exits are nodes on the rim of the network
        ii, jj = np.divmod(candidates, n)
        candidates = candidates[((ii == 0) | (ii == n - 1) | (jj == 0) | (jj == n - 1)) & (indeg[candidates] >= 2)]
    exits = rng.choice(candidates, n_exits, replace=False).astype(np.int64)
    next_link = next_link_tables(f, exits, router)
    dslot = rng.integers(0, n_exits, n_agents)
    origin = np.empty(n_agents, np.int64)
    for d in range(n_exits):
        ok = np.flatnonzero(next_link[d] >= 0)
        m = dslot == d
        origin[m] = rng.choice(ok, int(m.sum()))
    dept = np.sort(rng.integers(0, horizon, n_agents))[rng.permutation(n_agents)]
    f.next_link_table = next_link
    add_agents(f, origin, exits[dslot], dept, rng.choice([6, 8, 8, 8, 12], n_agents), None if device_routes else next_link, exits)
    return f


def routing_graph(f):
    """The routing graph of a flat grid network (what the loader builds with interface.from_dataframe on the real
    links, queue_class.py:123): edge k = real link N + k, weight = fft."""
    from . import sp_interface
    return sp_interface.from_arrays(f.edge_start, f.edge_end, f.edge_fft, max_vertex_id=len(f.node_key) - 1)


# ---- stand-ins for the two configurations whose inputs the reference does not ship -------------------

def berkeley_standin(router=None, device_routes=False, n_agents=50_000):
    """BASELINE config 2 (benchmark/profile_berkeley.py: 1,401 ticks, no rerouting; link ids up to 37,806 at
    model/dta_meso_berkeley.py:92; the OD and network files are not shipped): an OSM-like network of ~38k links on
    ~15k nodes -- the jittered lattice with 36 % of its streets removed, 8 % one-way, diagonals, hubs, 1-4 lanes --
    and 50,000 agents leaving over twenty minutes."""
    return town_scenario(122, n_agents, horizon=1200, seed=21, n_exits=12, router=router, device_routes=device_routes,
                         exits_in_core=True, drop=0.36, oneway=0.08, diag=0.06, hubs=4, hub_links=8)


def butte_standin(router=None, device_routes=True, n_agents=100_000):
    """BASELINE config 3 (model/dta_meso_butte.py: 18,001 ticks, link ids up to 30,175 at :99; inputs not shipped):
    ~31k links on ~12k nodes, 100,000 evacuees leaving over two hours for six exits on the rim."""
    return town_scenario(110, n_agents, horizon=7200, seed=31, n_exits=6, router=router, device_routes=device_routes,
                         exits_on_rim=True, exits_in_core=True, drop=0.36, oneway=0.08, diag=0.06, hubs=3, hub_links=8)


def butte_closure_schedule(f, seed=7, last_fire_tick=340 * 60, fire_share=0.005, p_close=0.15, reopen_after=3600):
    """The closure schedule of model/dta_meso_butte.py:56-115 with the fire raster replaced by host draws (SURVEY.md
    8d): every 300 s until t = 20,400 the fire reaches 0.5 % of the residential links (here: one-lane links); a link
    it reaches for the first time burns, and with probability 0.15 it is closed to newcomers together with its
    opposite link, to reopen 3,600 s later (:111-115; the reopening acts on the link model of the following second);
    three timed closures of arterials (two, two and eight links with three or more lanes) at 1,800 / 3,600 / 10,800 s
    stay for good (:88-101).  Returns [(tick, link index, closed)] sorted by tick -- the input of
    sq_set_closure_schedule and of the host paths it is compared with."""
    rng = np.random.default_rng(seed)
    N = len(f.node_key)
    real = np.arange(N, len(f.link_end_node))
    lanes = np.asarray(f.link_lanes_floor)
    residential = real[lanes[real] == 1]
    arterial = real[lanes[real] >= 3]
    rev = np.asarray(f.link_reverse)
    burnt = np.zeros(len(f.link_end_node), bool)
    forever = np.zeros(len(f.link_end_node), bool)
    ev = []
    timed = {1800: 2, 3600: 2, 10800: 8}
    picks = rng.choice(arterial, sum(timed.values()), replace=False)
    k = 0
    timed_links = {}
    for t, cnt in timed.items():
        timed_links[t] = picks[k:k + cnt]
        k += cnt
    for t in range(0, last_fire_tick + 1, 300):
        for l in rng.choice(residential, max(1, int(fire_share * len(residential))), replace=False):
            if burnt[l]:
                continue
            burnt[l] = True
            opp = rev[l]
            if opp >= 0:
                burnt[opp] = True
            if rng.random() < p_close:
                for x in ([l] if opp < 0 else [l, opp]):
                    if not forever[x]:
                        ev.append((t, int(x), 1))
                        ev.append((t + reopen_after + 1, int(x), 0))
        if t in timed_links:
            for x in timed_links[t]:
                forever[x] = True
                ev.append((t, int(x), 1))
    # a timed closure wins over a later reopening of the same link
    ev = [e for e in ev if not (e[2] == 0 and forever[e[1]])]
    return sorted(ev, key=lambda e: e[0])
