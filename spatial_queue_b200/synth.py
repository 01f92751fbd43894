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
