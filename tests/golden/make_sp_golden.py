"""Generate the HP-2 golden vectors from the REFERENCE ITSELF (runs only in the build container).

Inputs : /root/reference/projects/bolinas_civic/sp/network.mtx (the graph of sp/tests/graph_test.cc:106-128)
         oracle/_ref/liblsp.so  (the reference's sp library compiled from its own sources by oracle/Makefile)
Outputs: tests/golden/sp_network.npz   edge list of network.mtx (int32 starts/ends, f64 weights, insertion order)
         tests/golden/sp_golden.json   known answers of graph_test.cc + reference outputs for a few trees
                                       (sha256 of the f64 distance array, sampled entries, reference parents
                                       at sampled vertices), produced by calling the reference library.
Run:  python tests/golden/make_sp_golden.py
"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import sssp as osssp  # noqa: E402

MTX = "/root/reference/projects/bolinas_civic/sp/network.mtx"
OUT = os.path.dirname(os.path.abspath(__file__))


def read_mtx(path):
    s, e, w = [], [], []
    nv = None
    with open(path) as f:
        for line in f:
            if "%" in line or "#" in line or not line.strip():
                continue
            parts = line.split()
            if nv is None:
                nv = int(parts[0])
                continue
            s.append(int(parts[0])); e.append(int(parts[1])); w.append(float(parts[2]))
    return nv, np.array(s, np.int32), np.array(e, np.int32), np.array(w, np.float64)


def main():
    osssp.build()
    nv, s, e, w = read_mtx(MTX)
    np.savez_compressed(os.path.join(OUT, "sp_network.npz"), header_n=np.int64(nv), starts=s, ends=e, weights=w)
    # the reference reads the file itself (readgraph -> nvertices = header + 1)
    L = osssp.ref_lib()
    g = osssp.RefGraph(handle=L.readgraph(MTX.encode(), True))
    g.V = nv + 1
    gold = {"source": "oracle/_ref/liblsp.so built from /root/reference sp sources; graph_test.cc known answers",
            "network_mtx": {"header_n": nv, "nvertices": nv + 1, "nedges": int(len(s))},
            "known_answers": {  # sp/tests/graph_test.cc
                "directed_simple_1_3": 7.2, "directed_update_1_3_to_3.7": 3.7,
                "directed_remove_3_1": 7.2, "directed_remove_2_4": 9.1,
                "undirected_simple_1_3": 5.6, "undirected_remove_3_1": 9.1,
                "network_1020_20": 12409.660000000002},
            "trees": []}
    rng = np.random.default_rng(7)
    sample = np.sort(rng.choice(nv + 1, size=48, replace=False)).tolist() + [20]
    for origin in [1020, 1, 17, 30000]:
        d, p = g.dijkstra(origin, -1)
        og = osssp.OracleGraph(s, e, w, max_vertex_id=nv, directed=True)
        canon = og.canonical_parents(origin, d)
        gold["trees"].append({
            "origin": origin,
            "dist_sha256": hashlib.sha256(d.tobytes()).hexdigest(),
            "canonical_parent_sha256": hashlib.sha256(canon.tobytes()).hexdigest(),
            "reached": int((d < 1e300).sum()),
            "sample_vertices": sample,
            "sample_dist_hex": [float(d[v]).hex() for v in sample],
            "sample_ref_parent": [int(p[v]) for v in sample],
            "ref_parent_equals_canonical": bool((p == canon).all()),
        })
    d1 = g.one_to_one(1020, 20)
    gold["one_to_one_1020_20_hex"] = float(d1).hex()
    with open(os.path.join(OUT, "sp_golden.json"), "w") as f:
        json.dump(gold, f, indent=1)
    print("wrote", OUT, "1020->20 =", repr(d1))


if __name__ == "__main__":
    main()
