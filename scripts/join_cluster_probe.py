#!/usr/bin/env python3
"""GPU probe: wall time of Index.join and Index.cluster beside the unmodified reference (oracle/_ref, all host threads) on
the same seeded sets.   python scripts/join_cluster_probe.py [ROWS_A ROWS_B]   -> one JSON line per measurement"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O  # noqa: E402
from rad_b200 import join as J, synth  # noqa: E402
from rad_b200.index import Index  # noqa: E402


def main():
    n_a = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
    n_b = int(sys.argv[2]) if len(sys.argv) > 2 else 300_000
    ref_scale = 10                                    # the reference gets 1/10 of the rows (bounded CPU time)
    db = synth.named_set("dense", n_a + n_b, seed=5)
    db = db[np.random.default_rng(1).permutation(len(db))]
    va, vb = db[:n_a], db[n_a:]
    a = Index(ndim=1024, connectivity=8, expansion_add=400, expansion_search=64)
    b = Index(ndim=1024, connectivity=8, expansion_add=400, expansion_search=64)
    t = time.time()
    a.add(np.arange(n_a, dtype=np.uint64), va)
    b.add(np.arange(n_b, dtype=np.uint64) + 10**9, vb)
    print(json.dumps({"what": "gpu build of both indexes", "rows": [n_a, n_b], "s": round(time.time() - t, 3)}), flush=True)
    for exact in (True, False):
        J.join(a, b, max_proposals=12, exact=exact)   # warm-up
        stats = {}
        t = time.time()
        pairs = J.join(a, b, max_proposals=12, exact=exact, stats=stats)
        dt = time.time() - t
        print(json.dumps({"what": "gpu join", "exact": exact, "rows": [n_a, n_b], "proposals": 12, "pairs": len(pairs),
                          "generations": stats["generations"], "engagements": stats["engagements"], "s": round(dt, 3),
                          "men_per_s": round(n_a / dt)}), flush=True)
    for level_min in (100, 2000):
        a.cluster(vectors=vb[:1000], min_count=level_min, max_count=10**9)
        t = time.time()
        c = a.cluster(vectors=vb, min_count=level_min, max_count=10**9)
        dt = time.time() - t
        print(json.dumps({"what": "gpu cluster", "queries": n_b, "index_rows": n_a, "min_count": level_min,
                          "centroids": int(len(c.centroids_popularity[0])), "s": round(dt, 3),
                          "queries_per_s": round(n_b / dt)}), flush=True)
    if not O.ref_available():
        return
    ui = O.ref_index_module()
    ra, rb_ = n_a // ref_scale, n_b // ref_scale
    x = ui.Index(ndim=1024, dtype="b1", metric="tanimoto", connectivity=8, expansion_add=400, expansion_search=64)
    y = ui.Index(ndim=1024, dtype="b1", metric="tanimoto", connectivity=8, expansion_add=400, expansion_search=64)
    t = time.time()
    x.add(np.arange(ra, dtype=np.uint64), va[:ra])
    y.add(np.arange(rb_, dtype=np.uint64) + 10**9, vb[:rb_])
    print(json.dumps({"what": "reference build of both indexes", "rows": [ra, rb_], "cores": os.cpu_count(),
                      "s": round(time.time() - t, 3)}), flush=True)
    for exact in (True, False):
        t = time.time()
        pairs = x.join(y, max_proposals=12, exact=exact)
        dt = time.time() - t
        print(json.dumps({"what": "reference join", "exact": exact, "rows": [ra, rb_], "proposals": 12, "pairs": len(pairs),
                          "cores": os.cpu_count(), "s": round(dt, 3), "men_per_s": round(ra / dt)}), flush=True)
    t = time.time()
    c = x.cluster(vectors=vb[:rb_], min_count=100, max_count=10**9)
    dt = time.time() - t
    print(json.dumps({"what": "reference cluster", "queries": rb_, "index_rows": ra, "min_count": 100,
                      "centroids": int(len(c.centroids_popularity[0])), "cores": os.cpu_count(), "s": round(dt, 3),
                      "queries_per_s": round(rb_ / dt)}), flush=True)


if __name__ == "__main__":
    main()
