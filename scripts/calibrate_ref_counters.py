#!/usr/bin/env python3
"""Work counters of the REFERENCE algorithm (oracle/rad_oracle.c == usearch index.hpp:2183-2191 counters) per inserted
vector, for the bench workload's generator and parameters (M=8, ef_add=400), at a few sizes; least-squares fit
D(n) = d0 + d1 ln n, H(n) = h0 + h1 ln n -> profiles/ref_counters.json.  bench.py's roofline uses this fit so that
the GPU build's own (possibly wasteful) work never inflates the achieved-bandwidth figure.
CPU only; single-threaded oracle; the 300k point takes about five minutes."""
import json
import os
import sys
import time
from concurrent.futures import ProcessPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one(n):
    from oracle import oracle as O
    from rad_b200 import synth
    db = synth.clustered_fingerprints(n, seed=0)
    ox = O.OracleIndex(1024, 8, 400, 64)
    t = time.time()
    ox.add(np.arange(n, dtype=np.uint64), db)
    c = ox.counters()
    return {"rows": n, "seconds": time.time() - t, "distances_per_insert": c["computed_distances"] / n,
            "expansions_per_insert": c["iteration_cycles"] / n,
            "refine_distances_per_insert": c["distances_in_refines"] / n,
            "reverse_refine_distances_per_insert": c["distances_in_reverse_refines"] / n}


if __name__ == "__main__":
    args = sys.argv[1:]
    append = "--append" in args  # keep the points already in profiles/ref_counters.json and refit over all of them
    sizes = [int(a) for a in args if a != "--append"] or [10000, 30000, 100000, 300000]
    with ProcessPoolExecutor(len(sizes)) as ex:
        rows = list(ex.map(one, sizes))
    path = os.path.join(ROOT, "profiles", "ref_counters.json")
    if append and os.path.exists(path):
        with open(path) as fh:
            old = json.load(fh)["table"]
        rows = sorted([r for r in old if r["rows"] not in sizes] + rows, key=lambda r: r["rows"])
    x = np.log([r["rows"] for r in rows])
    A = np.stack([np.ones_like(x), x], axis=1)
    d = np.linalg.lstsq(A, np.array([r["distances_per_insert"] for r in rows]), rcond=None)[0]
    h = np.linalg.lstsq(A, np.array([r["expansions_per_insert"] for r in rows]), rcond=None)[0]
    out = {"parameters": {"connectivity": 8, "expansion_add": 400, "generator": "rad_b200.synth seed 0"},
           "table": rows, "fit": {"d0": float(d[0]), "d1": float(d[1]), "h0": float(h[0]), "h1": float(h[1])},
           "note": "computed_distances includes the refine and reverse-refine distances (index.hpp:2183-2191)"}
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    resid = {str(r["rows"]): {"distances": float(r["distances_per_insert"] - (d[0] + d[1] * np.log(r["rows"]))),
                              "expansions": float(r["expansions_per_insert"] - (h[0] + h[1] * np.log(r["rows"])))}
             for r in rows}
    out["fit_residuals"] = resid
    with open(path, "w") as fh:
        json.dump(out, fh, indent=1)
    print(json.dumps(out, indent=1))
