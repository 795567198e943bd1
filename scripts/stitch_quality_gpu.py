#!/usr/bin/env python3
"""GPU probe: recall@10 of the stitched graph against the one-piece graph for several (k_remote, ef_merge), with the
shards as handles on ONE GPU (threads + sharded.LocalComm).   python scripts/stitch_quality_gpu.py KIND ROWS SHARDS kr:ef ..."""
import os
import sys
import threading
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rad_b200 import sharded, synth  # noqa: E402
from rad_b200.index import Index  # noqa: E402


def recall(d, exact, k=10):
    return float((d[:, :k] <= exact[:, k - 1:k]).mean())


def main():
    kind, n, shards = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    combos = [tuple(int(x) for x in a.split(":")) for a in sys.argv[4:]] or [(16, 128), (16, 64)]
    m, ef_add = 8, 400
    db = synth.named_set(kind, n, seed=17)
    bounds = [n * g // shards for g in range(shards + 1)]
    q = synth.queries_from_db(db, 2000, seed=9)
    one = Index(ndim=1024, connectivity=m, expansion_add=ef_add, expansion_search=64)
    one.add(np.arange(n, dtype=np.uint64), db)
    exact = one.search(q, 10, exact=True).distances
    base = {}
    for ef in (64, 128):
        one.expansion_search = ef
        base[ef] = recall(one.search(q, 10).distances, exact)
    print(f"{kind} {n} rows: one-piece {base}", flush=True)
    parts = []
    for g in range(shards):
        ix = Index(ndim=1024, connectivity=m, expansion_add=ef_add, expansion_search=64)
        ix.seek_level_stream(bounds[g])
        ix.add(np.arange(bounds[g], bounds[g + 1], dtype=np.uint64), db[bounds[g]:bounds[g + 1]])
        parts.append(ix)
    for kr, efm in combos:
        hub = sharded.LocalComm(shards)
        results, errors = [None] * shards, []

        def run(g):
            try:
                import torch
                torch.cuda.set_device(0)
                results[g] = sharded.build_global_graph(sharded.DeviceShard(parts[g]), bounds[g], expansion_add=ef_add,
                                                        comm=hub.rank_view(g), k_remote=kr, ef_merge=efm)
            except Exception as e:  # noqa: BLE001
                errors.append(e)
                hub._bar.abort()

        t0 = time.time()
        th = [threading.Thread(target=run, args=(g,)) for g in range(shards)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        assert not errors, errors
        res = results[0]
        whole = Index(ndim=1024, connectivity=m, expansion_add=ef_add, expansion_search=64)
        whole.import_graph(np.arange(n, dtype=np.uint64), res["levels"], res["adj0"], res["adj_upper"], db, res["entry"])
        out = {}
        for ef in (64, 128):
            whole.expansion_search = ef
            out[ef] = recall(whole.search(q, 10).distances, exact)
        print(f"  {shards} shards k_remote={kr} ef_merge={efm}: stitched {out}  ({time.time() - t0:.1f} s)", flush=True)


if __name__ == "__main__":
    main()
