"""CPU probe: recall@10 of the BATCHED build algorithm (tests/model/batched_model.c -- what the CUDA kernels compute,
bit for bit) against the reference (oracle/_ref, the unmodified usearch) on the routable and the adversarial sets.
Test infrastructure; run where /root/reference was compiled:  python scripts/model_recall_probe.py analogue 200000"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O  # noqa: E402
from rad_b200 import synth  # noqa: E402
from tests.model import model as M  # noqa: E402


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "analogue"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
    ef_add = int(sys.argv[3]) if len(sys.argv) > 3 else 400
    extra = dict(a.split("=") for a in sys.argv[4:])
    gen = {"analogue": synth.analogue_set, "dense": synth.dense_set, "sparse": synth.clustered_fingerprints}[kind]
    db = gen(n, seed=3)
    q = synth.queries_from_db(db, 1000, seed=9)
    _, exact_d, _ = O.exact_search(db, q, 10)
    kth = exact_d[:, 9:10]
    lv, _ = O.levels(n, 8)
    out = {}
    if "noref" not in extra:
        ui = O.ref_index_module()
        ref = ui.Index(ndim=1024, dtype="b1", metric="tanimoto", connectivity=8, expansion_add=ef_add)
        t0 = time.time()
        ref.add(np.arange(n, dtype=np.uint64), db, threads=int(extra.get("threads", os.cpu_count())))
        t_ref = time.time() - t0
        for ef in (64, 128):
            ref.expansion_search = ef
            d = ref.search(q, 10).distances
            out[f"ref@{ef}"] = float((d <= kth).mean())
        out["ref_s"] = round(t_ref, 1)
    kw = {k: int(v) for k, v in extra.items() if k in ("batch_max", "growth_den", "batch_k")}
    t0 = time.time()
    model = M.BatchedModel(db, lv, connectivity=8, expansion_add=ef_add, **kw)
    model.build()
    out["model_s"] = round(time.time() - t0, 1)
    for ef in (64, 128):
        _, d, c = model.search(q, 10, ef)
        out[f"model@{ef}"] = float((d <= kth).mean())
    out["model_counters"] = model.counters()
    print(kind, n, ef_add, kw, out)


if __name__ == "__main__":
    main()
