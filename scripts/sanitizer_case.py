#!/usr/bin/env python3
"""The small case compute-sanitizer runs (profiles/r2_sanitizer_*.txt): a 20 k-row build (batched insert: search kernel,
batch tile, select kernel, radix sort, reverse links), an ef-bounded search, the exact search on both kernels, and a
two-shard stitch.  `python scripts/sanitizer_case.py [rows]` -- exits non-zero when an answer differs from the oracle."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O  # noqa: E402  (the checker)
from rad_b200 import synth  # noqa: E402
from rad_b200.index import Index  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
db = synth.clustered_fingerprints(n, seed=5)
q = synth.queries_from_db(db, 256, seed=6)
ix = Index(ndim=1024, connectivity=8, expansion_add=128, expansion_search=64)
ix.add(np.arange(n, dtype=np.uint64), db)
slots, dists, _ = O.exact_search(db, q, 10)
for kernel in ("popc", "i8"):
    os.environ["RB_EXACT_KERNEL"] = kernel
    got = ix.search(q, 10, exact=True)
    assert np.array_equal(got.keys, slots) and np.array_equal(got.distances.view(np.uint32), dists.view(np.uint32)), kernel
del os.environ["RB_EXACT_KERNEL"]
approx = ix.search(q, 10)
recall = float((approx.distances <= dists[:, 9:10]).mean())
assert recall > 0.9, recall
nb = ix.get_neighbors(0, 0)
print(f"sanitizer case ok: rows={n} recall={recall:.3f} neighbours(0)={len(nb)}")
