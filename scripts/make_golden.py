#!/usr/bin/env python3
"""Generate tests/golden/ from the UNMODIFIED reference binding (oracle/_ref, built by oracle/build_ref.sh from
/root/reference).  Run in the build container only; the fixtures travel, the reference does not.

Outputs:
  ref_n1000_m8.usearch   the reference's own file for 1000 synthetic fingerprints (threads=1, M=8, ef_add=64)
  ref_n1000_m8.npz       exact top-10 (keys, distance bits), approximate top-10 at ef=64, per-level stats,
                         neighbour lists of a few nodes, top-level nodes, levels
  ref_n64_ndim64_m4.usearch / .npz   a 64-bit / M=4 index like tests/test_hnsw_service.py of the reference builds
  tanimoto_pairs.npz     known-answer distances (float32 bits) for 512 fingerprint pairs incl. the u == 0 case
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from rad_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
os.makedirs(OUT, exist_ok=True)
ui = O.ref_index_module()


def dump(name, db, ndim, m, ef_add, queries):
    n = len(db)
    idx = ui.Index(ndim=ndim, dtype="b1", metric="tanimoto", connectivity=m, expansion_add=ef_add, expansion_search=64)
    idx.add(np.arange(n, dtype=np.uint64), db, threads=1)
    path = os.path.join(OUT, name + ".usearch")
    idx.save(path)
    ex = idx.search(queries, 10, exact=True, threads=1)
    ap = idx.search(queries, 10, threads=1)
    stats = np.array([[s.nodes, s.edges, s.max_edges, s.allocated_bytes] for s in idx.levels_stats], dtype=np.int64)
    levels = O.levels(n, m)[0]
    probe = [0, 1, 2, n // 2, n - 1] + [int(s) for s in np.flatnonzero(levels > 0)[:5]]
    nb = {}
    for s in probe:
        for l in range(int(levels[s]) + 1):
            nb[f"nb_{s}_{l}"] = np.asarray(idx.get_neighbors(s, l), dtype=np.uint64)
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"),
        queries=queries, exact_keys=ex.keys, exact_dist_bits=ex.distances.view(np.uint32), exact_counts=ex.counts,
        approx_keys=ap.keys, approx_dist_bits=ap.distances.view(np.uint32), approx_counts=ap.counts,
        level_stats=stats, levels=levels, max_level=np.int64(idx.max_level),
        top_level_nodes=np.asarray(idx.get_top_level_nodes(), dtype=np.uint64), size=np.int64(len(idx)),
        probe=np.array(probe, dtype=np.int64), **nb)
    print(name, "size", len(idx), "max_level", idx.max_level, os.path.getsize(path), "bytes")


db = synth.clustered_fingerprints(1000, seed=7)
dump("ref_n1000_m8", db, 1024, 8, 64, synth.queries_from_db(db, 32, seed=3))

rng = np.random.default_rng(42)
small = np.packbits(rng.random((64, 64)) < 0.3, axis=1)
small[5] = 0
dump("ref_n64_ndim64_m4", small, 64, 4, 20, small[:8].copy())

# known-answer distances through the reference's own metric (exact search of one row against one row)
a = synth.clustered_fingerprints(512, seed=11)
b = synth.clustered_fingerprints(512, seed=12)
a[0] = 0
b[0] = 0            # u == 0 -> 1.0
b[1] = a[1]         # identical -> 0.0
bits = np.empty(512, dtype=np.uint32)
for i in range(512):
    one = ui.Index(ndim=1024, dtype="b1", metric="tanimoto", connectivity=2)
    one.add(np.array([0], dtype=np.uint64), b[i:i + 1], threads=1)
    bits[i] = one.search(a[i], 1, exact=True).distances.view(np.uint32)[0]
np.savez_compressed(os.path.join(OUT, "tanimoto_pairs.npz"), a=a, b=b, dist_bits=bits)
print("tanimoto_pairs", bits[:4])
