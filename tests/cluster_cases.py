"""Shared by the cluster tests: the cases of scripts/make_cluster_golden.py and the committed answers of the reference."""
import os

import numpy as np

from rad_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
GRAPH = os.path.join(HERE, "golden", "ref_n1000_m8.usearch")
GOLDEN = os.path.join(HERE, "golden", "ref_cluster.npz")
CASES = [("vectors_default", "vectors", None, None), ("vectors_5_1000", "vectors", 5, 1000),
         ("vectors_40_1000", "vectors", 40, 1000), ("keys_default", "keys", None, None),
         ("keys_subset_3_1000", "subset", 3, 1000), ("vectors_merge_5_8", "vectors", 5, 8),
         ("vectors_merge_40_25", "vectors", 40, 25)]


def indexed_vectors():
    return synth.clustered_fingerprints(1000, seed=7)   # scripts/make_golden.py


def queries():
    return synth.clustered_fingerprints(700, seed=21)


def run_case(index, what, lo, hi):
    if what == "vectors":
        return index.cluster(vectors=queries(), min_count=lo, max_count=hi)
    if what == "keys":
        return index.cluster(min_count=lo, max_count=hi)
    return index.cluster(keys=np.arange(0, 1000, 3, dtype=np.uint64), min_count=lo, max_count=hi)


def check_against_the_reference(index):
    g = np.load(GOLDEN)
    assert [s.nodes for s in index.levels_stats] == g["levels_nodes"].tolist()
    for name, what, lo, hi in CASES:
        c = run_case(index, what, lo, hi)
        keys = np.asarray(c.matches.keys).reshape(-1)
        bits = np.asarray(c.matches.distances, dtype=np.float32).reshape(-1).view(np.uint32)
        assert np.array_equal(keys, g[name + "_keys"]), name
        assert np.array_equal(bits, g[name + "_dist_bits"]), name
        assert (np.asarray(c.matches.counts) == 1).all()
