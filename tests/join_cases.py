"""Shared by the join tests: the seeded sets of scripts/make_join_golden.py and the committed answers of the reference."""
import os

import numpy as np

from rad_b200 import synth

KEY_BASE = 1_000_000
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_join.npz")


def sets(kind, n_a, n_b, seed=5):
    db = synth.named_set(kind, n_a + n_b, seed=seed)
    db = db[np.random.default_rng(1).permutation(len(db))]
    return db[:n_a], db[n_a:]


def golden_cases():
    g = np.load(GOLDEN)
    c = 0
    while f"case{c}_pairs" in g:
        n_a, n_b, p = (int(x) for x in g[f"case{c}_spec"])
        yield str(g[f"case{c}_kind"]), n_a, n_b, p, g[f"case{c}_pairs"]
        c += 1
