"""The join oracle (tests/model/join_model.py: the reference's loop on one thread) against the answers of the unmodified
reference (tests/golden/ref_join.npz, scripts/make_join_golden.py) -- no GPU.  The GPU tests compare
rb_stable_match_device with this model bit for bit (tests/test_gpu_join.py)."""
import numpy as np

from oracle import oracle as O
from tests.join_cases import KEY_BASE, golden_cases, sets
from tests.model.join_model import sequential_join


def test_sequential_model_reproduces_the_reference_join():
    cases = list(golden_cases())
    assert len(cases) >= 5
    for kind, n_a, n_b, p, pairs in cases:
        va, vb = sets(kind, n_a, n_b)
        men, women, swapped = (va, vb, False) if n_a <= n_b else (vb, va, True)   # index.hpp:4343-4353
        slots, dists, counts = O.exact_search(women, men, p)
        wife, husband, _ = sequential_join(slots.astype(np.int64), dists, counts, len(women), p)
        m = np.flatnonzero(wife >= 0)
        if swapped:   # men are b's rows (keys + KEY_BASE), women a's
            got = sorted(zip(wife[m].tolist(), (m + KEY_BASE).tolist()))
        else:
            got = sorted(zip(m.tolist(), (wife[m] + KEY_BASE).tolist()))
        assert got == [tuple(x) for x in pairs.tolist()], (kind, n_a, n_b, p)
        # one-to-one, and consistent both ways
        assert len(set(wife[m].tolist())) == len(m)
        assert (husband[wife[m]] == m).all()


def test_model_edge_cases():
    # everybody wants woman 0; ties keep the earlier man; ragged candidate lists
    slots = np.array([[0, 1], [0, 1], [0, 2], [0, 0]])
    dists = np.array([[0.5, 0.9], [0.5, 0.6], [0.25, 0.7], [0.5, 0.0]], dtype=np.float32)
    counts = np.array([2, 2, 2, 1])
    wife, husband, engagements = sequential_join(slots, dists, counts, 3, 2)
    assert wife.tolist() == [-1, 1, 0, -1] and husband.tolist() == [2, 1, -1]
    assert engagements == 3
