"""The batched-build algorithm (tests/model/batched_model.c, the CPU statement of what the CUDA kernels compute)
against the reference algorithm (oracle/rad_oracle.c): recall@10 at equal M / ef within 0.5 points (SURVEY.md §8c
rule 5), plus the structural invariants.  CPU only -- this is how algorithmic choices are evaluated without a GPU;
tests/test_gpu_build.py then pins the kernels to the model bit for bit."""
import numpy as np
import pytest

from oracle import oracle as O
from rad_b200 import synth
from tests.model import model as M


@pytest.mark.parametrize("n,ef_add", [(12000, 128)])
def test_batched_algorithm_recall_vs_reference_algorithm(n, ef_add):
    db = synth.clustered_fingerprints(n, seed=3)
    q = synth.queries_from_db(db, 500, seed=9)
    _, exact_d, _ = O.exact_search(db, q, 10)
    lv, _ = O.levels(n, 8)
    model = M.BatchedModel(db, lv, connectivity=8, expansion_add=ef_add)
    model.build()
    ref = O.OracleIndex(1024, 8, ef_add, 64)
    ref.add(np.arange(n, dtype=np.uint64), db)
    assert model.max_level == ref.max_level and model.entry == ref.entry_slot
    for ef in (64, 128):
        _, d, c = model.search(q, 10, ef)
        assert (c == 10).all()
        ours = float((d <= exact_d[:, 9:10]).mean())
        theirs = float((ref.search(q, 10, ef)[2] <= exact_d[:, 9:10]).mean())
        assert ours >= theirs - 0.005, (ef, ours, theirs)
    adj0 = model.adj0()
    cnt = (adj0 != M.EMPTY).sum(axis=1)
    assert cnt.min() >= 1 and cnt.max() <= 16
    assert not (adj0 == np.arange(n, dtype=np.uint32)[:, None]).any()


def test_model_work_is_comparable_to_the_reference():
    """The roofline uses the REFERENCE's distance counts; the batched algorithm must not do wildly
    more traversal work than the reference for the same input."""
    n = 6000
    db = synth.clustered_fingerprints(n, seed=13)
    lv, _ = O.levels(n, 8)
    model = M.BatchedModel(db, lv, connectivity=8, expansion_add=64)
    model.build()
    ref = O.OracleIndex(1024, 8, 64, 64)
    ref.add(np.arange(n, dtype=np.uint64), db)
    ratio = model.counters()["computed_distances"] / ref.counters()["computed_distances"]
    assert 0.5 < ratio < 1.5, ratio
