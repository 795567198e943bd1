"""Fingerprint ingest on the GPU: duplicates and popcounts through the C ABI."""
import numpy as np
import pytest

from rad_b200 import synth

pytestmark = pytest.mark.gpu


def _first_occurrence(db: np.ndarray) -> np.ndarray:
    """numpy statement of rb_dedup_device: the lowest row with the same bytes"""
    rows = np.ascontiguousarray(db).view(np.dtype((np.void, db.shape[1]))).ravel()
    _, first, inverse = np.unique(rows, return_index=True, return_inverse=True)
    return first[inverse].astype(np.uint32)


@pytest.mark.parametrize("n", [1, 7, 4097, 60000])
def test_duplicate_representatives_equal_the_numpy_statement(n):
    from rad_b200 import ingest
    db = synth.analogue_set(n, seed=11)  # runs of 400 exact duplicates, row 1 == row 2, the all-zero row
    db[n // 3] = db[0]                   # far-apart copies as well
    if n > 10:
        db[n - 1] = db[5]
    rep = ingest.duplicate_representatives(db)
    assert np.array_equal(rep, _first_occurrence(db))


def test_every_row_identical():
    from rad_b200 import ingest
    db = np.tile(synth.clustered_fingerprints(1, seed=3), (5000, 1))
    assert np.array_equal(ingest.duplicate_representatives(db), np.zeros(5000, dtype=np.uint32))


def test_short_fingerprints_and_popcounts():
    from rad_b200 import ingest
    rng = np.random.default_rng(5)
    db = rng.integers(0, 256, size=(3000, 64), dtype=np.uint8)
    db[100:200] = db[7]
    assert np.array_equal(ingest.duplicate_representatives(db), _first_occurrence(db))
    assert np.array_equal(ingest.popcounts(db), np.unpackbits(db, axis=1).sum(axis=1).astype(np.uint16))


def test_deduplicated_add_keeps_every_row_and_every_exact_answer():
    from rad_b200.index import Index
    n = 60000
    db = synth.analogue_set(n, seed=4)
    keys = np.arange(n, dtype=np.uint64) * 3 + 1
    plain = Index(ndim=1024, connectivity=8, expansion_add=128, expansion_search=64)
    plain.add(keys, db)
    dd = Index(ndim=1024, connectivity=8, expansion_add=128, expansion_search=64)
    dd.add(keys, db, dedup=True)
    assert len(dd) == n and np.array_equal(dd.keys, keys)
    q = synth.queries_from_db(db, 500, seed=2)
    a, b = plain.search(q, 10, exact=True), dd.search(q, 10, exact=True)
    assert np.array_equal(a.keys, b.keys) and np.array_equal(a.distances.view(np.uint32), b.distances.view(np.uint32))
    # the graph: every row reachable from the entry point on level 0; a duplicate spends at most two links
    rep = _first_occurrence(db)
    dups = np.flatnonzero(rep != np.arange(n))
    assert len(dups) > 1000
    adj0, _ = dd._debug_adjacency()
    assert ((adj0[dups] != 0xFFFFFFFF).sum(axis=1) <= 2).all()
    assert np.array_equal(adj0[dups, 0], rep[dups])

    def reachable(ix):
        rows, _ = ix._debug_adjacency()
        seen = np.zeros(n, dtype=bool)
        frontier = np.array([ix.entry_slot])
        seen[frontier] = True
        while len(frontier):
            nb = rows[frontier].ravel()
            nb = np.unique(nb[nb != 0xFFFFFFFF])
            frontier = nb[~seen[nb]]
            seen[frontier] = True
        return seen

    seen = reachable(dd)
    assert seen.mean() >= reachable(plain).mean() - 0.001 and seen.mean() > 0.995
    # a duplicate is reachable whenever its representative is (the chain)
    assert np.array_equal(seen[dups], seen[rep[dups]])
    # recall of the ef-bounded search against the exact answers: not below the plain build
    truth = a.distances[:, 9:10]
    r_plain = float((plain.search(q, 10).distances <= truth).mean())
    r_dd = float((dd.search(q, 10).distances <= truth).mean())
    assert r_dd >= r_plain - 0.005, (r_dd, r_plain)
    # queries that ARE duplicated rows: the chained graph does not wade through cliques of 400 identical rows.  (In this
    # set a duplicated fingerprint is the ONLY member of its cluster, so without its copies it is a lone point far from
    # everything -- a smaller target for greedy routing than a clique of 400; its hit rate is reported, not gated.)
    qd = np.ascontiguousarray(db[dups[:: max(1, len(dups) // 300)]])
    hit_plain, hit_dd = plain.search(qd, 10), dd.search(qd, 10)
    print("d = 0 found:", float((hit_plain.distances[:, 0] == 0).mean()), float((hit_dd.distances[:, 0] == 0).mean()))
    found = hit_dd.distances[:, 0] == 0
    assert found.mean() > 0.5
    # ... and when the representative is found, the chain behind it delivers its copies
    assert (hit_dd.distances[found][:, :2] == 0).all()
