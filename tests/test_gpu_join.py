"""Index.join on the GPU: rb_stable_match_device against the one-thread statement of the
reference's loop (tests/model/join_model.py, itself pinned to the reference by tests/test_join_host.py), and
Index.join(exact=True) against the committed answers of the unmodified reference (tests/golden/ref_join.npz)."""
import numpy as np
import pytest

from tests.join_cases import KEY_BASE, golden_cases, sets
from tests.model.join_model import sequential_join

pytestmark = pytest.mark.gpu


def _match_on_gpu(slots, dists, counts, women):
    import torch
    from rad_b200 import join as J
    dev = torch.device("cuda", 0)
    m2w, w2m, info = J.stable_match_device(torch.from_numpy(slots.astype(np.int32)).to(dev).contiguous(),
                                           torch.from_numpy(dists.astype(np.float32)).to(dev).contiguous(),
                                           torch.from_numpy(counts.astype(np.int32)).to(dev).contiguous(), women)
    return m2w.cpu().numpy().astype(np.int64), w2m.cpu().numpy().astype(np.int64), info


@pytest.mark.parametrize("men,women,proposals,spread", [(1, 1, 1, 1), (5, 3, 2, 3), (300, 300, 4, 40), (5000, 7000, 8, 7000),
                                                        (5000, 7000, 3, 50), (20000, 20000, 12, 2000), (4000, 9, 9, 9)])
def test_matching_equals_the_one_thread_statement(men, women, proposals, spread):
    """random candidate lists with heavy collisions (spread = how many distinct women are ever named), coarse distances
    (many ties) and ragged list lengths"""
    rng = np.random.default_rng(men * 31 + proposals)
    proposals = min(proposals, women)
    slots = np.stack([rng.choice(min(spread, women), size=proposals, replace=False) for _ in range(men)]).astype(np.int64)
    dists = np.sort(rng.integers(0, 12, size=(men, proposals)).astype(np.float32) / 16, axis=1)
    counts = rng.integers(0, proposals + 1, size=men) if men > 5 else np.full(men, proposals)
    want_m, want_w, want_e = sequential_join(slots, dists, counts, women, proposals)
    got_m, got_w, info = _match_on_gpu(slots, dists, counts, women)
    assert np.array_equal(got_m, want_m)
    assert np.array_equal(got_w, want_w)
    assert info["engagements"] == want_e


def test_everybody_names_the_same_woman_first():
    men, women = 3000, 3000
    slots = np.stack([np.array([0, 1 + m % 7, 8 + m]) % women for m in range(men)]).astype(np.int64)
    dists = np.tile(np.array([0.25, 0.5, 0.75], dtype=np.float32), (men, 1))
    dists[:, 0] += (np.arange(men) % 5).astype(np.float32) / 64     # five distance classes at the first woman
    counts = np.full(men, 3)
    want_m, want_w, _ = sequential_join(slots, dists, counts, women, 3)
    got_m, got_w, _ = _match_on_gpu(slots, dists, counts, women)
    assert np.array_equal(got_m, want_m) and np.array_equal(got_w, want_w)


def _index(vectors, key_base=0):
    from rad_b200.index import Index
    ix = Index(ndim=1024, connectivity=8, expansion_add=400, expansion_search=64)
    ix.add(np.arange(len(vectors), dtype=np.uint64) + key_base, vectors)
    return ix


def test_exact_join_equals_the_reference_join():
    built = {}
    for kind, n_a, n_b, p, pairs in golden_cases():
        if (kind, n_a, n_b) not in built:
            va, vb = sets(kind, n_a, n_b)
            built[(kind, n_a, n_b)] = (_index(va), _index(vb, KEY_BASE))
        a, b = built[(kind, n_a, n_b)]
        got = a.join(b, max_proposals=p, exact=True)
        assert sorted(got.items()) == [tuple(x) for x in pairs.tolist()], (kind, n_a, n_b, p)


def test_join_through_the_graph_search_and_argument_errors():
    from rad_b200 import join as J
    va, vb = sets("dense", 20000, 30000, seed=8)
    a, b = _index(va), _index(vb, KEY_BASE)
    exact = a.join(b, max_proposals=8, exact=True)
    stats = {}
    approx = J.join(a, b, max_proposals=8, exact=False, stats=stats)
    assert stats["proposals"] == 8 and stats["pairs"] == len(approx) and stats["generations"] >= 1
    # one-to-one, keys of the right sides
    assert len(set(approx.values())) == len(approx)
    assert all(k < KEY_BASE <= v for k, v in approx.items())
    # the ef-bounded candidates are nearly the exact ones: about as many pairs, mostly the same ones
    assert len(approx) >= 0.97 * len(exact)
    same = sum(1 for k, v in approx.items() if exact.get(k) == v)
    assert same >= 0.80 * len(exact)
    # every matched pair is one of the man's candidates
    found = b.search(va[:200], 8, exact=False)
    for m in range(200):
        if m in approx:
            assert approx[m] in set(np.asarray(found.keys[m]).tolist())
    # the reverse call is the same matching, mirrored (the smaller index plays the men either way)
    assert {v: k for k, v in b.join(a, max_proposals=8, exact=True).items()} == exact
    with pytest.raises(RuntimeError):
        a.join(a)
    calls = []

    def report(done: int, total: int) -> bool:   # the reference checks the annotations (index.py:32-36)
        calls.append((done, total))
        return True

    assert a.join(b, max_proposals=1, exact=True, progress=report)
    assert calls and calls[-1][0] == calls[-1][1]
    with pytest.raises(AssertionError):
        a.join(b, progress=lambda done, total: True)
