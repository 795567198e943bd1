"""The oracle (oracle/rad_oracle.c + numpy twin) against the golden vectors generated from the UNMODIFIED reference
(scripts/make_golden.py) and, when oracle/_ref is present, against the reference itself.  CPU only."""
import os

import numpy as np
import pytest

from oracle import oracle as O
from rad_b200 import synth


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def test_known_answer_distances(golden_dir):
    g = np.load(os.path.join(golden_dir, "tanimoto_pairs.npz"))
    a, b, want = g["a"], g["b"], g["dist_bits"]
    got = np.array([_bits(np.array([O.tanimoto(a[i], b[i])]))[0] for i in range(len(a))], dtype=np.uint32)
    assert np.array_equal(got, want)
    assert want[0] == np.float32(1.0).view(np.uint32)  # u == 0 -> 1 (binary.h:104)
    assert want[1] == 0                                   # identical fingerprints
    m = O.tanimoto_matrix(a[:64], b[:64])
    assert np.array_equal(_bits(np.diagonal(m).copy()), want[:64])


def test_levels_match_reference_stream(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_n1000_m8.npz"))
    lv, state = O.levels(1000, 8)
    assert np.array_equal(lv, g["levels"])
    more, _ = O.levels(10, 8, state)
    assert np.array_equal(O.levels(1010, 8)[0][1000:], more)  # the stream continues across calls


@pytest.mark.parametrize("name,ndim,m,ef", [("ref_n1000_m8", 1024, 8, 64), ("ref_n64_ndim64_m4", 64, 4, 20)])
def test_oracle_reproduces_reference_file_and_answers(golden_dir, tmp_path, name, ndim, m, ef):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    ref_bytes = open(os.path.join(golden_dir, name + ".usearch"), "rb").read()
    n = int(g["size"])
    rows = np.frombuffer(ref_bytes[8:8 + n * ndim // 8], dtype=np.uint8).reshape(n, ndim // 8)
    ix = O.OracleIndex(ndim, m, ef, 64)
    ix.add(np.arange(n, dtype=np.uint64), rows)
    out = tmp_path / "oracle.usearch"
    ix.save(str(out))
    assert out.read_bytes() == ref_bytes  # graph, levels, head, vectors: byte-identical to the reference build
    assert ix.max_level == int(g["max_level"])
    q = g["queries"]
    slots, dists, counts = O.exact_search(rows, q, 10)
    assert np.array_equal(slots, g["exact_keys"]) and np.array_equal(_bits(dists), g["exact_dist_bits"])
    assert np.array_equal(counts, g["exact_counts"])
    if ndim == 1024:
        k2, d2 = O.exact_topk_numpy(q, rows, 10)
        assert np.array_equal(k2, g["exact_keys"]) and np.array_equal(_bits(d2), g["exact_dist_bits"])
    keys, _, ad, ac = ix.search(q, 10)
    assert np.array_equal(ac, g["approx_counts"])
    for i in range(len(q)):
        c = int(ac[i])
        assert np.array_equal(keys[i, :c], g["approx_keys"][i, :c])
        assert np.array_equal(_bits(ad)[i, :c], g["approx_dist_bits"][i, :c])
    for s in g["probe"]:
        for l in range(int(g["levels"][s]) + 1):
            assert np.array_equal(ix.get_neighbors(int(s), l), g[f"nb_{s}_{l}"][0::2].astype(np.uint32))


def test_exact_tie_rule_descending_slot():
    # four identical rows: equal distances must come back highest slot first, and the highest slots win the cut
    db = synth.clustered_fingerprints(40, seed=5)
    db[10] = db[20] = db[30] = db[35]
    slots, dists, counts = O.exact_search(db, db[35:36], 3)
    assert list(slots[0]) == [35, 30, 20] and np.all(dists[0] == 0)
    k2, _ = O.exact_topk_numpy(db[35:36], db, 3)
    assert list(k2[0]) == [35, 30, 20]


def test_synth_is_slice_reproducible_and_has_edge_cases():
    full = synth.clustered_fingerprints(5000, seed=3)
    part = synth.clustered_fingerprints(700, seed=3, start=2100, n_total=5000)
    assert np.array_equal(full[2100:2800], part)
    assert not full[0].any() and np.array_equal(full[1], full[2])
    density = np.unpackbits(full, axis=1).mean()
    assert 0.03 < density < 0.08


@pytest.mark.ref
@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_against_live_reference(tmp_path):
    ui = O.ref_index_module()
    n, m, ef = 3000, 8, 400
    db = synth.clustered_fingerprints(n, seed=0)
    keys = np.arange(n, dtype=np.uint64)
    ref = ui.Index(ndim=1024, dtype="b1", metric="tanimoto", connectivity=m, expansion_add=ef, expansion_search=64)
    ref.add(keys, db, threads=1)
    orc = O.OracleIndex(1024, m, ef, 64)
    orc.add(keys, db)
    ref.save(str(tmp_path / "r.usearch"))
    orc.save(str(tmp_path / "o.usearch"))
    assert (tmp_path / "r.usearch").read_bytes() == (tmp_path / "o.usearch").read_bytes()
    q = synth.queries_from_db(db, 64)
    mres = ref.search(q, 10, exact=True)
    s, d, _ = O.exact_search(db, q, 10)
    assert np.array_equal(mres.keys, s) and np.array_equal(_bits(mres.distances), _bits(d))
