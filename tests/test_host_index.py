"""Host side of the drop-in through the C ABI: .usearch reader / writer and the graph reads RAD performs, against the
reference's own files and answers (tests/golden).  CPU only -- none of these calls touches CUDA."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from rad_b200 import _native
from rad_b200.index import Index

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(native):
    header = open(os.path.join(ROOT, "include", "rad_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(rb_[a-z0-9_]+)\s*\(", header)))
    declared = [d for d in declared if d not in ("rb_progress_t",)]
    out = subprocess.check_output(["nm", "-D", "--defined-only", _native.LIB_PATH], text=True)
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
    assert [d for d in declared if d not in exported] == []
    assert sorted(_native.SIGNATURES) == declared
    assert b"rad_b200" in native.rb_version()


@pytest.mark.parametrize("name,ndim,m", [("ref_n1000_m8", 1024, 8), ("ref_n64_ndim64_m4", 64, 4)])
@pytest.mark.parametrize("mode", ["load", "view", "view_exclude"])
def test_reads_reference_file(native, golden_dir, tmp_path, name, ndim, m, mode):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    path = os.path.join(golden_dir, name + ".usearch")
    ix = Index(ndim=ndim, connectivity=m, path=path, view=mode != "load", exclude_vectors=mode == "view_exclude")
    assert len(ix) == int(g["size"]) and ix.max_level == int(g["max_level"])
    assert ix.connectivity == m and ix.ndim == ndim and ix.multi is False
    assert str(ix.dtype) == "ScalarKind.B1"
    assert np.array_equal(ix.levels, g["levels"])
    assert np.array_equal(ix.get_top_level_nodes(), g["top_level_nodes"])
    for s in g["probe"]:
        for l in range(int(g["levels"][s]) + 1):
            assert np.array_equal(ix.get_neighbors(int(s), l), g[f"nb_{s}_{l}"])
    st = np.array([[s.nodes, s.edges, s.max_edges, s.allocated_bytes] for s in ix.levels_stats], dtype=np.int64)
    assert np.array_equal(st, g["level_stats"])
    assert np.array_equal(ix.get_keys_from_node_ids(np.arange(5)), np.arange(5, dtype=np.uint64))
    assert list(ix.get_node_ids_from_keys([3, 10**9])) == [3, -1]
    assert ix.contains(7) and not ix.contains(10**9)
    # the writer reproduces the reference's bytes (with vectors when we have them)
    if mode != "view_exclude":
        out = tmp_path / "again.usearch"
        ix.save(str(out))
        assert out.read_bytes() == open(path, "rb").read()
        assert ix.serialized_length == os.path.getsize(path)
        row = ix.get(2)
        raw = open(path, "rb").read()
        assert bytes(row) == raw[8 + 2 * ndim // 8: 8 + 3 * ndim // 8]


def test_graph_read_errors_and_immutability(native, golden_dir):
    path = os.path.join(golden_dir, "ref_n1000_m8.usearch")
    ix = Index(ndim=1024, connectivity=8, path=path, view=True, exclude_vectors=True)
    lv = ix.levels
    with pytest.raises(ValueError):
        ix.get_neighbors(len(ix), 0)
    with pytest.raises(ValueError, match="out of range for node"):
        ix.get_neighbors(0, int(lv[0]) + 1)
    with pytest.raises(ValueError):
        ix.get_neighbors(0, -1)
    with pytest.raises(RuntimeError, match="immutable"):
        ix.add(None, np.zeros((1, 128), dtype=np.uint8))
    with pytest.raises(RuntimeError):
        ix.search(np.zeros(128, dtype=np.uint8), 3)
    md = Index.metadata(path)
    assert md["dimensions"] == 1024 and md["count_present"] == 1000 and md["matrix_included"]
    assert str(md["kind_metric"]) == "MetricKind.Tanimoto" and md["version"] == "2.17.2"
    assert Index.metadata(path + ".missing") is None
    with pytest.raises(RuntimeError):
        Index(ndim=1024, connectivity=8).load(path + ".missing")
    with pytest.raises(RuntimeError):
        bad = os.path.join(os.path.dirname(path), "tanimoto_pairs.npz")
        Index(ndim=1024, connectivity=8).load(bad)


def test_argument_validation_without_gpu(native):
    with pytest.raises(ValueError):
        Index(ndim=1024, metric="cos")
    with pytest.raises(ValueError):
        Index(ndim=1024, dtype="f32")
    with pytest.raises(ValueError):
        Index(ndim=2048)
    with pytest.raises(ValueError):
        Index(ndim=1024, connectivity=64)
    ix = Index(ndim=1024, connectivity=8, expansion_add=400)
    assert len(ix) == 0 and ix.max_level == 0 and ix.expansion_add == 400
    ix.expansion_search = 128
    assert ix.expansion_search == 128
    with pytest.raises(ValueError):
        ix.add(None, np.zeros((2, 64), dtype=np.uint8))
    with pytest.raises(ValueError):
        ix.add(None, np.zeros((2, 128), dtype=np.float32))
    assert len(ix.get_top_level_nodes()) == 0
    empty = ix.search(np.zeros((3, 128), dtype=np.uint8), 5)
    assert list(empty.counts) == [0, 0, 0]


def test_reads_survive_fork(native, golden_dir):
    """rad/hnsw_service.py:129-134 forks after the index exists; the child answers from host memory."""
    import multiprocessing as mp
    path = os.path.join(golden_dir, "ref_n1000_m8.usearch")
    g = np.load(os.path.join(golden_dir, "ref_n1000_m8.npz"))
    ix = Index(ndim=1024, connectivity=8, path=path)
    ctx = mp.get_context("fork")
    q = ctx.Queue()

    def child():
        q.put((ix.get_neighbors(0, 0).tolist(), ix.get_top_level_nodes().tolist(), len(ix), ix.max_level))

    p = ctx.Process(target=child)
    p.start()
    nb, top, n, ml = q.get(timeout=30)
    p.join(30)
    assert nb == g["nb_0_0"].tolist() and top == g["top_level_nodes"].tolist() and n == 1000 and ml == int(g["max_level"])


def test_batched_neighbour_reads_equal_single_reads(native, golden_dir):
    """rb_neighbors_many (additive API): every list equals the single-call answer; out-of-range
    requests come back empty instead of raising."""
    import os

    import numpy as np

    from rad_b200.index import Index
    ix = Index(path=os.path.join(golden_dir, "ref_n1000_m8.usearch"), view=True, exclude_vectors=True)
    gold = np.load(os.path.join(golden_dir, "ref_n1000_m8.npz"))
    levels = gold["levels"]
    slots = np.arange(len(levels))
    pairs, off = ix.get_neighbors_many(slots)
    assert len(off) == len(slots) + 1 and off[0] == 0
    for s in range(0, len(slots), 11):
        assert (pairs[off[s]:off[s + 1]] == ix.get_neighbors(int(s), 0)).all()
    upper = np.nonzero(levels > 0)[0]
    pairs, off = ix.get_neighbors_many(upper, levels[upper])
    for i, s in enumerate(upper):
        assert (pairs[off[i]:off[i + 1]] == ix.get_neighbors(int(s), int(levels[s]))).all()
    pairs, off = ix.get_neighbors_many([0, len(levels) + 7, 1], [0, 0, 99])
    assert off[1] > 0 and off[2] == off[1] and off[3] == off[2]
