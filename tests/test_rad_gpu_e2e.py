"""BASELINE.json configs[0] end to end ON THE GPU BOX: 100 k synthetic fingerprints built by the CUDA kernels
(connectivity 8, expansion_add 400) -> `.usearch` file ->

  (a) SURVEY.md §8c rule 5, directly: the UNMODIFIED reference binding (oracle/_ref, compiled from the reference's own
      sources) loads the GPU-written file and returns identical get_neighbors for every (slot, level), identical
      levels_stats, identical exact-search answers (keys and float bits), and ef-bounded answers of the same quality;
  (b) the reference's UNMODIFIED create_local_hnsw_service (forks a server process around the index object:
      rad/hnsw_service.py:129-134) + RADTraverser.prime() / traverse(n_workers=1, n_to_score=10 000)
      (rad/traverser.py:128-245, the flow of the reference's tests/test_integration.py:89-288) run on the GPU-built
      rad_b200 Index and score the same molecules in the same order as the same walk over the reference binding's
      view of that file.

redis: neither the `redis` wheel nor a redis-server exists in this image, so tests/fake_redis.py (test infrastructure)
stands in for the client package; RAD's own code is untouched (BASELINE.md says so too)."""
import os
import sqlite3
import sys
import zlib

import numpy as np
import pytest

from oracle import oracle as O
from rad_b200 import synth
from rad_b200.index import Index
from tests.refpath import rad_root

pytestmark = pytest.mark.gpu
N, M, EF_ADD = 100_000, 8, 400


@pytest.fixture(scope="module")
def built(tmp_path_factory):
    """(our index, path of the file it wrote, fingerprints)"""
    db = synth.named_set("dense", N, seed=11)
    ix = Index(ndim=1024, dtype="b1", metric="tanimoto", connectivity=M, expansion_add=EF_ADD, expansion_search=64)
    ix.add(np.arange(N, dtype=np.uint64), db)
    path = str(tmp_path_factory.mktemp("e2e") / "gpu_100k.usearch")
    ix.save(path)
    return ix, path, db


@pytest.fixture(scope="module")
def ref_index(built):
    if not O.ref_available():
        pytest.skip("compiled reference binding (oracle/_ref) not available")
    _, path, _ = built
    return O.ref_index_module().Index(path=path)  # load: fingerprints included, so it can be searched


def test_reference_binding_reads_the_gpu_written_file(built, ref_index):
    ix, path, db = built
    ref = ref_index
    assert len(ref) == N and ref.max_level == ix.max_level and ref.connectivity == M and ref.ndim == 1024
    lv = ix.levels
    for slot in range(N):
        for level in range(int(lv[slot]) + 1):
            a, b = ix.get_neighbors(slot, level), ref.get_neighbors(slot, level)
            assert np.array_equal(np.asarray(a, dtype=np.uint64), np.asarray(b, dtype=np.uint64)), (slot, level)
    assert [int(x) for x in ix.get_top_level_nodes()] == [int(x) for x in ref.get_top_level_nodes()]
    mine, theirs = ix.levels_stats, ref.levels_stats
    assert len(mine) == len(theirs)
    for a, b in zip(mine, theirs):
        assert (a.nodes, a.edges, a.max_edges, a.allocated_bytes) == (b.nodes, b.edges, b.max_edges, b.allocated_bytes)
    # the out-of-range errors RAD's server relies on (rad/hnsw_server.py:481-487)
    for bad in ((N, 0), (0, int(lv[0]) + 1)):
        with pytest.raises(ValueError):
            ix.get_neighbors(*bad)
        with pytest.raises(ValueError):
            ref.get_neighbors(*bad)
    q = synth.queries_from_db(db, 500, seed=5)
    ours = ix.search(q, 10, exact=True)
    want = ref.search(q, 10, exact=True)
    assert np.array_equal(ours.keys, want.keys)
    assert np.array_equal(ours.distances.view(np.uint32), want.distances.view(np.uint32))
    # ef-bounded search of the SAME graph by the two implementations: same quality, nearly all rows bit-identical
    kth = want.distances[:, 9:10]
    for ef in (64, 128):
        ix.expansion_search = ef
        ref.expansion_search = ef
        g, r = ix.search(q, 10), ref.search(q, 10)
        ours_recall = float((g.distances <= kth).mean())
        ref_recall = float((r.distances <= kth).mean())
        assert ours_recall >= ref_recall - 0.005, (ef, ours_recall, ref_recall)
        same = float((g.distances.view(np.uint32) == r.distances.view(np.uint32)).mean())
        assert same >= 0.95, (ef, same)
    # and the other way round: the reference saves what it loaded, byte for byte what the GPU build wrote
    back = path + ".ref"
    ref.save(back)
    assert open(back, "rb").read() == open(path, "rb").read()


def score_fn(smiles, **_):
    """The mock of SURVEY.md §8d: a docking score that is a pure function of the SMILES string."""
    return (zlib.crc32(str(smiles).encode()) % 100000) / 1000.0 - 50.0


@pytest.fixture(scope="module")
def rad():
    root = rad_root()
    if root is None:
        pytest.skip("the reference's rad package is not staged (oracle/build_ref.sh)")
    from tests import fake_redis
    had = sys.modules.get("redis")
    sys.modules["redis"] = fake_redis
    sys.path.insert(0, root)
    try:
        import rad.hnsw_service as svc  # noqa: PLC0415
        import rad.traverser as trv  # noqa: PLC0415
        yield svc, trv
    finally:
        sys.path.remove(root)
        if had is None:
            sys.modules.pop("redis", None)
        else:
            sys.modules["redis"] = had


def _walk(svc, trv, index, db_path, host, n_to_score):
    service = svc.create_local_hnsw_service(index, database_path=db_path)
    traverser = trv.RADTraverser(hnsw_service=service, scoring_fn=score_fn, redis_host=host, redis_port=6379,
                                 namespace=f"ns_{host}")
    try:
        traverser.prime()
        traverser.traverse(n_workers=1, n_to_score=n_to_score, timeout=600)
        return traverser.get_molecules()
    finally:
        traverser.shutdown()
        service.shutdown()


def test_unmodified_rad_traversal_on_the_gpu_built_index(built, rad, tmp_path):
    ix, path, _ = built
    svc, trv = rad
    db_path = str(tmp_path / "nodes.db")
    with sqlite3.connect(db_path) as conn:
        conn.execute("CREATE TABLE nodes (node_key INTEGER PRIMARY KEY, smi TEXT)")
        conn.executemany("INSERT INTO nodes VALUES (?, ?)", [(k, f"C{k}N") for k in range(N)])
    n_to_score = 10_000
    mine = _walk(svc, trv, ix, db_path, "fake-gpu", n_to_score)  # the index object the GPU build left behind
    assert len(mine) >= n_to_score
    assert all(abs(s - score_fn(smi)) < 1e-9 for _, s, smi in mine)
    assert len({node for node, _, _ in mine}) == len(mine)  # nothing scored twice
    if O.ref_available():
        ref = O.ref_index_module().Index(path=path, view=True, exclude_vectors=True)  # scripts/start_hnsw_server.py:69
        theirs = _walk(svc, trv, ref, db_path, "fake-ref", n_to_score)
        common = min(len(mine), len(theirs))
        assert common >= n_to_score and mine[:common] == theirs[:common]
