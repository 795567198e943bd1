"""BASELINE.json configs[0], CPU part: the reference's UNMODIFIED RADTraverser (rad/traverser.py: prime + traverse)
walking a graph served by rad_b200.index.Index, beside the same traversal over the reference's usearch Index on the
same `.usearch` file.  One worker makes the walk deterministic, so the two scored sequences must be identical.

The container has no `redis` wheel and no redis-server: tests/fake_redis.py stands in for the client package (test
infrastructure only).  The walk over a GPU-BUILT index runs on the GPU box: tests/test_rad_gpu_e2e.py."""
import os
import sqlite3
import sys
import zlib

import pytest

from tests.refpath import rad_root

REFERENCE = rad_root()
pytestmark = pytest.mark.skipif(REFERENCE is None, reason="the reference's rad package is not present on this machine")


@pytest.fixture(scope="module")
def rad():
    from tests import fake_redis
    had = sys.modules.get("redis")
    sys.modules["redis"] = fake_redis
    sys.path.insert(0, REFERENCE)
    try:
        import rad.hnsw_service as svc  # noqa: PLC0415
        import rad.traverser as trv  # noqa: PLC0415
        yield svc, trv
    finally:
        sys.path.remove(REFERENCE)
        if had is None:
            sys.modules.pop("redis", None)
        else:
            sys.modules["redis"] = had


def score_fn(smiles, **_):
    """The mock of SURVEY.md §8d: a docking score that is a pure function of the SMILES string."""
    return (zlib.crc32(str(smiles).encode()) % 100000) / 1000.0 - 50.0


def _walk(svc, trv, index, db_path, host, n_to_score):
    service = svc.LocalHNSWService(index, database_path=db_path)
    traverser = trv.RADTraverser(hnsw_service=service, scoring_fn=score_fn, redis_host=host, redis_port=6379,
                                 namespace=f"ns_{host}")
    try:
        traverser.prime()
        traverser.traverse(n_workers=1, n_to_score=n_to_score, timeout=120)
        return traverser.get_molecules()
    finally:
        traverser.shutdown()
        service.shutdown()


def test_traversal_over_our_index_equals_traversal_over_the_reference_index(rad, golden_dir, tmp_path):
    svc, trv = rad
    from oracle import oracle as O
    if not O.ref_available():
        pytest.skip("compiled reference binding not available")
    from rad_b200.index import Index
    path = os.path.join(golden_dir, "ref_n1000_m8.usearch")
    ours = Index(path=path, view=True, exclude_vectors=True)
    ref = O.ref_index_module().Index(path=path, view=True, exclude_vectors=True)
    db_path = str(tmp_path / "nodes.db")
    with sqlite3.connect(db_path) as conn:
        conn.execute("CREATE TABLE nodes (node_key INTEGER PRIMARY KEY, smi TEXT)")
        conn.executemany("INSERT INTO nodes VALUES (?, ?)", [(k, f"C{k}N") for k in range(len(ours))])
    n_to_score = 300
    mine = _walk(svc, trv, ours, db_path, "fake-ours", n_to_score)
    theirs = _walk(svc, trv, ref, db_path, "fake-ref", n_to_score)
    assert len(mine) >= n_to_score and len(theirs) >= n_to_score
    # same nodes, same scores, same SMILES, in the same order, up to the point where the stop condition cut in
    common = min(len(mine), len(theirs))
    assert mine[:common] == theirs[:common]
    assert all(abs(s - score_fn(smi)) < 1e-9 for _, s, smi in mine)
    assert len({node for node, _, _ in mine}) == len(mine)  # nothing scored twice
