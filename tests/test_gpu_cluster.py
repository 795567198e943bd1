"""Index.cluster on the GPU: rb_cluster_device (the greedy descent of index.hpp:3046-3079,
one warp per query) on the reference-built golden graph against the answers of the unmodified reference, and on a
GPU-built graph against the CPU statement of the descent."""
import numpy as np
import pytest

from rad_b200 import synth
from tests import cluster_cases as cases
from tests.model import cluster_model

pytestmark = pytest.mark.gpu


def test_cluster_on_the_reference_graph_equals_the_reference():
    from rad_b200.index import Index
    cases.check_against_the_reference(Index.restore(cases.GRAPH))


def test_descent_on_a_gpu_built_graph_equals_the_cpu_statement():
    from rad_b200 import cluster as cluster_module
    from rad_b200.index import Index
    db = synth.clustered_fingerprints(60000, seed=4)
    ix = Index(ndim=1024, connectivity=8, expansion_add=400, expansion_search=64)
    ix.add(np.arange(len(db), dtype=np.uint64) + 7, db)
    assert ix.max_level >= 3
    q = synth.clustered_fingerprints(300, seed=99)
    q[0] = 0                                   # the all-zero query: distance 1 to everything but zero rows
    for level in (0, 1, 2, ix.max_level):
        slots, dists, visited, computed = cluster_module.descend(ix, q, level)
        want_s, want_d, want_v, want_c = cluster_model.descend(ix, db, q, level)
        assert np.array_equal(slots, want_s), level
        assert np.array_equal(dists.view(np.uint32), want_d.view(np.uint32)), level
        assert (visited, computed) == (want_v, want_c), level
        assert (np.asarray(ix.levels)[slots] >= max(level, 1)).all()
    # level 0 stops at level 1 like the reference (neighbors_non_base_)
    assert np.array_equal(cluster_module.descend(ix, q, 0)[0], cluster_module.descend(ix, q, 1)[0])
    # the user-facing call: every centroid is a key of a node of the chosen level, at most max_count of them
    c = ix.cluster(vectors=q, min_count=10, max_count=40)
    keys, sizes = c.centroids_popularity
    assert 1 <= len(keys) <= 40 and sizes.sum() == len(q)
    assert (np.asarray(ix.levels)[np.asarray(ix.get_node_ids_from_keys(keys))] >= 1).all()
    sub = ix.cluster(keys=np.arange(7, 507, dtype=np.uint64))
    assert len(sub.matches.keys) == 500 and repr(sub) == "usearch.Clustering(for 500 queries)"
    sub_keys, sub_sizes = sub.centroids_popularity
    members = sub.members_of(sub_keys[0])
    assert len(members) == sub_sizes[0] and set(members.tolist()) <= set(range(7, 507))
    again = sub.subcluster(sub_keys[0])
    assert len(again.matches.keys) == len(members)
    with pytest.raises(RuntimeError, match="Key missing"):
        ix.cluster(keys=np.array([1], dtype=np.uint64))
    with pytest.raises(AssertionError):
        ix.cluster(vectors=q, keys=np.array([7], dtype=np.uint64))


def test_too_small_an_index_cannot_cluster():
    from rad_b200.index import Index
    db = synth.clustered_fingerprints(20, seed=1)
    ix = Index(ndim=1024, connectivity=8)
    ix.add(None, db)
    if ix.max_level < 2:
        with pytest.raises(RuntimeError, match="too small"):
            ix.cluster(vectors=db)
