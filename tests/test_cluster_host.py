"""Index.cluster's bookkeeping (level choice, refinement, merging of centroids: rad_b200/cluster.py) against the
answers of the unmodified reference on the reference-built golden graph (tests/golden/ref_cluster.npz,
scripts/make_cluster_golden.py) -- no GPU: the descents come from the CPU statement tests/model/cluster_model.py here;
tests/test_gpu_cluster.py runs the same cases with rb_cluster_device."""
from rad_b200 import cluster as cluster_module
from rad_b200.index import Index
from tests import cluster_cases as cases
from tests.model import cluster_model


def test_cluster_bookkeeping_reproduces_the_reference(monkeypatch):
    index = Index.restore(cases.GRAPH)
    vectors = cases.indexed_vectors()
    monkeypatch.setattr(cluster_module, "descend",
                        lambda ix, q, level, batch=0: cluster_model.descend(ix, vectors, q, level))
    cases.check_against_the_reference(index)
