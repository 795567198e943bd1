"""rad_b200 -- B200-native drop-in for RAD's data-parallel hot path (HNSW build + Tanimoto k-NN over packed
fingerprints).  ``from rad_b200.index import Index`` mirrors ``usearch.index.Index(dtype='b1', metric='tanimoto')``;
put ``rad_b200/shim`` on ``sys.path`` to serve ``from usearch.index import Index`` to unmodified RAD scripts."""
from .index import BatchMatches, Index, Match, Matches, MetricKind, ScalarKind  # noqa: F401

__version__ = "0.1.0"
