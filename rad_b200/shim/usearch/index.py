"""``usearch.index`` as served by rad_b200 (see ``rad_b200/index.py`` for the reference anchors)."""
from rad_b200.index import *  # noqa: F401,F403
from rad_b200.index import (BatchMatches, Index, IndexStats, Key, Match, Matches, MetricKind, NodeID,  # noqa: F401
                            ScalarKind, DEFAULT_CONNECTIVITY, DEFAULT_EXPANSION_ADD, DEFAULT_EXPANSION_SEARCH,
                            USES_OPENMP, USES_SIMSIMD, USES_FP16LIB)
