"""Drop-in for the subset of ``usearch.index`` that RAD's hot path uses, served by the sm_100a kernels behind
``librad_b200.so``.

Mirrors (same names, argument meaning and error behaviour) the Python half of the reference boundary,
``usearch/python/usearch/index.py`` of the usearch fork 2.17.2 vendored by keiserlab/rad @ a21d60b:
``Index`` (:470-1482: ``__init__`` :480-593, ``add`` :631-678, ``search`` :680-720, ``get_neighbors`` :722-723,
``get_top_level_nodes`` :725-726, ``get_keys_from_node_ids`` :728-741, ``get_node_ids_from_keys`` :743-756,
``save/load/view`` :1067-1148, properties :895-1059 and :1315-1386), ``Match`` / ``Matches`` / ``BatchMatches``
(:282-380) and ``Index.metadata`` / ``Index.restore`` (:595-626).

Only ``dtype='b1'`` with ``metric='tanimoto'`` exists here; anything else raises (no fallback, per the north star).
Features no RAD code path reaches (remove / rename / copy / join / cluster / multi / JIT metrics / buffer
serialisation) raise ``NotImplementedError`` instead of being emulated.
"""
from __future__ import annotations

import ctypes as C
import enum
import math
import os
from collections.abc import Sequence
from dataclasses import dataclass
from inspect import signature
from typing import Callable, Dict, Iterable, List, Optional, Union

import numpy as np

from . import _native

Key = np.uint64
NodeID = np.uint64
ProgressCallback = Callable[[int, int], bool]

DEFAULT_CONNECTIVITY = 16        # usearch/include/usearch/index.hpp:1335-1345
DEFAULT_EXPANSION_ADD = 128
DEFAULT_EXPANSION_SEARCH = 64

USES_OPENMP = False
USES_SIMSIMD = False
USES_FP16LIB = False
VERSION_MAJOR, VERSION_MINOR, VERSION_PATCH = 2, 17, 2   # the .usearch file version this package reads and writes

_VALUE_ERRORS = (
    "Expects a matrix", "The number of vector dimensions", "Node ", "Level ", "Keys ", "k must",
    "rad_b200 supports", "connectivity must", "expansion", "bloom_bits_log2", "The level is out of range",
)


class MetricKind(enum.Enum):
    Unknown = 0
    Tanimoto = ord("t")

    def __str__(self) -> str:  # pybind11 enums print as "MetricKind.Tanimoto"
        return f"MetricKind.{self.name}"


class ScalarKind(enum.Enum):
    Unknown = 0
    B1 = 1

    def __str__(self) -> str:
        return f"ScalarKind.{self.name}"


def _normalize_metric(metric) -> MetricKind:
    if metric is None:
        return MetricKind.Tanimoto
    if isinstance(metric, MetricKind):
        return metric
    name = getattr(metric, "name", metric)
    if isinstance(name, str) and name.lower() == "tanimoto":
        return MetricKind.Tanimoto
    if isinstance(metric, int) and metric == ord("t"):
        return MetricKind.Tanimoto
    raise ValueError(f"rad_b200 only implements metric='tanimoto' (got {metric!r}); there is no fallback")


def _normalize_dtype(dtype) -> ScalarKind:
    if dtype is None or dtype == "":
        return ScalarKind.B1  # bitwise metrics default to b1 (usearch index.py:44-50)
    if isinstance(dtype, ScalarKind):
        return dtype
    name = getattr(dtype, "name", dtype)
    if isinstance(name, str) and name.lower() in ("b1", "b1x8", "bits"):
        return ScalarKind.B1
    if isinstance(dtype, int) and dtype == 1:
        return ScalarKind.B1
    raise ValueError(f"rad_b200 only implements dtype='b1' (got {dtype!r}); there is no fallback")


def _match_signature(func, arg_types, ret_type) -> bool:
    """The reference's own callback check (index.py:32-36): exact annotations."""
    sig = signature(func)
    return [p.annotation for p in sig.parameters.values()] == arg_types and sig.return_annotation == ret_type


@dataclass
class Match:
    key: int
    distance: float

    def to_tuple(self) -> tuple:
        return self.key, self.distance


@dataclass
class Matches:
    keys: np.ndarray
    distances: np.ndarray
    visited_members: int = 0
    computed_distances: int = 0

    def __len__(self) -> int:
        return len(self.keys)

    def __getitem__(self, index: int) -> Match:
        if isinstance(index, int) and index < len(self):
            return Match(key=self.keys[index], distance=self.distances[index])
        raise IndexError(f"`index` must be an integer under {len(self)}")

    def to_list(self) -> List[tuple]:
        return [(int(k), float(d)) for k, d in zip(self.keys, self.distances)]

    def __repr__(self) -> str:
        return f"usearch.Matches({len(self)})"


@dataclass
class BatchMatches(Sequence):
    keys: np.ndarray
    distances: np.ndarray
    counts: np.ndarray
    visited_members: int = 0
    computed_distances: int = 0

    def __len__(self) -> int:
        return len(self.counts)

    def __getitem__(self, index: int) -> Matches:
        if isinstance(index, int) and index < len(self):
            return Matches(
                keys=self.keys[index, : self.counts[index]],
                distances=self.distances[index, : self.counts[index]],
                visited_members=self.visited_members // len(self),
                computed_distances=self.computed_distances // len(self),
            )
        raise IndexError(f"`index` must be an integer under {len(self)}")

    def to_list(self) -> List[tuple]:
        rows = [self[i] for i in range(len(self))]
        return [match.to_tuple() for matches in rows for match in matches]

    def mean_recall(self, expected: np.ndarray, count: Optional[int] = None) -> float:
        return self.count_matches(expected, count=count) / len(expected)

    def count_matches(self, expected: np.ndarray, count: Optional[int] = None) -> int:
        assert len(expected) == len(self)
        if count is None:
            count = self.keys.shape[1]
        if count == 1:
            return int(np.sum(self.keys[:, 0] == expected))
        return int(sum(expected[i] in self.keys[i, :count] for i in range(len(self))))

    def __repr__(self) -> str:
        return f"usearch.BatchMatches({np.sum(self.counts)} across {len(self)} queries)"


class Clustering:
    """What ``Index.cluster`` returns (reference: index.py:384-409): the centroid of every query."""

    def __init__(self, index: "Index", matches: BatchMatches, queries: Optional[np.ndarray] = None) -> None:
        if queries is None:
            queries = index.keys
        self.index = index
        self.queries = queries
        self.matches = matches

    def __repr__(self) -> str:
        return f"usearch.Clustering(for {len(self.queries)} queries)"

    @property
    def centroids_popularity(self):
        return np.unique(self.matches.keys, return_counts=True)

    def members_of(self, centroid) -> np.ndarray:
        return self.queries[self.matches.keys.flatten() == centroid]

    def subcluster(self, centroid, **clustering_kwards) -> "Clustering":
        return self.index.cluster(keys=self.members_of(centroid), **clustering_kwards)


@dataclass
class IndexStats:
    """Per-level statistics (usearch/python/lib.cpp:1007-1013)."""
    nodes: int
    edges: int
    max_edges: int
    allocated_bytes: int


class Index:
    """GPU-built HNSW index over packed fingerprints under the Tanimoto metric -- the reference's ``Index`` surface."""

    def __init__(
        self,
        *,
        ndim: int = 0,
        metric="tanimoto",
        dtype=None,
        connectivity: Optional[int] = None,
        expansion_add: Optional[int] = None,
        expansion_search: Optional[int] = None,
        multi: bool = False,
        path: Optional[os.PathLike] = None,
        view: bool = False,
        enable_key_lookups: bool = True,
        exclude_vectors: bool = False,
        device: int = -1,
        lazy_host_sync: bool = False,
        time_kernels: bool = False,
        batch_growth_den: int = 0,
        batch_max: int = 0,
        bloom_bits_log2: int = 0,
        _debug_max_batches: int = 0,
        _no_batch_tile: bool = False,
        _connectivity_base: int = 0,
    ) -> None:
        if connectivity is None:
            connectivity = DEFAULT_CONNECTIVITY
        if expansion_add is None:
            expansion_add = DEFAULT_EXPANSION_ADD
        if expansion_search is None:
            expansion_search = DEFAULT_EXPANSION_SEARCH
        assert isinstance(connectivity, int), "Expects integer `connectivity`"
        assert isinstance(expansion_add, int), "Expects integer `expansion_add`"
        assert isinstance(expansion_search, int), "Expects integer `expansion_search`"
        self._metric_kind = _normalize_metric(metric)
        self._dtype = _normalize_dtype(dtype)
        if multi:
            raise NotImplementedError("multi=True is outside RAD's path and not implemented by rad_b200")
        self._exclude_vectors = bool(exclude_vectors)
        self._enable_key_lookups = bool(enable_key_lookups)
        if not ndim:
            ndim = 1024
        opts = _native.InitOptions(
            dimensions=ndim, connectivity=connectivity, expansion_add=expansion_add,
            expansion_search=expansion_search, device=device,
            flags=(_native.RB_FLAG_LAZY_HOST_SYNC if lazy_host_sync else 0)
            | (_native.RB_FLAG_TIME_KERNELS if time_kernels else 0),
            batch_growth_den=batch_growth_den, batch_max=batch_max, bloom_bits_log2=bloom_bits_log2,
            debug_max_batches=_debug_max_batches, no_batch_tile=1 if _no_batch_tile else 0,
            connectivity_base=_connectivity_base)
        self._connectivity_base = int(_connectivity_base)
        err = C.c_char_p()
        self._lib = _native.lib()
        self._h = self._lib.rb_init(C.byref(opts), C.byref(err))
        if not self._h:
            raise ValueError((err.value or b"rb_init failed").decode())
        self.path = path
        if path is not None and os.path.exists(path):
            if view:
                self.view(path)
            else:
                self.load(path)

    # ------------------------------------------------------------------ lifetime
    def reset(self):
        h = getattr(self, "_h", None)
        if h:
            self._lib.rb_free(h)
            self._h = None

    def __del__(self):
        try:
            self.reset()
        except Exception:
            pass

    def _handle(self):
        if not self._h:
            raise RuntimeError("The index has been reset")
        return self._h

    def _check(self, err):
        _native.check(err, _VALUE_ERRORS)

    # ------------------------------------------------------------------ build
    def __len__(self) -> int:
        return int(self._lib.rb_size(self._handle()))

    def add(self, keys, vectors, *, copy: bool = True, threads: int = 0, log: Union[str, bool] = False,
            progress: Optional[ProgressCallback] = None, dedup: bool = False):
        """Insert one or many fingerprints (reference: index.py:631-678, :224-279, lib.cpp:223-268).
        ``threads`` is accepted and ignored: the build runs as batched CUDA kernels.  ``dedup=True`` (additive, empty
        index only) keeps exact duplicates out of the graph search and chains them behind their first occurrence
        (``rad_b200.ingest.add_deduplicated``)."""
        assert isinstance(vectors, np.ndarray), "Expects a NumPy array"
        if dedup:
            from . import ingest
            ingest.add_deduplicated(self, keys, vectors.reshape(1, -1) if vectors.ndim == 1 else vectors)
            return self.keys
        assert not progress or _match_signature(progress, [int, int], bool), "Invalid callback"
        assert vectors.ndim == 1 or vectors.ndim == 2, "Expects a matrix or vector"
        if vectors.ndim == 1:
            vectors = vectors.reshape(1, len(vectors))
        count = vectors.shape[0]
        if keys is None:
            start = len(self)
            keys = np.arange(start, start + count, dtype=Key)
        else:
            if not isinstance(keys, Iterable):
                assert count == 1, "Each vector must have a key"
                keys = [keys]
            keys = np.asarray(keys, dtype=Key)  # no copy when the caller already holds uint64 keys
        assert len(keys) == count
        if vectors.dtype != np.uint8:
            raise ValueError("Incompatible scalars in the vectors matrix: rad_b200 stores packed bits (numpy uint8)")
        if vectors.shape[1] != self.ndim // 8:
            raise ValueError("The number of vector dimensions doesn't match!")
        if vectors.strides[1] != 1:
            raise ValueError("Matrix rows must be contiguous, try `ascontiguousarray`.")
        keys = np.ascontiguousarray(keys)

        bar = None
        if log:
            from tqdm import tqdm
            bar = tqdm(desc=log if isinstance(log, str) else "Add", total=count, unit="vector")

        def on_progress(done, total, _user):
            try:
                if bar is not None:
                    bar.update(int(done) - bar.n)
                return 1 if (progress(int(done), int(total)) if progress else True) else 0
            except BaseException:  # an exception in a ctypes callback cannot propagate: cancel instead
                return 0

        cb = _native.PROGRESS(on_progress) if (progress or bar is not None) else C.cast(None, _native.PROGRESS)
        err = C.c_char_p()
        self._lib.rb_add(self._handle(), keys.ctypes.data, vectors.ctypes.data, count, vectors.strides[0],
                         1 if copy else 0, cb, None, C.byref(err))
        if bar is not None:
            bar.close()
        self._check(err)
        return keys

    # ------------------------------------------------------------------ search
    def search(self, vectors, count: int = 10, radius: float = math.inf, *, threads: int = 0, exact: bool = False,
               log: Union[str, bool] = False, progress: Optional[ProgressCallback] = None):
        """Batched k-NN (reference: index.py:680-720, :181-221, lib.cpp:511-557).  ``exact=True`` is the brute-force
        scan with the reference's tie rule (ascending distance, ties by descending slot)."""
        assert isinstance(vectors, np.ndarray), "Expects a NumPy array"
        assert vectors.ndim == 1 or vectors.ndim == 2, "Expects a matrix or vector"
        assert not progress or _match_signature(progress, [int, int], bool), "Invalid callback"
        single = vectors.ndim == 1
        if single:
            vectors = vectors.reshape(1, len(vectors))
        if vectors.dtype != np.uint8:
            raise ValueError("Incompatible scalars in the query matrix: rad_b200 stores packed bits (numpy uint8)")
        if vectors.shape[1] != self.ndim // 8:
            raise ValueError("The number of vector dimensions doesn't match!")
        if vectors.strides[1] != 1:
            raise ValueError("Matrix rows must be contiguous, try `ascontiguousarray`.")
        nq = vectors.shape[0]
        keys = np.zeros((nq, count), dtype=Key)
        distances = np.zeros((nq, count), dtype=np.float32)
        counts = np.zeros(nq, dtype=np.int64)
        visited, computed = C.c_uint64(0), C.c_uint64(0)
        err = C.c_char_p()
        self._lib.rb_search(self._handle(), vectors.ctypes.data, nq, vectors.strides[0], count, 1 if exact else 0,
                            keys.ctypes.data, distances.ctypes.data, counts.ctypes.data, C.byref(visited),
                            C.byref(computed), C.byref(err))
        self._check(err)
        if progress:
            progress(nq, nq)
        batch = BatchMatches(keys, distances, counts, int(visited.value), int(computed.value))
        return batch[0] if single else batch

    # ------------------------------------------------------------------ graph reads (what RAD calls at run time)
    def get_neighbors(self, node_id, level):
        """``[slot, key, slot, key, ...]`` of (node_id, level); ValueError when out of range (lib.cpp:270-288)."""
        out = np.empty(2 * 2 * int(self.connectivity), dtype=np.uint64)
        err = C.c_char_p()
        n = self._lib.rb_neighbors(self._handle(), int(node_id), int(level), out.ctypes.data, out.size, C.byref(err))
        self._check(err)
        return out[:n].copy()

    def get_neighbors_many(self, node_ids, levels=None):
        """Neighbour lists of many (node_id, level) pairs in one call (additive API).

        Returns ``(pairs, offsets)``: ``pairs[offsets[i]:offsets[i + 1]]`` is the ``[slot, key, ...]`` list of request
        ``i``, exactly what ``get_neighbors(node_ids[i], levels[i])`` returns; an out-of-range request gives an empty
        list instead of raising.  ``levels=None`` means level 0 for every node."""
        ids = np.ascontiguousarray(np.asarray(node_ids).astype(np.uint64).reshape(-1))
        lv = None if levels is None else np.ascontiguousarray(
            np.broadcast_to(np.asarray(levels, dtype=np.int32), ids.shape))
        out = np.empty(max(1, 4 * int(self.connectivity) * len(ids)), dtype=np.uint64)
        offsets = np.zeros(len(ids) + 1, dtype=np.uint64)
        err = C.c_char_p()
        self._lib.rb_neighbors_many(self._handle(), ids.ctypes.data, None if lv is None else lv.ctypes.data, len(ids),
                                    out.ctypes.data, out.size, offsets.ctypes.data, C.byref(err))
        self._check(err)
        return out[:int(offsets[-1])].copy(), offsets.astype(np.int64)

    def get_top_level_nodes(self):
        """``[slot, key, ...]`` of every node on the top level (lib.cpp:290-305)."""
        err = C.c_char_p()
        need = self._lib.rb_top_level_nodes(self._handle(), None, 0, C.byref(err))
        out = np.empty(max(int(need), 1), dtype=np.uint64)
        n = self._lib.rb_top_level_nodes(self._handle(), out.ctypes.data, out.size, C.byref(err))
        self._check(err)
        return out[:n].copy()

    def get_keys_from_node_ids(self, node_ids):
        one = not isinstance(node_ids, Iterable)
        ids = np.ascontiguousarray(np.array([node_ids] if one else node_ids).astype(NodeID))
        out = np.empty(len(ids), dtype=Key)
        err = C.c_char_p()
        self._lib.rb_keys_from_slots(self._handle(), ids.ctypes.data, len(ids), out.ctypes.data, C.byref(err))
        self._check(err)
        return out[0] if one else out

    def get_node_ids_from_keys(self, keys):
        one = not isinstance(keys, Iterable)
        ks = np.ascontiguousarray(np.array([keys] if one else keys).astype(Key))
        out = np.empty(len(ks), dtype=np.int64)
        err = C.c_char_p()
        self._lib.rb_slots_from_keys(self._handle(), ks.ctypes.data, len(ks), out.ctypes.data, C.byref(err))
        self._check(err)
        return out[0] if one else out

    def contains(self, keys):
        if isinstance(keys, Iterable):
            return np.array([bool(self._lib.rb_contains(self._handle(), int(k))) for k in keys], dtype=bool)
        return bool(self._lib.rb_contains(self._handle(), int(keys)))

    def __contains__(self, keys):
        return self.contains(keys)

    def count(self, keys):
        c = self.contains(keys)
        return c.astype(np.int64) if isinstance(c, np.ndarray) else int(c)

    def get(self, keys, dtype=None):
        """Stored fingerprints by key (reference: index.py:773-817); ``None`` for a missing key."""
        one = not isinstance(keys, Iterable)
        ks = [keys] if one else list(keys)
        slots = np.atleast_1d(self.get_node_ids_from_keys(np.array(ks, dtype=Key)))
        results = []
        for s in slots:
            if s < 0:
                results.append(None)
                continue
            out = np.empty(self.ndim // 8, dtype=np.uint8)
            sl = np.array([s], dtype=np.uint64)
            err = C.c_char_p()
            self._lib.rb_get_vectors(self._handle(), sl.ctypes.data, 1, out.ctypes.data, C.byref(err))
            self._check(err)
            results.append(out)
        return results[0] if one else tuple(results)

    def __getitem__(self, keys):
        return self.get(keys)

    # ------------------------------------------------------------------ unsupported mutations (outside RAD's path)
    def remove(self, *a, **k):
        raise NotImplementedError("remove() is outside RAD's path and not implemented by rad_b200")

    def rename(self, *a, **k):
        raise NotImplementedError("rename() is outside RAD's path and not implemented by rad_b200")

    def copy(self):
        raise NotImplementedError("copy() is outside RAD's path and not implemented by rad_b200")

    def join(self, other: "Index", max_proposals: int = 0, exact: bool = False,
             progress: Optional[ProgressCallback] = None) -> Dict[int, int]:
        """One-to-one matching between the rows of ``self`` and ``other``: mapping from keys of ``self`` to keys of
        ``other`` with no key used twice (reference: index.py:1182-1212, lib.cpp:876-894, index.hpp:4330-4515).  The
        candidate searches and the proposals run on the GPU (``rad_b200/join.py``)."""
        assert not progress or _match_signature(progress, [int, int], bool), "Invalid callback signature"
        from .join import join as _join
        return _join(self, other, max_proposals=max_proposals, exact=exact, progress=progress)

    def cluster(self, *, vectors: Optional[np.ndarray] = None, keys: Optional[np.ndarray] = None,
                min_count: Optional[int] = None, max_count: Optional[int] = None, threads: int = 0,
                log: Union[str, bool] = False, progress: Optional[ProgressCallback] = None) -> "Clustering":
        """Clusters already indexed or provided ``vectors``, mapping them to centroids taken from an upper level of the
        graph (reference: index.py:1214-1273, lib.cpp:758-875, index_dense.hpp:1919-2080; ``rad_b200/cluster.py``)."""
        assert not progress or _match_signature(progress, [int, int], bool), "Invalid callback signature"
        from .cluster import cluster as _cluster
        if vectors is not None:
            assert keys is None, "You can either cluster vectors or member keys"
            if vectors.ndim != 2:
                raise ValueError("Expects a matrix of queries to add!")
            if vectors.shape[1] != self.ndim // 8:
                raise ValueError("The number of vector dimensions doesn't match!")
            queries, members = np.ascontiguousarray(vectors, dtype=np.uint8), None
        else:
            members = np.asarray(self.keys if keys is None else keys).astype(Key)
            slots = np.atleast_1d(self.get_node_ids_from_keys(members))
            if (slots < 0).any():
                raise RuntimeError("Key missing!")                            # index_dense.hpp:865-866
            queries = np.empty((len(members), self.ndim // 8), dtype=np.uint8)
            err = C.c_char_p()
            sl = np.ascontiguousarray(slots, dtype=np.uint64)
            self._lib.rb_get_vectors(self._handle(), sl.ctypes.data, len(sl), queries.ctypes.data, C.byref(err))
            self._check(err)
        ckeys, dists, visited, computed, _ = _cluster(self, queries, int(min_count or 0), int(max_count or 0))
        n = len(queries)
        if progress:
            progress(n, n)
        matches = BatchMatches(ckeys.reshape(n, 1).astype(Key), dists.reshape(n, 1), np.ones(n, dtype=np.int64),
                               visited, computed)
        return Clustering(self, matches, members)

    def clear(self):
        """Erases all vectors from the index, preserving the allocated space (reference: index.py:1150-1152)."""
        err = C.c_char_p()
        self._lib.rb_clear(self._handle(), C.byref(err))
        self._check(err)

    # ------------------------------------------------------------------ properties (index.py:895-1059, 1315-1386)
    @property
    def jit(self) -> bool:
        return False

    @property
    def hardware_acceleration(self) -> str:
        return "sm_100a"

    @property
    def size(self) -> int:
        return len(self)

    @property
    def ndim(self) -> int:
        return int(self._lib.rb_dimensions(self._handle()))

    @property
    def serialized_length(self) -> int:
        return int(self._lib.rb_serialized_length(self._handle(), 1 if self._exclude_vectors else 0))

    @property
    def metric_kind(self) -> MetricKind:
        return self._metric_kind

    @property
    def metric(self) -> MetricKind:
        return self._metric_kind

    @property
    def dtype(self) -> ScalarKind:
        return self._dtype

    @property
    def connectivity(self) -> int:
        return int(self._lib.rb_connectivity(self._handle()))

    @property
    def capacity(self) -> int:
        return int(self._lib.rb_capacity(self._handle()))

    @property
    def memory_usage(self) -> int:
        return int(self._lib.rb_memory_usage(self._handle()))

    @property
    def expansion_add(self) -> int:
        return int(self._lib.rb_expansion_add(self._handle()))

    @expansion_add.setter
    def expansion_add(self, v: int):
        self._lib.rb_set_expansion_add(self._handle(), int(v))

    @property
    def expansion_search(self) -> int:
        return int(self._lib.rb_expansion_search(self._handle()))

    @expansion_search.setter
    def expansion_search(self, v: int):
        self._lib.rb_set_expansion_search(self._handle(), int(v))

    @property
    def max_level(self) -> int:
        return int(self._lib.rb_max_level(self._handle()))

    @property
    def nlevels(self) -> int:
        return self.max_level + 1

    @property
    def multi(self) -> bool:
        return False

    @property
    def entry_slot(self) -> int:
        return int(self._lib.rb_entry_slot(self._handle()))

    def level_stats(self, level: int) -> IndexStats:
        st = _native.LevelStats()
        err = C.c_char_p()
        self._lib.rb_level_stats(self._handle(), int(level), C.byref(st), C.byref(err))
        self._check(err)
        return IndexStats(int(st.nodes), int(st.edges), int(st.max_edges), int(st.allocated_bytes))

    @property
    def levels_stats(self) -> List[IndexStats]:
        return [self.level_stats(l) for l in range(self.max_level + 1)]

    @property
    def stats(self) -> IndexStats:
        per = self.levels_stats
        return IndexStats(per[0].nodes if per else 0, sum(s.edges for s in per), sum(s.max_edges for s in per),
                          sum(s.allocated_bytes for s in per))

    @property
    def levels(self) -> np.ndarray:
        out = np.empty(len(self), dtype=np.int16)
        self._lib.rb_levels(self._handle(), out.ctypes.data, out.size)
        return out

    @property
    def keys(self) -> np.ndarray:
        n = len(self)
        return self.get_keys_from_node_ids(np.arange(n, dtype=NodeID)) if n else np.zeros(0, dtype=Key)

    @property
    def last_counters(self) -> dict:
        """Work counters of the last add/search (the reference's add_result_t / search_result_t counters)."""
        c = _native.Counters()
        self._lib.rb_last_counters(self._handle(), C.byref(c))
        return dict(computed_distances=int(c.computed_distances), visited_members=int(c.visited_members),
                    batches=int(c.batches), kernel_launches=int(c.kernel_launches), device_ms=float(c.device_ms),
                    insert_ms=float(c.insert_ms), sort_ms=float(c.sort_ms), reverse_ms=float(c.reverse_ms),
                    insert_launches=int(c.insert_launches), tile_ms=float(c.tile_ms),
                    select_ms=float(c.select_ms))

    def sync_host(self) -> None:
        err = C.c_char_p()
        self._lib.rb_sync_host(self._handle(), C.byref(err))
        self._check(err)

    def release_device(self, keep_host_vectors: bool = False) -> None:
        """Free every device allocation; graph reads keep working from host memory."""
        err = C.c_char_p()
        self._lib.rb_release_device(self._handle(), 1 if keep_host_vectors else 0, C.byref(err))
        self._check(err)

    # ------------------------------------------------------------------ pieces of the cross-shard stitch (sharded.py)
    def seek_level_stream(self, first_row: int) -> None:
        """This EMPTY index is the shard starting at global row ``first_row``: its nodes draw the levels a one-piece
        build gives those rows (the reference's single level stream, index.hpp:3867-3871)."""
        err = C.c_char_p()
        self._lib.rb_seek_level_stream(self._handle(), int(first_row), C.byref(err))
        self._check(err)

    def add_device(self, keys, d_vectors_ptr: int, count: int, levels=None) -> None:
        """Insert ``count`` fingerprints that are already in HBM (``d_vectors_ptr``: device address of count x 128
        bytes).  ``levels`` (int16 per row) prescribes the levels instead of drawing them from the level stream."""
        err = C.c_char_p()
        kptr = None
        if keys is not None:
            keys = np.ascontiguousarray(keys, dtype=np.uint64)
            kptr = keys.ctypes.data
        if levels is None:
            self._lib.rb_add_device(self._handle(), kptr, d_vectors_ptr, int(count), C.byref(err))
        else:
            levels = np.ascontiguousarray(levels, dtype=np.int16)
            assert len(levels) == count
            self._lib.rb_add_device_levels(self._handle(), kptr, d_vectors_ptr, int(count), levels.ctypes.data,
                                           C.byref(err))
        self._check(err)

    def extract_upper_levels(self, source: "Index") -> None:
        """This EMPTY index (created with ``_connectivity_base=connectivity``) becomes ``source``'s nodes of level >= 1,
        one level down, with the links the build already gave them (``rb_extract_level``; device to device)."""
        err = C.c_char_p()
        self._lib.rb_extract_level(source._handle(), self._handle(), C.byref(err))
        self._check(err)

    def import_graph(self, keys, levels, adj0, adj_upper, vectors, entry: int) -> None:
        """Replace the graph by one assembled elsewhere (global slots): see ``rb_import_graph``."""
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        levels = np.ascontiguousarray(levels, dtype=np.int16)
        n = len(keys)
        adj0 = np.ascontiguousarray(adj0, dtype=np.uint32).reshape(n, -1)
        adj_upper = np.ascontiguousarray(adj_upper, dtype=np.uint32).reshape(-1, self.connectivity)
        assert adj0.shape[1] == self.connectivity_base and len(adj_upper) == int(levels.astype(np.int64).sum())
        vptr = None
        if vectors is not None:
            vectors = np.ascontiguousarray(vectors, dtype=np.uint8).reshape(n, -1)
            assert vectors.shape[1] == self.ndim // 8
            vptr = vectors.ctypes.data
        err = C.c_char_p()
        self._lib.rb_import_graph(self._handle(), keys.ctypes.data, levels.ctypes.data, adj0.ctypes.data,
                                  adj_upper.ctypes.data if len(adj_upper) else None, vptr, n, int(entry), C.byref(err))
        self._check(err)

    @property
    def connectivity_base(self) -> int:
        return self._connectivity_base or 2 * self.connectivity

    def _debug_adjacency(self):
        """(adj0 [n, 2M], adj_upper [rows, M]) with 0xFFFFFFFF padding -- tests compare these with the CPU model."""
        n, m = len(self), self.connectivity
        rows = int(self._lib.rb_debug_adjacency(self._handle(), None, None))
        adj0 = np.empty((n, self.connectivity_base), dtype=np.uint32)
        adju = np.empty((rows, m), dtype=np.uint32)
        self._lib.rb_debug_adjacency(self._handle(), adj0.ctypes.data, adju.ctypes.data if rows else None)
        return adj0, adju

    # ------------------------------------------------------------------ serialisation (index.py:1067-1148)
    def save(self, path_or_buffer=None, progress: Optional[ProgressCallback] = None):
        assert not progress or _match_signature(progress, [int, int], bool), "Invalid callback signature"
        path_or_buffer = path_or_buffer if path_or_buffer else self.path
        if path_or_buffer is None or isinstance(path_or_buffer, (bytes, bytearray)):
            raise NotImplementedError("buffer serialisation is outside RAD's path; pass a file path")
        err = C.c_char_p()
        self._lib.rb_save(self._handle(), os.fsencode(os.fspath(path_or_buffer)), 1 if self._exclude_vectors else 0,
                          C.byref(err))
        self._check(err)

    def load(self, path_or_buffer=None, progress: Optional[ProgressCallback] = None):
        assert not progress or _match_signature(progress, [int, int], bool), "Invalid callback signature"
        path_or_buffer = path_or_buffer if path_or_buffer else self.path
        if path_or_buffer is None:
            raise Exception("Define the source")
        if isinstance(path_or_buffer, (bytes, bytearray)):
            raise NotImplementedError("buffer serialisation is outside RAD's path; pass a file path")
        path_or_buffer = os.fspath(path_or_buffer)
        if not os.path.exists(path_or_buffer):
            raise RuntimeError("Missing file!")
        err = C.c_char_p()
        self._lib.rb_load(self._handle(), os.fsencode(path_or_buffer), 1 if self._exclude_vectors else 0,
                          C.byref(err))
        self._check(err)

    def view(self, path_or_buffer=None, progress: Optional[ProgressCallback] = None):
        assert not progress or _match_signature(progress, [int, int], bool), "Invalid callback signature"
        path_or_buffer = path_or_buffer if path_or_buffer else self.path
        if path_or_buffer is None:
            raise Exception("Define the source")
        if isinstance(path_or_buffer, (bytes, bytearray)):
            raise NotImplementedError("buffer serialisation is outside RAD's path; pass a file path")
        err = C.c_char_p()
        self._lib.rb_view(self._handle(), os.fsencode(os.fspath(path_or_buffer)), 1 if self._exclude_vectors else 0,
                          C.byref(err))
        self._check(err)

    @staticmethod
    def metadata(path_or_buffer) -> Optional[dict]:
        if isinstance(path_or_buffer, (bytes, bytearray)):
            raise NotImplementedError("buffer serialisation is outside RAD's path; pass a file path")
        path = os.fspath(path_or_buffer)
        if not os.path.exists(path):
            return None
        md = _native.Metadata()
        err = C.c_char_p()
        _native.lib().rb_metadata(os.fsencode(path), C.byref(md), C.byref(err))
        _native.check(err)
        return {
            "matrix_included": bool(md.has_vectors),
            "version": f"{md.version_major}.{md.version_minor}.{md.version_patch}",
            "kind_metric": MetricKind.Tanimoto if md.kind_metric == ord("t") else MetricKind.Unknown,
            "kind_scalar": ScalarKind.B1 if md.kind_scalar == 1 else ScalarKind.Unknown,
            "kind_key": int(md.kind_key),
            "kind_compressed_slot": int(md.kind_compressed_slot),
            "count_present": int(md.count_present),
            "count_deleted": int(md.count_deleted),
            "dimensions": int(md.dimensions),
        }

    @staticmethod
    def restore(path_or_buffer, view: bool = False) -> Optional["Index"]:
        meta = Index.metadata(path_or_buffer)
        if not meta:
            return None
        index = Index(ndim=meta["dimensions"], dtype=meta["kind_scalar"], metric=meta["kind_metric"])
        if view:
            index.view(path_or_buffer)
        else:
            index.load(path_or_buffer)
        return index

    @property
    def specs(self) -> dict:
        return {
            "type": "rad_b200.Index",
            "ndim": self.ndim,
            "multi": self.multi,
            "connectivity": self.connectivity,
            "expansion_add": self.expansion_add,
            "expansion_search": self.expansion_search,
            "size": self.size,
            "jit": self.jit,
            "hardware_acceleration": self.hardware_acceleration,
            "metric_kind": self.metric_kind,
            "dtype": self.dtype,
            "path": self.path,
        }

    def __repr__(self) -> str:
        if not getattr(self, "_h", None):
            return "rad_b200.Index(reset)"
        return (f"rad_b200.Index({self.dtype} x {self.ndim}, {self.metric_kind}, connectivity: {self.connectivity}, "
                f"expansion: {self.expansion_add} & {self.expansion_search}, {len(self):,} vectors in "
                f"{self.nlevels} levels, {self.hardware_acceleration})")

