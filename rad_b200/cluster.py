"""``Index.cluster``: map vectors (or stored keys) to centroids taken from an upper level of the HNSW graph
(SURVEY.md section 8f row 4).

Reference: ``usearch/python/usearch/index.py:1214-1273`` -> ``usearch/python/lib.cpp:758-875`` ->
``usearch/include/usearch/index_dense.hpp:1919-2080``: pick the level whose node count exceeds ``min_count``; send
every query down the graph to that level (``cluster(query, level)``, ``index.hpp:3046-3079``: the greedy descent
alone); go one level deeper while fewer than ``min_count`` distinct centroids were hit; while there are more than
``max_count``, merge the least popular centroid into its nearest other centroid; queries of merged centroids are
re-measured against the centroid that absorbed them.

The descents run on the GPU (``rb_cluster_device``, one warp per query, deterministic given the graph -- on a graph the
reference built the answers are the reference's); the bookkeeping over the (few) distinct centroids is numpy.  Where
the reference leaves an order open (``std::sort`` by popularity is not stable) ties go to the lower centroid key."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native


def descend(index, queries: np.ndarray, level: int, batch: int = 1 << 20):
    """(slots uint32 [Q], distances float32 [Q]): where the greedy descent to ``level`` ends for every query."""
    import torch
    lib = _native.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    nq = len(queries)
    slots = np.empty(nq, dtype=np.uint32)
    dists = np.empty(nq, dtype=np.float32)
    visited = computed = 0
    for first in range(0, nq, batch):
        part = queries[first:first + batch]
        q = torch.zeros((len(part), 128), dtype=torch.uint8, device=dev)
        q[:, :part.shape[1]] = torch.from_numpy(np.ascontiguousarray(part)).to(dev)
        s = torch.empty((len(part),), dtype=torch.int32, device=dev)
        d = torch.empty((len(part),), dtype=torch.float32, device=dev)
        torch.cuda.current_stream(dev).synchronize()
        err = C.c_char_p()
        lib.rb_cluster_device(index._handle(), q.data_ptr(), len(part), level, s.data_ptr(), d.data_ptr(), C.byref(err))
        _native.check(err)
        slots[first:first + len(part)] = s.cpu().numpy().view(np.uint32)
        dists[first:first + len(part)] = d.cpu().numpy()
        counters = index.last_counters
        visited += int(counters.get("visited_members", 0))
        computed += int(counters.get("computed_distances", 0))
    return slots, dists, visited, computed


def _tanimoto(a: np.ndarray, b: np.ndarray) -> np.float32:
    """one pair, the f64 formula narrowed to f32 (simsimd binary.h:99-105, index_plugins.hpp:1892)"""
    inter = int(np.unpackbits(a & b).sum())
    union = int(np.unpackbits(a | b).sum())
    return np.float32(1.0 - inter / union) if union else np.float32(0.0)


def cluster(index, queries: np.ndarray, min_count: int, max_count: int):
    """-> (centroid keys uint64 [Q], distances float32 [Q], visited_members, computed_distances, level used)"""
    if index.max_level < 2:
        raise RuntimeError("Index too small to cluster!")                     # index_dense.hpp:1941-1942
    level = index.max_level
    if min_count:
        while level > 1 and not index.level_stats(level).nodes > min_count:   # index_dense.hpp:1931-1937
            level -= 1
    else:
        level, max_count, min_count = 1, index.level_stats(1).nodes, 2        # index_dense.hpp:1938-1939
    keys_of = np.asarray(index.keys)
    visited = computed = 0
    while True:
        slots, dists, v, c = descend(index, queries, level)
        visited, computed = visited + v, computed + c
        centroid_keys = keys_of[slots]
        uniq, first, popularity = np.unique(centroid_keys, return_index=True, return_counts=True)
        if len(uniq) < min_count and level > 1:                               # index_dense.hpp:2004-2009
            level -= 1
            continue
        break
    if len(uniq) <= max_count:
        return centroid_keys, dists, visited, computed, level
    # ---- merge the least popular centroid into its nearest one until max_count are left (index_dense.hpp:2011-2042)
    order = np.lexsort((uniq, -popularity))           # popularity descending; ties: lower key first (see module doc)
    c_key = uniq[order].tolist()
    c_pop = popularity[order].tolist()
    c_vec = [index.get(int(k)) for k in c_key]
    merged_into = {}
    live = len(c_key)
    while live > max_count:
        src = live - 1
        best, best_d = 0, np.float32(np.finfo(np.float32).max)
        for cand in range(src):
            d = _tanimoto(c_vec[src], c_vec[cand])
            if d < best_d:
                best, best_d = cand, d
        merged_into[c_key[src]] = c_key[best]
        c_pop[best] += c_pop[src]
        c_pop[src] = 0
        while best and c_pop[best - 1] < c_pop[best]:
            for arr in (c_key, c_pop, c_vec):
                arr[best - 1], arr[best] = arr[best], arr[best - 1]
            best -= 1
        live -= 1
    # ---- queries of merged centroids follow the chain and are measured again (index_dense.hpp:2044-2069)
    final = centroid_keys.copy()
    out_d = dists.copy()
    resolved = {}
    for k in merged_into:
        t = k
        while t in merged_into:
            t = merged_into[t]
        resolved[k] = t
    vec_of = {}
    for qi in np.flatnonzero(np.isin(centroid_keys, np.fromiter(resolved.keys(), dtype=centroid_keys.dtype))):
        t = resolved[int(centroid_keys[qi])]
        final[qi] = t
        if t not in vec_of:
            vec_of[t] = index.get(int(t))
        out_d[qi] = _tanimoto(vec_of[t], np.ascontiguousarray(queries[qi]))
    return final, out_d, visited, computed, level
