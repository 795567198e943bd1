"""``Index.join``: one-to-one matching between the rows of two indexes.

Reference: ``usearch/python/usearch/index.py:1182-1212`` -> ``usearch/python/lib.cpp:876-894`` ->
``usearch/include/usearch/index.hpp:4330-4515`` (male-optimal stable marriage over nearest-neighbour searches: a free
"man" proposes to his k-th nearest "woman", k growing with every refusal; a woman leaves her husband only for a strictly
nearer man; the smaller index plays the men).

Here every man's ``max_proposals`` candidates come from ONE batched search on the GPU (``rb_search_device`` on the
women's index: the exact tensor-core scan or the ef-bounded graph search, k = max_proposals -- the k-th answer of a
search for k rows is the k-th answer of the search for max_proposals rows), and the proposals run as data-parallel
generations (``rb_stable_match_device``) that reproduce the reference's loop on one thread.  PyTorch holds the device
buffers; the work is the C ABI's."""
from __future__ import annotations

import ctypes as C
import math
import os

import numpy as np

from . import _native

NONE = 0xFFFFFFFF


def default_proposals(men: int, threads: int | None = None) -> int:
    """``log(men) + threads`` truncated (index.hpp:4361-4362; the reference adds the size of its thread pool)."""
    threads = (os.cpu_count() or 1) if threads is None else threads
    return int(math.log(men) + threads) if men else 0


def candidates_device(men_index, women_index, proposals: int, exact: bool, expansion: int, batch: int = 1 << 20):
    """(slots int32 [men, proposals], distances float32, counts int32) on the GPU: every man's nearest women."""
    import torch
    lib = _native.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    n = len(men_index)
    slots = torch.empty((n, proposals), dtype=torch.int32, device=dev)
    dists = torch.empty((n, proposals), dtype=torch.float32, device=dev)
    counts = torch.zeros((n,), dtype=torch.int32, device=dev)
    saved = women_index.expansion_search
    women_index.expansion_search = max(int(expansion), proposals)
    try:
        for first in range(0, n, batch):
            cnt = min(batch, n - first)
            q = torch.empty((cnt, 128), dtype=torch.uint8, device=dev)
            torch.cuda.current_stream(dev).synchronize()
            err = C.c_char_p()
            lib.rb_copy_vectors_device(men_index._handle(), first, cnt, q.data_ptr(), C.byref(err))
            _native.check(err)
            lib.rb_search_device(women_index._handle(), q.data_ptr(), cnt, proposals, 1 if exact else 0,
                                 slots[first:first + cnt].data_ptr(), dists[first:first + cnt].data_ptr(),
                                 counts[first:first + cnt].data_ptr(), C.byref(err))
            _native.check(err)
    finally:
        women_index.expansion_search = saved
    return slots, dists, counts


def stable_match_device(slots, dists, counts, women: int):
    """``rb_stable_match_device`` on torch CUDA tensors -> (man_to_woman int32 [men], woman_to_man int32 [women],
    {"generations", "engagements"}); -1 = unmatched."""
    import torch
    lib = _native.lib()
    men, proposals = int(slots.shape[0]), int(slots.shape[1])
    dev = slots.device
    m2w = torch.empty((men,), dtype=torch.int32, device=dev)
    w2m = torch.empty((women,), dtype=torch.int32, device=dev)
    need = int(lib.rb_stable_match_workspace_bytes(men, women))
    ws = torch.empty((max(need, 1),), dtype=torch.uint8, device=dev)
    generations, engagements = C.c_uint64(0), C.c_uint64(0)
    err = C.c_char_p()
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream()
        lib.rb_stable_match_device(slots.data_ptr(), dists.data_ptr(), counts.data_ptr(), men, women, proposals,
                                   m2w.data_ptr(), w2m.data_ptr(), ws.data_ptr(), need, stream.cuda_stream,
                                   C.byref(generations), C.byref(engagements), C.byref(err))
        _native.check(err)
        stream.synchronize()
    return m2w, w2m, {"generations": int(generations.value), "engagements": int(engagements.value)}


def join(a, b, max_proposals: int = 0, exact: bool = False, progress=None, stats: dict | None = None) -> dict:
    """Mapping from keys of ``a`` to keys of ``b`` with no key used twice (the reference's ``a.join(b, ...)``)."""
    if a is b:
        raise RuntimeError("Can't join with itself, consider copying")  # index.hpp:4356-4357
    men, women = (a, b) if len(a) <= len(b) else (b, a)                 # index.hpp:4343-4353
    if not len(men):
        return {}
    proposals = min(len(men), max_proposals or default_proposals(len(men)))
    expansion = max(a.expansion_search, b.expansion_search)             # lib.cpp:888
    slots, dists, counts = candidates_device(men, women, proposals, exact, expansion)
    m2w, _, info = stable_match_device(slots, dists, counts, len(women))
    m2w = m2w.cpu().numpy()
    matched = np.flatnonzero(m2w >= 0)
    men_keys = np.asarray(men.keys)[matched]
    women_keys = np.asarray(women.keys)[m2w[matched]]
    if stats is not None:
        stats.update(info, proposals=proposals, pairs=len(matched))
    if progress:
        progress(len(men), len(men))
    left, right = (men_keys, women_keys) if men is a else (women_keys, men_keys)
    return dict(zip(left.tolist(), right.tolist()))
