"""Fingerprint ingest pre-passes on the GPU: popcounts and exact duplicates.

The reference inserts identical fingerprints like any other row (usearch/include/usearch/index.hpp:2734-2834); the
d = 0 cliques they form fill the M links of their members and inflate every ef-bounded search that meets them.
``duplicate_representatives`` names the first row of every set of identical fingerprints (``rb_dedup_device``), and
``add_deduplicated`` builds the HNSW graph over the distinct fingerprints only and hangs every duplicate off its
representative as a CHAIN (representative -> first duplicate -> second duplicate ...): every row keeps its slot, key
and vector, every row stays reachable, and no row spends more than two links on copies of itself.

PyTorch is used for device memory only; the work is the C ABI's (``include/rad_b200.h``)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native

EMPTY = np.uint32(0xFFFFFFFF)


def _device_rows(vectors: np.ndarray, device):
    import torch
    vectors = np.ascontiguousarray(vectors, dtype=np.uint8)
    if vectors.ndim != 2 or vectors.shape[1] > 128:
        raise ValueError("expects a matrix of packed fingerprints, at most 128 bytes per row")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    d = torch.zeros((len(vectors), 128), dtype=torch.uint8, device=dev)  # rows are zero-padded to one 128-byte line
    d[:, :vectors.shape[1]] = torch.from_numpy(vectors).to(dev)
    return torch, dev, d


def duplicate_representatives_device(d_vectors, out=None):
    """``d_vectors``: uint8 [n, 128] CUDA tensor -> int32 CUDA tensor [n]: the lowest row with the same bytes."""
    import torch
    lib = _native.lib()
    n = int(d_vectors.shape[0])
    rep = torch.empty((n,), dtype=torch.int32, device=d_vectors.device) if out is None else out
    if not n:
        return rep
    need = int(lib.rb_dedup_workspace_bytes(n))
    ws = torch.empty((need,), dtype=torch.uint8, device=d_vectors.device)
    err = C.c_char_p()
    with torch.cuda.device(d_vectors.device):
        lib.rb_dedup_device(d_vectors.data_ptr(), n, rep.data_ptr(), ws.data_ptr(), need,
                            torch.cuda.current_stream().cuda_stream, C.byref(err))
        _native.check(err)
        torch.cuda.current_stream().synchronize()
    return rep


def duplicate_representatives(vectors: np.ndarray, device=None) -> np.ndarray:
    """uint32 [n]: for every row the LOWEST row holding the same fingerprint (== its own index when it is the first)."""
    _, _, d = _device_rows(vectors, device)
    return duplicate_representatives_device(d).cpu().numpy().view(np.uint32)


def popcounts(vectors: np.ndarray, device=None) -> np.ndarray:
    """uint16 [n]: set bits per fingerprint (``rb_popcounts_device``)."""
    torch, dev, d = _device_rows(vectors, device)
    out = torch.empty((len(d),), dtype=torch.int16, device=dev)
    err = C.c_char_p()
    with torch.cuda.device(dev):
        _native.lib().rb_popcounts_device(d.data_ptr(), len(d), out.data_ptr(), torch.cuda.current_stream().cuda_stream,
                                          C.byref(err))
        _native.check(err)
        torch.cuda.current_stream().synchronize()
    return out.cpu().numpy().view(np.uint16)


def chain_duplicates(representative: np.ndarray, adj0: np.ndarray) -> None:
    """Hang the duplicates off their representatives, in place: ``adj0`` [n, 2M] holds the level-0 rows of the distinct
    rows (global slots, EMPTY padded, duplicates' rows EMPTY).  Row of a duplicate = [representative, next duplicate];
    the representative gets its first duplicate in a free place, or in its LAST place when the row is full."""
    rep = np.asarray(representative).astype(np.int64)
    n = len(rep)
    dups = np.flatnonzero(rep != np.arange(n))
    if not len(dups):
        return
    order = dups[np.argsort(rep[dups], kind="stable")]          # grouped by representative, ascending slot inside
    same_next = np.append(rep[order[1:]] == rep[order[:-1]], False)
    nxt = np.where(same_next, np.append(order[1:], 0), int(EMPTY)).astype(np.uint32)
    adj0[order, 0] = rep[order].astype(np.uint32)
    adj0[order, 1] = nxt
    first = np.append(True, rep[order[1:]] != rep[order[:-1]])
    heads, reps = order[first], rep[order[first]]
    deg = (adj0[reps] != EMPTY).sum(axis=1)
    at = np.minimum(deg, adj0.shape[1] - 1)
    adj0[reps, at] = heads.astype(np.uint32)


def add_deduplicated(index, keys, vectors: np.ndarray) -> np.ndarray:
    """``index.add(keys, vectors)`` for an EMPTY ``rad_b200.index.Index`` with the duplicates taken out of the graph
    search: the graph is built over the distinct fingerprints, duplicates are chained behind their representatives
    (``chain_duplicates``), and the index ends up with every row in its slot (keys, vectors, exact search, file format
    unchanged).  Returns the representative of every row (uint32)."""
    from .index import Index
    if len(index):
        raise ValueError("add_deduplicated fills an empty index")
    vectors = np.ascontiguousarray(vectors, dtype=np.uint8)
    n = len(vectors)
    keys = np.arange(n, dtype=np.uint64) if keys is None else np.ascontiguousarray(keys, dtype=np.uint64)
    assert len(keys) == n
    rep = duplicate_representatives(vectors)
    uniq = np.flatnonzero(rep == np.arange(n, dtype=np.uint32))
    if len(uniq) == n:
        index.add(keys, vectors)
        return rep
    sub = Index(ndim=index.ndim, connectivity=index.connectivity, expansion_add=index.expansion_add,
                expansion_search=index.expansion_search)
    sub.add(None, vectors[uniq])
    a0, au = sub._debug_adjacency()
    table = np.append(uniq.astype(np.uint32), EMPTY)           # compact slot -> row, EMPTY -> EMPTY

    def to_rows(a):
        return table[np.where(a == EMPTY, len(uniq), a).astype(np.int64)]

    levels = np.zeros((n,), dtype=np.int16)
    levels[uniq] = np.asarray(sub.levels, dtype=np.int16)
    adj0 = np.full((n, a0.shape[1]), EMPTY, dtype=np.uint32)
    adj0[uniq] = to_rows(a0)
    chain_duplicates(rep, adj0)
    # only distinct rows have upper levels, in the same (ascending) order as in the compact build
    index.import_graph(keys, levels, adj0, to_rows(au), vectors, int(uniq[sub.entry_slot]))
    return rep
