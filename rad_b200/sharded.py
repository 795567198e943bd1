"""Row-sharded Tanimoto k-NN across the GPUs of one box: one process per GPU, NCCL all-gather of the per-shard
``[Q, k]`` top-k keys, device merge.

This is the reference's multi-index model -- ``Indexes`` searching every shard and merging per query
(usearch/python/lib.cpp:417-498, merge rule usearch/include/usearch/index.hpp:2625-2645) -- with shards = contiguous
row ranges, one per rank.  The exchange is tiny (Q x k x 8 bytes per rank), so it is a plain collective after the
shard-local kernel; the shard-local work has no data-path communication.

Key format (the same 64-bit word the kernels emit): ``float32 distance bits << 32 | ~global_slot``.  Distances are
non-negative floats, so ascending unsigned key order IS the reference's result order: ascending distance, ties by
DESCENDING slot, and the k smallest keys are exactly the reference's top-k (index.hpp:923-934).
"""
from __future__ import annotations

import ctypes as C
from typing import Tuple

import numpy as np

EMPTY_KEY = np.uint64(0xFFFFFFFFFFFFFFFF)


def shard_bounds(n_total: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous rows ``[lo, hi)`` of rank ``rank``: the first ``n_total % world`` ranks hold one extra row."""
    base, extra = divmod(int(n_total), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def pack_topk_keys(distances: np.ndarray, slots: np.ndarray, counts: np.ndarray | None = None) -> np.ndarray:
    """(float32 distances, global slots) -> uint64 keys; entries past ``counts[q]`` become EMPTY_KEY."""
    bits = np.ascontiguousarray(distances, dtype=np.float32).view(np.uint32).astype(np.uint64)
    inv = (~np.asarray(slots).astype(np.uint64)) & np.uint64(0xFFFFFFFF)
    keys = (bits << np.uint64(32)) | inv
    if counts is not None:
        k = keys.shape[1]
        keys[np.arange(k)[None, :] >= np.asarray(counts)[:, None]] = EMPTY_KEY
    return keys


def unpack_topk_keys(keys: np.ndarray) -> Tuple[np.ndarray, np.ndarray, np.ndarray]:
    """uint64 keys -> (global slots uint64, float32 distances, counts int64)."""
    keys = np.asarray(keys).view(np.uint64)
    valid = keys != EMPTY_KEY
    slots = ((~keys) & np.uint64(0xFFFFFFFF)).astype(np.uint64)
    dists = (keys >> np.uint64(32)).astype(np.uint32).view(np.float32)
    return np.where(valid, slots, 0), np.where(valid, dists, np.float32(0)), valid.sum(axis=1).astype(np.int64)


def merge_topk_host(parts: np.ndarray, k: int) -> np.ndarray:
    """Host statement of the k-way merge: ``parts`` [P, Q, k] -> [Q, k] smallest keys per query (tests, tiny inputs;
    the product path merges on the device with ``rb_merge_topk_device``)."""
    parts = np.asarray(parts).view(np.uint64)
    p, q, kk = parts.shape
    flat = np.transpose(parts, (1, 0, 2)).reshape(q, p * kk)
    return np.sort(flat, axis=1)[:, :k]


class ShardedExactSearch:
    """Exact Tanimoto top-k over a database whose rows are sharded across the ranks of ``torch.distributed``.

    ``d_shard`` is this rank's ``uint8 [rows, 128]`` CUDA tensor, ``slot_base`` the global slot of its first row.
    ``search(d_queries, k)`` returns the merged ``int64 [Q, k]`` key tensor, identical on every rank."""

    def __init__(self, d_shard, slot_base: int):
        import torch
        from . import _native
        self._torch = torch
        self._native = _native
        self._lib = _native.lib()
        self.d_shard = d_shard
        self.rows = int(d_shard.shape[0])
        self.slot_base = int(slot_base)
        dist = torch.distributed
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self._ws = None

    def search(self, d_queries, k: int = 10):
        return self.merge(self.search_local(d_queries, k), k)

    def search_local(self, d_queries, k: int = 10):
        """This rank's shard only: ``int64 [Q, k]`` keys carrying GLOBAL slots, before the exchange."""
        torch = self._torch
        nq = int(d_queries.shape[0])
        need = int(self._lib.rb_exact_topk_workspace_bytes(self.rows, nq, k))
        if self._ws is None or self._ws.numel() < need:
            self._ws = torch.empty(need, dtype=torch.uint8, device=d_queries.device)
        part = torch.empty((nq, k), dtype=torch.int64, device=d_queries.device)
        err = C.c_char_p()
        stream = torch.cuda.current_stream().cuda_stream
        self._lib.rb_exact_topk_device(self.d_shard.data_ptr(), self.rows, self.slot_base, d_queries.data_ptr(), nq,
                                       k, part.data_ptr(), self._ws.data_ptr(), need, stream, C.byref(err))
        self._native.check(err)
        return part

    def merge(self, part, k: int):
        """All-gather the per-rank ``[Q, k]`` keys and merge them on the device (no-op for a single rank)."""
        if self.world == 1:
            return part
        torch = self._torch
        nq = int(part.shape[0])
        gathered = torch.empty((self.world, nq, k), dtype=torch.int64, device=part.device)
        torch.distributed.all_gather_into_tensor(gathered.view(-1), part.contiguous().view(-1))
        merged = torch.empty((nq, k), dtype=torch.int64, device=part.device)
        err = C.c_char_p()
        self._lib.rb_merge_topk_device(gathered.data_ptr(), self.world, nq, k, merged.data_ptr(),
                                       torch.cuda.current_stream().cuda_stream, C.byref(err))
        self._native.check(err)
        return merged


class ShardedIndex:
    """The reference's multi-index search (``Indexes``: usearch/python/usearch/index.py:1389-1482, lib.cpp:417-498,
    merge rule index.hpp:2625-2645) for one process per GPU: every rank holds a sub-index over a contiguous shard of
    the rows; ``search`` runs on every shard (exact or ef-bounded, on that rank's GPU), all-gathers the per-shard
    ``[Q, k]`` answers and merges them, so every rank returns the same ``BatchMatches``.

    ``local_index`` is a ``rad_b200.index.Index`` (or anything with its ``add`` / ``search`` signature) whose slots are
    this shard's rows in order; ``slot_base`` is the global slot of its first row.  Results carry the user's keys;
    ordering is the reference's: ascending distance, ties by descending GLOBAL slot.  The exchange is
    ``2 x Q x k x 8`` bytes per rank through ``torch.distributed`` (NCCL on GPUs, gloo in the CPU tests)."""

    def __init__(self, local_index, slot_base: int):
        import torch
        self._torch = torch
        self.local = local_index
        self.slot_base = int(slot_base)
        dist = torch.distributed
        self._dist = dist if dist.is_available() and dist.is_initialized() else None
        self.world = self._dist.get_world_size() if self._dist else 1

    def __len__(self) -> int:
        n = len(self.local)
        if self._dist is None:
            return n
        t = self._tensor(np.array([n], dtype=np.int64))
        self._dist.all_reduce(t)
        return int(t.cpu().item())

    def add(self, keys, vectors, **kw):
        """Insert this rank's rows (``keys`` are the user's keys of those rows) into the local sub-index."""
        return self.local.add(keys, vectors, **kw)

    def _tensor(self, a: np.ndarray):
        t = self._torch.from_numpy(np.ascontiguousarray(a))
        if self._dist is not None and self._dist.get_backend() == "nccl":
            t = t.cuda()
        return t

    def search(self, vectors, count: int = 10, *, exact: bool = False):
        from .index import BatchMatches
        vectors = np.atleast_2d(vectors)
        got = self.local.search(vectors, count, exact=exact)
        nq = len(vectors)
        keys = np.asarray(got.keys, dtype=np.uint64).reshape(nq, count)
        dists = np.asarray(got.distances, dtype=np.float32).reshape(nq, count)
        counts = np.asarray(got.counts, dtype=np.int64).reshape(nq)
        # merge key = distance bits << 32 | ~global slot; the local slot of a result is recovered from its key
        slots = np.asarray(self.local.get_node_ids_from_keys(keys.reshape(-1)), dtype=np.uint64).reshape(nq, count)
        mkeys = pack_topk_keys(dists, slots + np.uint64(self.slot_base), counts)
        if self._dist is None:
            return got
        both = np.stack([mkeys.view(np.int64), keys.view(np.int64)])            # [2, Q, k]
        part = self._tensor(both)
        gathered = self._torch.empty((self.world,) + tuple(part.shape), dtype=part.dtype, device=part.device)
        self._dist.all_gather_into_tensor(gathered.view(-1), part.contiguous().view(-1))
        g = gathered.cpu().numpy()                                               # [world, 2, Q, k]
        flat_m = np.transpose(g[:, 0], (1, 0, 2)).reshape(nq, self.world * count).view(np.uint64)
        flat_k = np.transpose(g[:, 1], (1, 0, 2)).reshape(nq, self.world * count).view(np.uint64)
        order = np.argsort(flat_m, axis=1, kind="stable")[:, :count]
        merged = np.take_along_axis(flat_m, order, axis=1)
        out_keys = np.take_along_axis(flat_k, order, axis=1)
        _, out_d, out_c = unpack_topk_keys(merged)
        out_keys = np.where(merged != EMPTY_KEY, out_keys, 0)
        return BatchMatches(keys=out_keys, distances=out_d, counts=out_c,
                            visited_members=int(getattr(got, "visited_members", 0)),
                            computed_distances=int(getattr(got, "computed_distances", 0)))
