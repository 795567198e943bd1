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
        if len(local_index) == 0 and hasattr(local_index, "seek_level_stream"):
            # the shard draws ITS rows of the reference's single level stream: the global graph gets the levels of a
            # one-piece build (index.hpp:3867-3871)
            local_index.seek_level_stream(self.slot_base)

    def build_global(self, *, with_vectors: bool = True, expansion_add: int | None = None, timings: dict | None = None):
        """ONE graph over all shards (``build_global_graph``: cross-shard stitch + upper levels on rank 0).  Collective.
        Returns on rank 0 a ``rad_b200.index.Index`` holding the global graph in global slots with the users' keys --
        ``save(path)`` writes the single `.usearch` file RAD serves (scripts/start_hnsw_server.py:69) -- and ``None``
        on the other ranks.  ``with_vectors=False`` leaves the fingerprints on their GPUs (a file for
        ``Index(path, view=True, exclude_vectors=True)``)."""
        import torch
        from .index import Index
        shard = DeviceShard(self.local)
        comm = TorchComm()
        ef_add = int(expansion_add or self.local.expansion_add)
        res = build_global_graph(shard, self.slot_base, expansion_add=ef_add, timings=timings, comm=comm)
        n_local = len(self.local)
        sizes = comm.all_gather(torch.tensor([n_local], dtype=torch.int64, device=shard.device)).view(-1).tolist()
        pad = max(sizes)
        keys = torch.zeros((pad,), dtype=torch.int64, device=shard.device)
        keys[:n_local] = torch.from_numpy(np.asarray(self.local.keys, dtype=np.uint64).view(np.int64)).to(shard.device)
        g_keys = comm.gather0(keys)
        g_fp = None
        if with_vectors:
            fp = torch.zeros((pad, 128), dtype=torch.uint8, device=shard.device)
            if n_local:
                shard.vectors(0, n_local, fp)
            g_fp = comm.gather0(fp)
        if comm.rank != 0:
            return None
        all_keys = np.concatenate([g_keys[h][:sizes[h]].cpu().numpy() for h in range(comm.world)]).view(np.uint64)
        vectors = None
        if with_vectors:
            vb = self.local.ndim // 8
            vectors = np.concatenate([g_fp[h][:sizes[h], :vb].cpu().numpy() for h in range(comm.world)])
        whole = Index(ndim=self.local.ndim, connectivity=self.local.connectivity, expansion_add=ef_add,
                      expansion_search=self.local.expansion_search,
                      device=shard.device.index if shard.device.index is not None else -1, exclude_vectors=not with_vectors)
        whole.import_graph(all_keys, res["levels"], res["adj0"], res["adj_upper"], vectors, res["entry"])
        return whole

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
        """Merged k-NN over all shards, the same ``BatchMatches`` on every rank.  Under NCCL the answers never leave
        the GPUs until they are merged (``search_device``); under gloo (CPU tests) the merge is the numpy statement."""
        vectors = np.atleast_2d(vectors)
        if self._dist is not None and self._dist.get_backend() == "nccl" and hasattr(self.local, "_handle"):
            return self._search_on_device(vectors, count, exact)
        return self._search_on_host(vectors, count, exact)

    # ---- device-resident path: rb_search_device -> keys on the GPU -> all-gather -> rb_merge_topk_device -> user keys
    def _device_keys(self, device):
        """this shard's user keys as an int64 CUDA tensor (uploaded once per index size)"""
        n = len(self.local)
        cached = getattr(self, "_d_user_keys", None)
        if cached is None or cached.numel() != n or cached.device != device:
            keys = np.ascontiguousarray(np.asarray(self.local.keys, dtype=np.uint64)).view(np.int64)
            cached = self._torch.from_numpy(keys).to(device)
            self._d_user_keys = cached
        return cached

    def local_answers_device(self, d_queries, count: int, exact: bool):
        """This shard's answers for CUDA queries ``uint8 [Q, 128]`` -> (merge keys int64 [Q, k] = distance bits << 32 |
        ~GLOBAL slot, -1 padded; the users' keys of those answers int64 [Q, k])."""
        torch = self._torch
        from . import _native
        dev = d_queries.device
        nq = int(d_queries.shape[0])
        slots = torch.zeros((nq, count), dtype=torch.int32, device=dev)
        dists = torch.zeros((nq, count), dtype=torch.float32, device=dev)
        counts = torch.zeros((nq,), dtype=torch.int32, device=dev)
        if nq and len(self.local):
            torch.cuda.current_stream(dev).synchronize()  # the library works on the handle's own stream
            err = C.c_char_p()
            _native.lib().rb_search_device(self.local._handle(), d_queries.data_ptr(), nq, count, 1 if exact else 0,
                                           slots.data_ptr(), dists.data_ptr(), counts.data_ptr(), C.byref(err))
            _native.check(err)
        valid = torch.arange(count, device=dev)[None, :] < counts[:, None]
        local = torch.where(valid, slots, torch.zeros_like(slots)).long()
        inv_global = (~(local + self.slot_base)) & 0xFFFFFFFF
        mkeys = (dists.view(torch.int32).long() << 32) | inv_global
        mkeys = torch.where(valid, mkeys, torch.full_like(mkeys, -1))
        ukeys = self._device_keys(dev)[local] if len(self.local) else torch.zeros_like(mkeys)
        return mkeys, torch.where(valid, ukeys, torch.zeros_like(ukeys))

    @staticmethod
    def merge_answers_device(all_mkeys, all_ukeys, count: int):
        """``[W, Q, k]`` merge keys and user keys of all shards -> (user keys [Q, k], distances f32 [Q, k], counts
        int64 [Q]) after ``rb_merge_topk_device`` (the merge rule of index.hpp:2625-2645), all on the device."""
        import torch
        from . import _native
        world, nq, k = all_mkeys.shape
        merged = torch.empty((nq, count), dtype=torch.int64, device=all_mkeys.device)
        err = C.c_char_p()
        _native.lib().rb_merge_topk_device(all_mkeys.contiguous().data_ptr(), world, nq, k, merged.data_ptr(),
                                           torch.cuda.current_stream(all_mkeys.device).cuda_stream, C.byref(err))
        _native.check(err)
        # where each merged key came from among the world * k gathered ones (global slots are unique)
        flat_m = all_mkeys.permute(1, 0, 2).reshape(nq, world * k)
        flat_u = all_ukeys.permute(1, 0, 2).reshape(nq, world * k)
        src = (flat_m[:, None, :] == merged[:, :, None]).int().argmax(dim=2)
        valid = merged != -1
        out_keys = torch.where(valid, torch.gather(flat_u, 1, src), torch.zeros_like(merged))
        out_d = torch.where(valid, (merged >> 32).int().view(torch.float32), torch.zeros((), device=merged.device))
        return out_keys, out_d, valid.sum(dim=1)

    def search_device(self, d_queries, count: int = 10, *, exact: bool = False):
        """CUDA queries in, CUDA answers out: (user keys int64 [Q, k], distances f32 [Q, k], counts int64 [Q])."""
        torch = self._torch
        mkeys, ukeys = self.local_answers_device(d_queries, count, exact)
        both = torch.stack([mkeys, ukeys])                                      # [2, Q, k]
        gathered = torch.empty((self.world,) + tuple(both.shape), dtype=both.dtype, device=both.device)
        if self.world > 1:
            self._dist.all_gather_into_tensor(gathered.view(-1), both.view(-1))
        else:
            gathered[0] = both
        return self.merge_answers_device(gathered[:, 0].contiguous(), gathered[:, 1].contiguous(), count)

    def _search_on_device(self, vectors, count, exact):
        from .index import BatchMatches
        torch = self._torch
        dev = torch.device("cuda", torch.cuda.current_device())
        q = np.ascontiguousarray(vectors, dtype=np.uint8)
        d_q = torch.zeros((len(q), 128), dtype=torch.uint8, device=dev)
        d_q[:, :q.shape[1]] = torch.from_numpy(q).to(dev)
        keys, dists, counts = self.search_device(d_q, count, exact=exact)
        c = self.local.last_counters if hasattr(self.local, "last_counters") else {}
        return BatchMatches(keys=keys.cpu().numpy().view(np.uint64), distances=dists.cpu().numpy(),
                            counts=counts.cpu().numpy(), visited_members=int(c.get("visited_members", 0)),
                            computed_distances=int(c.get("computed_distances", 0)))

    def _search_on_host(self, vectors, count, exact):
        from .index import BatchMatches
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


# ================================================================================================ one graph from G shards
EMPTY_SLOT = -1  # 0xFFFFFFFF in the int32 tensors that carry u32 slots between torch and the C ABI


def stitch_parameters(world: int) -> Tuple[int, int]:
    """(remote candidates per node and shard, ef of the remote searches): with few shards every other shard matters
    (half of a node's true neighbours live there), with many the union of their answers does (measured with
    scripts/stitch_probe.py: recall of the stitched graph within 0.5 points of the one-piece graph at 2..8 shards;
    at 8 shards ef 32 gives 0.991 / 0.995 on the routable set where the one-piece graph has 0.979 / 0.996)."""
    if world <= 3:
        return 16, 128
    return (8, 64) if world <= 5 else (8, 32)


class DeviceShard:
    """The C-ABI calls the stitch needs from one rank's sub-index, on torch CUDA tensors (``rb_copy_vectors_device``,
    ``rb_search_device``, ``rb_gather_vectors_device``, ``rb_stitch_device``)."""

    def __init__(self, index, device=None):
        import torch
        from . import _native
        self._torch, self._native, self.index = torch, _native, index
        self._lib = _native.lib()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.m, self.m0 = index.connectivity, index.connectivity_base

    def __len__(self):
        return len(self.index)

    def _settle(self):
        """The library works on the handle's own (non-blocking) stream and returns when its work is done; what torch
        queued on ITS stream -- the tensors handed in -- must be complete before the library reads them."""
        self._torch.cuda.current_stream(self.device).synchronize()

    def levels(self) -> np.ndarray:
        return self.index.levels

    def vectors(self, first: int, count: int, out):
        self._settle()
        err = C.c_char_p()
        self._lib.rb_copy_vectors_device(self.index._handle(), first, count, out.data_ptr(), C.byref(err))
        self._native.check(err)

    def gather_vectors(self, slots):
        torch = self._torch
        out = torch.empty((slots.numel(), 128), dtype=torch.uint8, device=self.device)
        self._settle()
        err = C.c_char_p()
        self._lib.rb_gather_vectors_device(self.index._handle(), slots.data_ptr(), slots.numel(), out.data_ptr(),
                                           C.byref(err))
        self._native.check(err)
        return out

    def remote_candidates(self, queries, k: int, ef: int):
        """The k nearest rows of THIS shard for foreign queries -> (local slots int32 [n, k], EMPTY_SLOT = none;
        their fingerprints uint8 [n, k, 128])."""
        torch = self._torch
        n = int(queries.shape[0])
        slots = torch.full((n, k), EMPTY_SLOT, dtype=torch.int32, device=self.device)
        dists = torch.empty((n, k), dtype=torch.float32, device=self.device)
        counts = torch.zeros((n,), dtype=torch.int32, device=self.device)
        if n and len(self.index):
            self.index.expansion_search = ef
            self._settle()
            err = C.c_char_p()
            self._lib.rb_search_device(self.index._handle(), queries.data_ptr(), n, k, 0, slots.data_ptr(),
                                       dists.data_ptr(), counts.data_ptr(), C.byref(err))
            self._native.check(err)
            slots = torch.where(torch.arange(k, device=self.device)[None, :] < counts[:, None], slots,
                                torch.full_like(slots, EMPTY_SLOT))
        return slots, self.gather_vectors(slots.reshape(-1).contiguous()).reshape(n, k, 128)

    def stitch(self, first: int, count: int, slot_base: int, cand_slots, cand_fp, out_rows):
        self._settle()
        err = C.c_char_p()
        self._lib.rb_stitch_device(self.index._handle(), first, count, slot_base, cand_slots.data_ptr(),
                                   cand_fp.data_ptr(), int(cand_slots.shape[1]), out_rows.data_ptr(), C.byref(err))
        self._native.check(err)

    def build_upper(self, fingerprints, levels_minus_one: np.ndarray, expansion_add: int):
        """A fresh build over the gathered upper-level nodes, one level down, level-0 rows M wide ->
        (adj0 [n, M], adj_upper, entry, max_level) of that build."""
        from .index import Index
        up = Index(ndim=self.index.ndim, connectivity=self.m, expansion_add=expansion_add,
                   device=self.device.index if self.device.index is not None else -1, _connectivity_base=self.m)
        fp = fingerprints.to(self.device).contiguous()
        self._settle()
        up.add_device(None, fp.data_ptr(), len(levels_minus_one), levels=levels_minus_one)
        a0, au = up._debug_adjacency()
        return a0, au, up.entry_slot, up.max_level


class TorchComm:
    """The collectives of the stitch over ``torch.distributed`` (NCCL between GPUs, gloo in the CPU tests)."""

    def __init__(self):
        import torch.distributed as dist
        self._dist = dist if dist.is_available() and dist.is_initialized() else None
        self.world = self._dist.get_world_size() if self._dist else 1
        self.rank = self._dist.get_rank() if self._dist else 0

    # every tensor travels as raw bytes: NCCL has no 16-bit integer type, and the payloads are slots, levels and
    # fingerprints, never numbers to reduce
    @staticmethod
    def _bytes(t):
        import torch
        return t.contiguous().view(-1).view(torch.uint8)

    def all_gather(self, t):
        import torch
        if self.world == 1:
            return t.unsqueeze(0)
        out = torch.empty((self.world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
        self._dist.all_gather_into_tensor(self._bytes(out), self._bytes(t))
        return out

    def all_to_all(self, send):
        """send[h] goes to rank h; returns recv with recv[h] = what rank h sent here"""
        import torch
        send = send.contiguous()
        recv = torch.empty_like(send)
        self._dist.all_to_all_single(self._bytes(recv), self._bytes(send))
        return recv

    def gather0(self, t):
        import torch
        if self.world == 1:
            return [t]
        out = [torch.empty_like(t) for _ in range(self.world)] if self.rank == 0 else None
        self._dist.gather(self._bytes(t), None if out is None else [self._bytes(o) for o in out], dst=0)
        return out

    def barrier(self):
        if self._dist is not None:
            self._dist.barrier()


class LocalComm:
    """The same collectives between THREADS of one process: several shards (handles) on one GPU, as the single-GPU
    tests run the stitch.  ``LocalComm(world).rank_view(r)`` is rank r's end."""

    def __init__(self, world: int):
        import threading
        self.world = world
        self._box = [None] * world
        self._bar = threading.Barrier(world)

    def rank_view(self, rank: int):
        return _LocalRank(self, rank)


class _LocalRank:
    def __init__(self, hub: LocalComm, rank: int):
        self._hub, self.rank, self.world = hub, rank, hub.world

    def _exchange(self, t):
        self._hub._box[self.rank] = t
        self._hub._bar.wait()
        got = list(self._hub._box)
        self._hub._bar.wait()
        return got

    def all_gather(self, t):
        import torch
        return torch.stack(self._exchange(t))

    def all_to_all(self, send):
        import torch
        return torch.stack([part[self.rank] for part in self._exchange(send)])

    def gather0(self, t):
        got = self._exchange(t)
        return got if self.rank == 0 else None

    def barrier(self):
        self._hub._bar.wait()


def _stitch_rows(shard, slot_base: int, comm, batch: int, kr: int, efm: int):
    """The stitch of ONE graph level held as level 0 of ``shard``'s index: every rank's rows after pruning own links +
    the other shards' answers -> int32 cuda tensor [n_local, row width] of slots in the numbering ``slot_base`` belongs
    to (EMPTY_SLOT padded).  Collective; the sub-graphs themselves stay untouched."""
    import torch
    world, rank, dev = comm.world, comm.rank, shard.device
    n_local, m0 = len(shard), shard.m0
    sizes = comm.all_gather(torch.tensor([n_local], dtype=torch.int64, device=dev)).view(-1).tolist()
    rows_global = torch.full((n_local, m0), EMPTY_SLOT, dtype=torch.int32, device=dev)
    others = [h for h in range(world) if h != rank]
    batch = max(1, min(batch, max(sizes)))
    for r0 in range(0, max(sizes), batch):
        cnt = max(0, min(batch, n_local - r0))
        q = torch.zeros((batch, 128), dtype=torch.uint8, device=dev)
        if cnt:
            shard.vectors(r0, cnt, q)
        if world == 1:
            cand_slots = torch.full((cnt, 0), EMPTY_SLOT, dtype=torch.int32, device=dev)
            cand_fp = torch.zeros((cnt, 0, 128), dtype=torch.uint8, device=dev)
        else:
            allq = comm.all_gather(q)                                       # [world, batch, 128]
            counts = [max(0, min(batch, sizes[h] - r0)) for h in range(world)]
            send_slots = torch.full((world, batch, kr), EMPTY_SLOT, dtype=torch.int32, device=dev)
            send_fp = torch.zeros((world, batch, kr, 128), dtype=torch.uint8, device=dev)
            for src in others:
                if not counts[src] or not n_local:
                    continue
                s, f = shard.remote_candidates(allq[src, :counts[src]].contiguous(), kr, efm)
                send_slots[src, :counts[src]] = torch.where(s != EMPTY_SLOT, s + int(slot_base), s)
                send_fp[src, :counts[src]] = f
            if dev.type == "cuda":
                torch.cuda.synchronize(dev)
            recv_slots, recv_fp = comm.all_to_all(send_slots), comm.all_to_all(send_fp)
            # recv[h] = what shard h found for MY rows; own block dropped
            cand_slots = recv_slots[others].permute(1, 0, 2).reshape(batch, len(others) * kr)[:cnt].contiguous()
            cand_fp = recv_fp[others].permute(1, 0, 2, 3).reshape(batch, len(others) * kr, 128)[:cnt].contiguous()
        if cnt:
            shard.stitch(r0, cnt, int(slot_base), cand_slots, cand_fp, rows_global[r0:r0 + cnt])
    return rows_global, sizes


def stitch_round_size(world: int, k_remote: int) -> int:
    """Nodes per rank and round of the stitch.  The graph does not depend on it; larger rounds mean fewer launch tails
    and synchronisations (262 144 instead of 65 536: 9 % less stitch time at two shards).  The candidate exchange of a
    round is world x round x k_remote x 128 bytes per rank and is kept to 1 GiB, the largest measured."""
    return 262144 if world * k_remote * 128 * 262144 <= (1 << 30) else 65536


def build_global_graph(shard, slot_base: int, *, expansion_add: int = 128, batch: int | None = None,
                       k_remote: int | None = None, ef_merge: int | None = None, timings: dict | None = None,
                       comm=None, assemble: bool = True, gather_below: int = 262144, rebuild_upper: bool = False):
    """ONE HNSW graph with global slots from the sub-graphs every rank built over its own row shard (SURVEY.md §8e
    row 3).  Collective: every rank calls it with its ``DeviceShard`` and the global slot of its first row; shards
    are contiguous and ordered by rank.

    LEVEL 0 -- the stitch (``_stitch_rows``), in rounds of ``batch`` nodes per rank: all-gather of the round's
    fingerprints (NCCL); every rank answers the other ranks' rows from its own graph (``k_remote`` nearest at ef
    ``ef_merge``) and sends the answers WITH their fingerprints back (all-to-all); every rank prunes own links + remote
    candidates with the reference's heuristic into rows of global slots (``rb_stitch_device``).
    LEVEL d >= 1 -- the same, recursively: every rank takes the sub-index of ITS nodes with level >= d, one level down
    (its level 0 = graph level d, rows M wide) -- the links its own build already made for that level
    (``rb_extract_level``; ``rebuild_upper=True`` builds it afresh instead, the first form of this code, 1 s slower per
    10 M rows) -- and those sub-indexes are stitched like level 0; 1/M of the work of the level below, on all GPUs at
    once.  Once at most ``gather_below`` nodes remain in total they are gathered on rank 0
    and all remaining levels are built there as one graph.

    Returns on rank 0 ``dict(levels, adj0, adj_upper, entry, max_level)`` (host arrays, global slots), ``None`` on the
    other ranks.  With ``assemble=False`` nothing is gathered: every rank gets ``dict(rows=int32 cuda tensor
    [n_local, 2M], upper=[(graph level, rows int32 [n_d, M], global slots of those nodes int64 [n_d]), ...])`` -- all
    rows in GLOBAL slots, resident on the GPU that owns the nodes -- and rank 0 also ``top`` (the rank-0 build) --
    the device-resident result ``bench.py`` times."""
    import time

    import torch
    from .index import Index
    comm = TorchComm() if comm is None else comm
    world, rank = comm.world, comm.rank
    dev = shard.device
    kr, efm = stitch_parameters(world)
    kr, efm = k_remote or kr, ef_merge or efm
    if batch is None:
        batch = stitch_round_size(world, kr)
    n_local, m = len(shard), shard.m
    t_start = time.perf_counter()

    def sync():
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)

    def counts_of(n):
        return comm.all_gather(torch.tensor([n], dtype=torch.int64, device=dev)).view(-1).tolist()

    def padded(t, cap, fill):
        out = torch.full((cap,) + tuple(t.shape[1:]), fill, dtype=t.dtype, device=dev)
        out[:len(t)] = t
        return out

    rows0, sizes = _stitch_rows(shard, int(slot_base), comm, batch, kr, efm)
    bases = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    assert int(bases[rank]) == int(slot_base), "shards must be contiguous and ordered by rank"
    assert int(bases[-1]) < (1 << 31), "the stitch addresses at most 2^31 - 1 global slots"
    sync()
    comm.barrier()
    t_stitch = time.perf_counter()
    # ---- upper levels
    lv = np.asarray(shard.levels(), dtype=np.int16)
    pieces, piece_counts, keep, top, top_built = [], [], [], None, False
    detail = []  # per stitched upper level: nodes in total, seconds of this rank's sub-index build and of its stitch
    cur, cur_lv = shard, lv
    cur_glob = torch.arange(n_local, dtype=torch.int64, device=dev) + int(slot_base)
    depth = 0
    while True:
        up_idx = np.flatnonzero(cur_lv >= 1)
        n_up = counts_of(len(up_idx))
        total = int(sum(n_up))
        if total == 0:
            break
        depth += 1
        idx32 = torch.from_numpy(up_idx.astype(np.int32)).to(dev)
        up_glob = cur_glob[idx32.long()]
        sub_lv = (cur_lv[up_idx] - 1).astype(np.int16)
        cap = max(max(n_up), 1)
        gather_top = world == 1 or total <= gather_below
        fp = None
        if gather_top or rebuild_upper:
            fp = cur.gather_vectors(idx32) if len(up_idx) else torch.zeros((0, 128), dtype=torch.uint8, device=dev)
        if gather_top:
            top_built = True  # (on every rank: the collectives below must stay in step)
            g_glob = comm.gather0(padded(up_glob, cap, -1))
            g_fp = comm.gather0(padded(fp, cap, 0))
            g_lv = comm.gather0(padded(torch.from_numpy(sub_lv).to(dev), cap, 0))
            if rank == 0:
                t_a = time.perf_counter()
                t_glob = torch.cat([g_glob[h][:n_up[h]] for h in range(world)]).cpu().numpy()
                t_lv = torch.cat([g_lv[h][:n_up[h]] for h in range(world)]).cpu().numpy().astype(np.int16)
                t_fp = torch.cat([g_fp[h][:n_up[h]].to(dev) for h in range(world)])
                a0, au, sub_entry, _ = shard.build_upper(t_fp, t_lv, expansion_add)
                top = dict(level=depth, slots=t_glob, sub_levels=t_lv, a0=a0, au=au, entry=int(t_glob[sub_entry]))
                detail.append(dict(level=depth, nodes=total, build_s=time.perf_counter() - t_a, stitch_s=0.0,
                                   gathered_on_rank_0=True))
            break
        t_a = time.perf_counter()
        sub = Index(ndim=shard.index.ndim, connectivity=m, expansion_add=expansion_add,
                    device=dev.index if dev.index is not None else -1, _connectivity_base=m)
        if len(up_idx):
            shard._settle()
            if rebuild_upper:
                sub.add_device(None, fp.data_ptr(), len(up_idx), levels=sub_lv)
            else:
                sub.extract_upper_levels(cur.index)
                assert len(sub) == len(up_idx)
        ushard = DeviceShard(sub, dev)
        sync()
        t_b = time.perf_counter()
        rows_u, _ = _stitch_rows(ushard, int(sum(n_up[:rank])), comm, batch, kr, efm)
        sync()
        detail.append(dict(level=depth, nodes=total, build_s=t_b - t_a, stitch_s=time.perf_counter() - t_b))
        # this level's own numbering (rank-major, ascending slot) -> global slots
        allg = comm.all_gather(padded(up_glob, cap, -1))
        table = torch.cat([allg[h, :n_up[h]] for h in range(world)] + [torch.tensor([-1], dtype=torch.int64, device=dev)])
        rows_g = table[torch.where(rows_u < 0, torch.full_like(rows_u, total), rows_u).long()].to(torch.int32)
        pieces.append((depth, rows_g, up_glob))
        piece_counts.append(n_up)
        keep.append(sub)
        cur, cur_lv, cur_glob = ushard, sub_lv, up_glob
    entry = None
    if not top_built and pieces:  # every level was stitched: the entry is the lowest slot of the top level
        lo = comm.all_gather(torch.tensor([int(pieces[-1][2].min()) if len(pieces[-1][2]) else (1 << 62)],
                                          dtype=torch.int64, device=dev))
        entry = int(lo.min())
    sync()
    comm.barrier()
    t_upper = time.perf_counter()
    if timings is not None:
        timings["stitch_s"] = t_stitch - t_start
        timings["upper_levels_s"] = t_upper - t_stitch
        timings["k_remote"], timings["ef_merge"], timings["stitched_levels"] = kr, efm, 1 + len(pieces)
        timings["upper_detail"] = detail
    if not assemble:
        out = dict(rows=rows0, upper=pieces)
        if rank == 0:
            out["top"], out["entry"] = top, top["entry"] if top else (entry or 0)
        return out
    result = _assemble_global(comm, dev, sizes, bases, lv, rows0, pieces, piece_counts, top, entry, m)
    sync()
    comm.barrier()
    if timings is not None:
        timings["assemble_s"] = time.perf_counter() - t_upper
    return result


def _assemble_global(comm, dev, sizes, bases, lv, rows0, pieces, piece_counts, top, entry, m):
    """Gather every rank's rows on rank 0 and lay them out as the host arrays of ONE graph (global slots)."""
    import torch
    world, rank = comm.world, comm.rank
    n_local, pad = len(lv), max(max(sizes), 1)

    def padded(t, cap, fill):
        out = torch.full((cap,) + tuple(t.shape[1:]), fill, dtype=t.dtype, device=dev)
        out[:len(t)] = t
        return out

    g_rows = comm.gather0(padded(rows0, pad, EMPTY_SLOT))
    g_levels = comm.gather0(padded(torch.from_numpy(lv).to(dev), pad, 0))
    g_pieces = []
    for (level, rows_g, up_glob), n_up in zip(pieces, piece_counts):
        cap = max(max(n_up), 1)
        g_pieces.append((level, n_up, comm.gather0(padded(rows_g, cap, EMPTY_SLOT)),
                         comm.gather0(padded(up_glob, cap, -1))))
    if rank != 0:
        return None
    n_total = int(bases[-1])
    levels = np.concatenate([g_levels[h][:sizes[h]].cpu().numpy() for h in range(world)]).astype(np.int16)
    adj0 = np.concatenate([g_rows[h][:sizes[h]].cpu().numpy() for h in range(world)]).view(np.uint32)
    upper_off = np.concatenate([[0], np.cumsum(levels.astype(np.int64))])
    adj_upper = np.full((int(upper_off[-1]), m), 0xFFFFFFFF, dtype=np.uint32)
    for level, n_up, rows_list, glob_list in g_pieces:
        rows = np.concatenate([rows_list[h][:n_up[h]].cpu().numpy() for h in range(world)]).view(np.uint32)
        glob = np.concatenate([glob_list[h][:n_up[h]].cpu().numpy() for h in range(world)])
        adj_upper[upper_off[glob] + level - 1] = rows
    if top is not None:
        t_glob, t_lv, dt = top["slots"], top["sub_levels"].astype(np.int64), top["level"]
        table = np.append(t_glob, 0xFFFFFFFF).astype(np.uint32)  # sub slot -> global slot, EMPTY -> EMPTY

        def to_global(rows):
            return table[np.where(rows == 0xFFFFFFFF, len(t_glob), rows).astype(np.int64)]

        first_row = upper_off[t_glob] + dt - 1  # the node's row on graph level dt = level 0 of the rank-0 build
        adj_upper[first_row] = to_global(top["a0"])
        if len(top["au"]):
            sub_off = np.concatenate([[0], np.cumsum(t_lv)])
            owner = np.repeat(np.arange(len(t_glob)), t_lv)
            rel = np.arange(len(top["au"])) - sub_off[owner]  # 0 = graph level dt + 1
            adj_upper[first_row[owner] + 1 + rel] = to_global(top["au"])
        entry = top["entry"]
    max_level = int(levels.max()) if n_total else 0
    return dict(levels=levels, adj0=adj0, adj_upper=adj_upper, entry=int(entry or 0), max_level=max_level)
