"""Host-side logic of the N > 1 path (rad_b200/sharded.py) with world_size-2 gloo on CPU: shard bounds, the 64-bit key
format (ascending key order == the reference's result order), the all-gather layout and the k-way merge rule
(usearch index.hpp:2625-2645).  The per-shard top-k lists come from the oracle here (the checker); on the GPU the same
keys are produced by rb_exact_topk_device and merged by rb_merge_topk_device (tests/test_gpu_exact.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle as O
from rad_b200 import sharded, synth


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_bounds_cover_rows_exactly():
    for n, w in [(10, 3), (7, 8), (100_000_000, 8), (1, 2), (0, 4)]:
        spans = [sharded.shard_bounds(n, w, r) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
        sizes = [hi - lo for lo, hi in spans]
        assert max(sizes) - min(sizes) <= 1


def test_key_order_is_reference_order():
    db = synth.clustered_fingerprints(3000, seed=31)
    q = synth.queries_from_db(db, 40, seed=2)
    slots, dists, counts = O.exact_search(db, q, 10)
    keys = sharded.pack_topk_keys(dists, slots, counts)
    assert (keys[:, 1:] > keys[:, :-1]).all()  # the oracle's order is strictly ascending in key space
    s2, d2, c2 = sharded.unpack_topk_keys(keys)
    assert np.array_equal(s2, slots.astype(np.uint64)) and np.array_equal(d2.view(np.uint32), dists.view(np.uint32))
    assert np.array_equal(c2, counts)


def _worker(rank, world, port, n, nq, k, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        db = synth.clustered_fingerprints(n, seed=17)
        q = synth.heldout_queries(nq, seed=17, n_total=n)
        lo, hi = sharded.shard_bounds(n, world, rank)
        slots, dists, counts = O.exact_search(db[lo:hi], q, k)          # shard-local answers (local slots)
        keys = sharded.pack_topk_keys(dists, slots.astype(np.uint64) + np.uint64(lo), counts)
        part = torch.from_numpy(keys.view(np.int64).copy())
        gathered = torch.empty((world, nq, k), dtype=torch.int64)
        dist.all_gather_into_tensor(gathered.view(-1), part.view(-1))
        merged = sharded.merge_topk_host(gathered.numpy(), k)
        np.save(os.path.join(out_dir, f"merged_{rank}.npy"), merged)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,nq,k", [(2001, 33, 10), (5, 4, 10)])
def test_two_rank_gather_and_merge_equals_unsharded(tmp_path, n, nq, k):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n, nq, k, str(tmp_path)), nprocs=world, join=True)
    db = synth.clustered_fingerprints(n, seed=17)
    q = synth.heldout_queries(nq, seed=17, n_total=n)
    slots, dists, counts = O.exact_search(db, q, k)
    want = sharded.pack_topk_keys(dists, slots, counts)
    a = np.load(tmp_path / "merged_0.npy")
    b = np.load(tmp_path / "merged_1.npy")
    assert np.array_equal(a, b)      # every rank holds the same merged answer
    assert np.array_equal(a, want)   # and it is the unsharded reference answer, ties included


class _OracleShard:
    """Stands in for rad_b200.index.Index on a rank without a GPU: same search / key-lookup surface, answers from the
    oracle over this rank's rows; user keys are global row + 1000 (not the slots)."""

    def __init__(self, rows, lo):
        self.rows, self.lo = rows, lo

    def __len__(self):
        return len(self.rows)

    def search(self, vectors, count=10, *, exact=False):
        from rad_b200.index import BatchMatches
        slots, dists, counts = O.exact_search(self.rows, vectors, count)
        keys = slots.astype(np.uint64) + np.uint64(self.lo + 1000)
        return BatchMatches(keys=keys, distances=dists, counts=counts)

    def get_node_ids_from_keys(self, keys):
        return np.asarray(keys, dtype=np.int64) - (self.lo + 1000)


def _sharded_index_worker(rank, world, port, n, nq, k, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        db = synth.clustered_fingerprints(n, seed=23)
        q = synth.heldout_queries(nq, seed=23, n_total=n)
        lo, hi = sharded.shard_bounds(n, world, rank)
        ix = sharded.ShardedIndex(_OracleShard(db[lo:hi], lo), lo)
        assert len(ix) == n
        got = ix.search(q, k, exact=True)
        np.savez(os.path.join(out_dir, f"sharded_{rank}.npz"), keys=got.keys, distances=got.distances,
                 counts=got.counts)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,nq,k", [(1501, 21, 10), (7, 3, 10)])
def test_sharded_index_search_equals_unsharded(tmp_path, n, nq, k):
    """ShardedIndex.search over two ranks == one search over all rows: user keys, float bits, counts, tie order."""
    world = 2
    mp.spawn(_sharded_index_worker, args=(world, _free_port(), n, nq, k, str(tmp_path)), nprocs=world, join=True)
    db = synth.clustered_fingerprints(n, seed=23)
    q = synth.heldout_queries(nq, seed=23, n_total=n)
    slots, dists, counts = O.exact_search(db, q, k)
    a = np.load(tmp_path / "sharded_0.npz")
    b = np.load(tmp_path / "sharded_1.npz")
    for name in ("keys", "distances", "counts"):
        assert np.array_equal(a[name], b[name])
    assert np.array_equal(a["counts"], counts)
    for r in range(nq):
        c = int(counts[r])
        assert np.array_equal(a["keys"][r, :c], slots[r, :c].astype(np.uint64) + np.uint64(1000))
        assert np.array_equal(a["distances"][r, :c].view(np.uint32), dists[r, :c].view(np.uint32))


def _comm_worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from rad_b200.sharded import TorchComm
        comm = TorchComm()
        g = comm.all_gather(torch.tensor([rank * 10, rank * 10 + 1]))
        send = torch.stack([torch.tensor([100 * rank + h, 100 * rank + h + 50]) for h in range(world)])
        recv = comm.all_to_all(send)
        got0 = comm.gather0(torch.tensor([rank + 7]))
        comm.barrier()
        out.put((rank, g.tolist(), recv.tolist(), None if got0 is None else [t.tolist() for t in got0]))
    finally:
        dist.destroy_process_group()


def test_stitch_collectives_over_gloo():
    """TorchComm (what the stitch uses between the ranks) against the semantics LocalComm gives the single-GPU tests:
    all_gather stacks by rank, all_to_all delivers send[h] to rank h, gather0 collects on rank 0."""
    import multiprocessing as pymp
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = pymp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_comm_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted(out.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(30)
    for rank, g, recv, g0 in got:
        assert g == [[0, 1], [10, 11]]
        assert recv == [[100 * h + rank, 100 * h + rank + 50] for h in range(world)]
        assert g0 == ([[7], [8]] if rank == 0 else None)
