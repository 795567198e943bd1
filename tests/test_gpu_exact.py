"""GPU parity of the exact Tanimoto top-k path (K1 + K2 + K3) through the C ABI, against the oracle and the golden
vectors produced by the unmodified reference.  Bar: BIT-EXACT keys and float32 distance bits (SURVEY.md §8c rules
1-3: ascending distance, ties by descending slot, highest slots kept at the k-th boundary)."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import oracle as O
from rad_b200 import _native, synth
from rad_b200.index import Index

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["popc", "f4", "i8"])
def exact_kernel(request, monkeypatch):
    """Every test of this module runs against ALL exact-search kernels: the POPC-pipe kernel, the FP4 tensor-core
    kernel (tcgen05.mma kind::mxf4, rad_b200/csrc/exact_f4_kernels.cuh -- the default above 2e9 pairs) and the int8
    tensor-core kernel (kind::i8, exact_i8_kernels.cuh); the library reads RB_EXACT_KERNEL on every call."""
    monkeypatch.setenv("RB_EXACT_KERNEL", request.param)
    return request.param


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _check_exact(db, q, k, got, ndim=1024):
    slots, dists, counts = O.exact_search(db, q, k)
    assert np.array_equal(got.counts, counts)
    for r in range(len(q)):
        c = int(counts[r])
        assert np.array_equal(got.keys[r, :c], slots[r, :c]), f"query {r}: keys differ"
        assert np.array_equal(_bits(got.distances[r, :c]), _bits(dists[r, :c])), f"query {r}: distance bits differ"


def _index_with_rows(db, ndim=1024, m=8):
    """An index holding `db` WITHOUT running the graph build: write the file with the numpy writer and load it."""
    ix = Index(ndim=ndim, connectivity=m, expansion_add=16, expansion_search=16)
    ix.add(np.arange(len(db), dtype=np.uint64), db)
    return ix


def test_golden_reference_answers(golden_dir):
    """Load the reference's own file, search with the reference's queries, compare with the reference's answers."""
    g = np.load(os.path.join(golden_dir, "ref_n1000_m8.npz"))
    ix = Index(ndim=1024, connectivity=8, path=os.path.join(golden_dir, "ref_n1000_m8.usearch"))
    got = ix.search(g["queries"], 10, exact=True)
    assert np.array_equal(got.keys, g["exact_keys"])
    assert np.array_equal(_bits(got.distances), g["exact_dist_bits"])
    assert np.array_equal(got.counts, g["exact_counts"])


def test_golden_small_dimensions(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_n64_ndim64_m4.npz"))
    ix = Index(ndim=64, connectivity=4, path=os.path.join(golden_dir, "ref_n64_ndim64_m4.usearch"))
    got = ix.search(g["queries"], 10, exact=True)
    assert np.array_equal(got.keys, g["exact_keys"])
    assert np.array_equal(_bits(got.distances), g["exact_dist_bits"])


@pytest.mark.parametrize("n,nq,k", [(1, 1, 1), (1, 3, 10), (63, 5, 10), (64, 64, 10), (65, 65, 3), (1000, 1, 1),
                                    (5000, 130, 10), (20000, 200, 10), (20000, 7, 100), (3000, 9, 1000)])
def test_against_oracle(n, nq, k):
    db = synth.clustered_fingerprints(n, seed=n + 3)
    ix = _index_with_rows(db)
    q = synth.heldout_queries(nq, seed=n + 3, n_total=n) if nq > n else synth.queries_from_db(db, nq, seed=5)
    if len(q) < nq:
        q = synth.heldout_queries(nq, seed=n + 3, n_total=n)
    q[0] = 0  # the all-zero query: every distance is 1.0 (u == 0 or i == 0), the highest slots win
    got = ix.search(q, k, exact=True)
    _check_exact(db, q, k, got)


def test_many_ties_and_duplicates():
    """A database made of 16 distinct rows repeated: every distance is tied hundreds of times."""
    base = synth.clustered_fingerprints(16, seed=99, edge_cases=False)
    db = np.ascontiguousarray(np.tile(base, (400, 1)))
    ix = _index_with_rows(db)
    q = np.ascontiguousarray(base[:8])
    got = ix.search(q, 25, exact=True)
    _check_exact(db, q, 25, got)
    # ties are ordered by descending slot: the first hit of query j is the LAST copy of row j
    assert got.keys[3, 0] == 16 * 399 + 3 and got.distances[3, 0] == 0.0


def test_single_vector_query_returns_matches():
    db = synth.clustered_fingerprints(500, seed=2)
    ix = _index_with_rows(db)
    one = ix.search(db[17], 5, exact=True)
    slots, dists, _ = O.exact_search(db, db[17:18], 5)
    assert np.array_equal(one.keys, slots[0]) and np.array_equal(_bits(one.distances), _bits(dists[0]))


def test_strided_queries_and_nondefault_keys():
    db = synth.clustered_fingerprints(4000, seed=21)
    keys = (np.arange(4000, dtype=np.uint64) * 7 + 1000)
    ix = Index(ndim=1024, connectivity=8, expansion_add=16)
    ix.add(keys, db)
    wide = np.zeros((50, 160), dtype=np.uint8)
    wide[:, :128] = db[100:150]
    q = wide[:, :128]  # row stride 160
    got = ix.search(q, 10, exact=True)
    slots, dists, _ = O.exact_search(db, np.ascontiguousarray(q), 10)
    assert np.array_equal(got.keys, keys[slots])
    assert np.array_equal(_bits(got.distances), _bits(dists))


def _torch():
    import torch
    return torch


def test_stateless_sharded_topk_merge_equals_unsharded():
    """rb_exact_topk_device per row shard (slot_base) + rb_merge_topk_device == one pass over the whole database ==
    the oracle.  This is the N-GPU data path with the all-gather replaced by a concatenation."""
    torch = _torch()
    L = _native.lib()
    n, nq, k = 30000, 150, 10
    db = synth.clustered_fingerprints(n, seed=8)
    q = synth.heldout_queries(nq, seed=8, n_total=n)
    d_db = torch.from_numpy(db).cuda()
    d_q = torch.from_numpy(q).cuda()
    err = C.c_char_p()
    stream = torch.cuda.current_stream().cuda_stream

    def topk(lo, hi):
        ws = L.rb_exact_topk_workspace_bytes(hi - lo, nq, k)
        d_ws = torch.empty(ws, dtype=torch.uint8, device="cuda")
        out = torch.empty((nq, k), dtype=torch.int64, device="cuda")
        L.rb_exact_topk_device(d_db[lo:hi].data_ptr(), hi - lo, lo, d_q.data_ptr(), nq, k, out.data_ptr(),
                               d_ws.data_ptr(), ws, stream, C.byref(err))
        _native.check(err)
        return out

    whole = topk(0, n)
    bounds = [0, 7001, 7001 + 64, 19999, n]
    parts = torch.stack([topk(bounds[i], bounds[i + 1]) for i in range(4)])
    merged = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    L.rb_merge_topk_device(parts.data_ptr(), 4, nq, k, merged.data_ptr(), stream, C.byref(err))
    _native.check(err)
    torch.cuda.synchronize()
    assert torch.equal(whole, merged)
    slots_o = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    dist_o = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    cnt_o = torch.empty((nq,), dtype=torch.int64, device="cuda")
    L.rb_unpack_topk_device(merged.data_ptr(), nq, k, slots_o.data_ptr(), dist_o.data_ptr(), cnt_o.data_ptr(), stream,
                            C.byref(err))
    _native.check(err)
    torch.cuda.synchronize()
    slots, dists, counts = O.exact_search(db, q, k)
    assert np.array_equal(slots_o.cpu().numpy().astype(np.uint64), slots.astype(np.uint64))
    assert np.array_equal(_bits(dist_o.cpu().numpy()), _bits(dists))
    assert np.array_equal(cnt_o.cpu().numpy(), counts)


def test_device_generator_matches_numpy_generator():
    torch = _torch()
    L = _native.lib()
    n_total, first, rows, seed = 250000, 123457, 3000, 4
    centres = synth.centre_table(seed, n_total)
    d_c = torch.from_numpy(centres).cuda()
    out = torch.empty((rows, 128), dtype=torch.uint8, device="cuda")
    err = C.c_char_p()
    L.rb_synth_fingerprints_device(out.data_ptr(), first, rows, seed, d_c.data_ptr(), len(centres), 1,
                                   torch.cuda.current_stream().cuda_stream, C.byref(err))
    _native.check(err)
    torch.cuda.synchronize()
    want = synth.clustered_fingerprints(rows, seed, start=first, n_total=n_total, centres=centres)
    assert np.array_equal(out.cpu().numpy(), want)
    head = torch.empty((4, 128), dtype=torch.uint8, device="cuda")
    L.rb_synth_fingerprints_device(head.data_ptr(), 0, 4, seed, d_c.data_ptr(), len(centres), 1,
                                   torch.cuda.current_stream().cuda_stream, C.byref(err))
    _native.check(err)
    assert np.array_equal(head.cpu().numpy(), synth.clustered_fingerprints(4, seed, n_total=n_total, centres=centres))


def test_full_size_properties():
    """Properties that hold at any size (checked at 2M rows x 512 queries, beyond what the oracle scans quickly):
    self-queries come back first at distance 0; sharded == unsharded; the k-th distances are monotone in k."""
    torch = _torch()
    L = _native.lib()
    n, nq, k, seed = 2_000_000, 512, 10, 6
    centres = synth.centre_table(seed, n)
    d_c = torch.from_numpy(centres).cuda()
    d_db = torch.empty((n, 128), dtype=torch.uint8, device="cuda")
    err = C.c_char_p()
    stream = torch.cuda.current_stream().cuda_stream
    L.rb_synth_fingerprints_device(d_db.data_ptr(), 0, n, seed, d_c.data_ptr(), len(centres), 0, stream, C.byref(err))
    _native.check(err)
    pick = torch.arange(nq, device="cuda") * 3907 + 11
    d_q = d_db[pick].contiguous()

    def topk(lo, hi, kk):
        ws = L.rb_exact_topk_workspace_bytes(hi - lo, nq, kk)
        d_ws = torch.empty(ws, dtype=torch.uint8, device="cuda")
        out = torch.empty((nq, kk), dtype=torch.int64, device="cuda")
        L.rb_exact_topk_device(d_db[lo:hi].data_ptr(), hi - lo, lo, d_q.data_ptr(), nq, kk, out.data_ptr(),
                               d_ws.data_ptr(), ws, stream, C.byref(err))
        _native.check(err)
        return out

    whole = topk(0, n, k)
    halves = torch.stack([topk(0, n // 2 + 13, k), topk(n // 2 + 13, n, k)])
    merged = torch.empty_like(whole)
    L.rb_merge_topk_device(halves.data_ptr(), 2, nq, k, merged.data_ptr(), stream, C.byref(err))
    _native.check(err)
    torch.cuda.synchronize()
    assert torch.equal(whole, merged)
    w = whole.cpu().numpy().view(np.uint64)
    assert (np.diff(w.astype(np.uint64), axis=1).astype(np.int64) > 0).all() or (w[:, 1:] > w[:, :-1]).all()
    dist = (w >> np.uint64(32)).astype(np.uint32).view(np.float32)
    slot = (~w).astype(np.uint32)
    # each query is a database row: distance 0 first, and its own slot is among the zero-distance hits
    assert (dist[:, 0] == 0).all()
    own = pick.cpu().numpy().astype(np.uint32)
    assert all(own[r] in slot[r][dist[r] == 0] or (dist[r] == 0).all() for r in range(nq))
    # spot-check 8 queries against the oracle over the full 2M rows
    db_host = d_db.cpu().numpy()
    qs = d_q[:8].cpu().numpy()
    slots, dists, _ = O.exact_search(db_host, qs, k)
    assert np.array_equal(slot[:8], slots.astype(np.uint32))
    assert np.array_equal(dist[:8].view(np.uint32), _bits(dists))
    top1 = topk(0, n, 1).cpu().numpy().view(np.uint64)
    assert np.array_equal(top1[:, 0], w[:, 0])
