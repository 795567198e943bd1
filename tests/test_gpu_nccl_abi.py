"""rb_sharded_search_device: the multi-GPU search exchange behind the C ABI (NCCL through dlopen, no torch.distributed).
World size 1 runs on any GPU box; world size 2 needs two GPUs (skipped otherwise).  The communicators are made with
NCCL's own C calls through ctypes -- what a C or C++ host would do."""
import ctypes as C
import multiprocessing
import os

import numpy as np
import pytest

from rad_b200 import synth

pytestmark = pytest.mark.gpu


class _UniqueId(C.Structure):
    _fields_ = [("internal", C.c_byte * 128)]     # ncclUniqueId (nccl.h: NCCL_UNIQUE_ID_BYTES = 128)


def _nccl():
    import torch  # noqa: F401 -- its bundled libnccl.so.2 is then the copy in the process
    lib = C.CDLL("libnccl.so.2")
    lib.ncclGetUniqueId.argtypes = [C.POINTER(_UniqueId)]
    lib.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, _UniqueId, C.c_int]
    lib.ncclCommDestroy.argtypes = [C.c_void_p]
    return lib


def _rank_main(rank, world, id_bytes, db, queries, k, bounds, out_queue):
    try:
        import torch
        from rad_b200 import _native
        from rad_b200.index import Index
        torch.cuda.set_device(rank)
        nccl = _nccl()
        uid = _UniqueId()
        C.memmove(C.byref(uid), id_bytes, 128)
        comm = C.c_void_p()
        assert nccl.ncclCommInitRank(C.byref(comm), world, uid, rank) == 0
        lo, hi = bounds[rank], bounds[rank + 1]
        ix = Index(ndim=1024, connectivity=8, expansion_add=400, expansion_search=128, device=rank)
        if hi > lo:
            ix.add(np.arange(lo, hi, dtype=np.uint64), db[lo:hi])
        lib = _native.lib()
        dev = torch.device("cuda", rank)
        q = torch.from_numpy(queries).to(dev).contiguous()
        nq = len(queries)
        results = {}
        for exact in (1, 0):
            keys = torch.empty((nq, k), dtype=torch.int64, device=dev)
            need = int(lib.rb_sharded_search_workspace_bytes(nq, k, world))
            ws = torch.empty((need,), dtype=torch.uint8, device=dev)
            torch.cuda.synchronize(dev)
            err = C.c_char_p()
            lib.rb_sharded_search_device(ix._handle(), lo, q.data_ptr(), nq, k, exact, comm.value, keys.data_ptr(),
                                         ws.data_ptr(), need, C.byref(err))
            _native.check(err)
            results[exact] = keys.cpu().numpy().view(np.uint64)
        nccl.ncclCommDestroy(comm)
        out_queue.put((rank, results, None))
    except Exception as e:  # noqa: BLE001
        import traceback
        out_queue.put((rank, None, traceback.format_exc() + repr(e)))


def _run(world, empty_last=False):
    n, nq, k = 40000, 200, 10
    db = synth.clustered_fingerprints(n, seed=13)
    queries = synth.queries_from_db(db, nq, seed=2)
    bounds = [n * r // world for r in range(world + 1)]
    if empty_last:
        bounds[-2] = n                        # the last rank holds no rows
    uid = _UniqueId()
    assert _nccl().ncclGetUniqueId(C.byref(uid)) == 0
    ctx = multiprocessing.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_rank_main, args=(r, world, bytes(uid.internal), db, queries, k, bounds, out))
             for r in range(world)]
    for p in procs:
        p.start()
    got = {}
    try:
        for _ in range(world):
            rank, res, problem = out.get(timeout=240)
            assert problem is None, problem
            got[rank] = res
    finally:
        for p in procs:
            p.join(20)
            if p.is_alive():
                p.kill()
    # the exact answers are the answers of the whole set: ascending distance, ties by descending slot
    from oracle import oracle as O
    slots, dists, _ = O.exact_search(db, queries, k)
    want = (dists.view(np.uint32).astype(np.uint64) << np.uint64(32)) | (~slots.astype(np.uint32)).astype(np.uint64)
    for rank in range(world):
        assert np.array_equal(got[rank][1], want), rank
        assert np.array_equal(got[rank][0], got[0][0])       # every rank holds the same merged answer
    approx = got[0][0]
    found = (approx >> np.uint64(32)).astype(np.uint32).view(np.float32)
    recall = float((found <= dists[:, k - 1:k]).mean())
    assert recall > 0.9, recall


def test_world_of_one():
    _run(1)


def test_two_ranks_and_an_empty_shard():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    _run(2)
    _run(2, empty_last=True)
