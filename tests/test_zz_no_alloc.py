"""rb_exact_topk_device_noalloc (include/rad_b200.h): the exact Tanimoto top-k search with its workspace in a static
device array, for callers that cannot call cudaMalloc.  It has no counterpart in the reference;
the yardstick is the existing entry point, rb_exact_topk_device, which the other tests pin to
the oracle bit for bit.  Integer / bit work: the results must be IDENTICAL, on two calls in a row (the second sees
whatever the first left in the shared array).

Collected last on purpose; passes alone in a fresh process and after the others."""
import ctypes as C

import numpy as np
import pytest

from rad_b200 import _native, synth

pytestmark = pytest.mark.gpu


def _keys_by_existing_entry(L, torch, d_db, rows, slot_base, d_q, nq, k):
    need = int(L.rb_exact_topk_workspace_bytes(rows, nq, k))
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    out = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    err = C.c_char_p()
    stream = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()
    L.rb_exact_topk_device(d_db.data_ptr(), rows, slot_base, d_q.data_ptr(), nq, k, out.data_ptr(), ws.data_ptr(), need,
                           stream, C.byref(err))
    _native.check(err)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_no_alloc_entry_equals_the_existing_entry_on_two_calls_in_a_row():
    import torch
    L = _native.lib()
    assert int(L.rb_exact_topk_noalloc_max_bytes()) == 16 << 20
    rows = 300_000
    d_db = synth.named_set_device("dense", rows, seed=31)
    err = C.c_char_p()
    stream = torch.cuda.current_stream().cuda_stream
    # two different searches back to back through the SAME static array: the POPC kernel (few pairs) and the tensor-core
    # kernel (>= 2e9 pairs), different k, different slot bases; then the first one again
    cases = [(200, 10, 0, 3), (7000, 12, 1_000_000, 4), (200, 10, 0, 3)]
    outs = [torch.empty((nq, k), dtype=torch.int64, device="cuda") for nq, k, _, _ in cases]  # inputs and outputs are
    queries = []                                                                              # the caller's
    for nq, _, _, seed in cases:
        pick = torch.from_numpy(np.random.default_rng(seed).choice(rows, size=nq, replace=False)).cuda()
        queries.append(d_db[pick].contiguous())
    torch.cuda.synchronize()
    for (nq, k, base, _), d_q, out in zip(cases, queries, outs):
        assert int(L.rb_exact_topk_workspace_bytes(rows, nq, k)) <= 16 << 20
        L.rb_exact_topk_device_noalloc(d_db.data_ptr(), rows, base, d_q.data_ptr(), nq, k, out.data_ptr(), stream,
                                       C.byref(err))
        _native.check(err)
    torch.cuda.synchronize()
    for (nq, k, base, _), d_q, out in zip(cases, queries, outs):
        want = _keys_by_existing_entry(L, torch, d_db, rows, base, d_q, nq, k)
        assert np.array_equal(out.cpu().numpy(), want), (nq, k, base)
    assert np.array_equal(outs[0].cpu().numpy(), outs[2].cpu().numpy())
    # sanity of what is compared: every query is a row of the set and finds itself (or a duplicate) at distance 0
    keys = outs[0].cpu().numpy().view(np.uint64)
    assert ((keys[:, 0] >> np.uint64(32)) == 0).mean() > 0.99


def test_no_alloc_entry_rejects_what_does_not_fit():
    import torch
    L = _native.lib()
    rows, nq, k = 200_000, 4096, 64  # chunks x Q x k x 12 bytes: far beyond 16 MiB
    assert int(L.rb_exact_topk_workspace_bytes(rows, nq, k)) > 16 << 20
    d_db = torch.zeros((rows, 128), dtype=torch.uint8, device="cuda")
    d_q = torch.zeros((nq, 128), dtype=torch.uint8, device="cuda")
    out = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    err = C.c_char_p()
    torch.cuda.synchronize()
    L.rb_exact_topk_device_noalloc(d_db.data_ptr(), rows, 0, d_q.data_ptr(), nq, k, out.data_ptr(),
                                   torch.cuda.current_stream().cuda_stream, C.byref(err))
    assert err.value is not None and b"static array holds 16777216" in err.value
    with pytest.raises(Exception, match="static array"):
        _native.check(err)
