#!/usr/bin/env python3
"""Exploratory timings on the GPU box (not the bench): build and search throughput at a few sizes."""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from rad_b200 import _native, synth  # noqa: E402
from rad_b200.index import Index  # noqa: E402

L = _native.lib()
err = C.c_char_p()


def gen(n, seed=0, first=0, n_total=None, edge=1):
    n_total = n_total or n
    centres = synth.centre_table(seed, n_total)
    d_c = torch.from_numpy(centres).cuda()
    out = torch.empty((n, 128), dtype=torch.uint8, device="cuda")
    L.rb_synth_fingerprints_device(out.data_ptr(), first, n, seed, d_c.data_ptr(), len(centres), edge,
                                   torch.cuda.current_stream().cuda_stream, C.byref(err))
    _native.check(err)
    torch.cuda.synchronize()
    return out


def exact(d_db, d_q, k=10, reps=3):
    n, nq = d_db.shape[0], d_q.shape[0]
    ws = L.rb_exact_topk_workspace_bytes(n, nq, k)
    d_ws = torch.empty(ws, dtype=torch.uint8, device="cuda")
    out = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        L.rb_exact_topk_device(d_db.data_ptr(), n, 0, d_q.data_ptr(), nq, k, out.data_ptr(), d_ws.data_ptr(), ws, s,
                               C.byref(err))
        _native.check(err)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best, out


sizes = [int(a) for a in sys.argv[1:]] or [100_000, 1_000_000]
for n in sizes:
    d_db = gen(n)
    for nq in (1000, 10000):
        d_q = gen(nq, first=1 << 40, n_total=n, edge=0)
        ms, _ = exact(d_db, d_q)
        print(f"exact  n={n} q={nq}: {ms:.2f} ms  {n * nq / ms / 1e6:.1f} G pairs/s", flush=True)
    for ef_add in (400,):
        ix = Index(ndim=1024, connectivity=8, expansion_add=ef_add, expansion_search=64, lazy_host_sync=True,
                   bloom_bits_log2=int(os.environ.get("RB_BLOOM", "0")))
        t = time.perf_counter()
        L.rb_add_device(ix._handle(), None, d_db.data_ptr(), n, C.byref(err))
        _native.check(err)
        dt = time.perf_counter() - t
        c = ix.last_counters
        print(f"build  n={n} ef_add={ef_add}: wall {dt:.2f} s device {c['device_ms'] / 1e3:.2f} s "
              f"{n / dt:.0f} vec/s  counters={c}", flush=True)
        t = time.perf_counter()
        ix.sync_host()
        print(f"sync_host {time.perf_counter() - t:.2f} s", flush=True)
        d_q = d_db[torch.randperm(n, device="cuda")[:10000]].contiguous()
        ms, truth = exact(d_db, d_q, reps=1)
        tk = truth.cpu().numpy().view(np.uint64)
        kth = (tk[:, 9] >> np.uint64(32)).astype(np.uint32).view(np.float32)
        for ef in (64, 128):
            ix.expansion_search = ef
            sl = torch.empty((10000, 10), dtype=torch.int32, device="cuda")
            ds = torch.empty((10000, 10), dtype=torch.float32, device="cuda")
            ct = torch.empty((10000,), dtype=torch.int32, device="cuda")
            for _ in range(2):
                L.rb_search_device(ix._handle(), d_q.data_ptr(), 10000, 10, 0, sl.data_ptr(), ds.data_ptr(),
                                   ct.data_ptr(), C.byref(err))
                _native.check(err)
            c = ix.last_counters
            rec = float((ds.cpu().numpy() <= kth[:, None]).mean())
            print(f"search n={n} ef={ef}: {c['device_ms']:.2f} ms {10000 / c['device_ms'] * 1e3:.0f} QPS "
                  f"recall@10={rec:.4f} dist/q={c['computed_distances'] / 1e4:.0f}", flush=True)
        del ix
    del d_db
    torch.cuda.empty_cache()
