#!/usr/bin/env python3
"""Quick GPU check of the tcgen05 exact kernel against the POPC kernel (same C ABI, RB_EXACT_KERNEL switches)."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from rad_b200 import _native, synth  # noqa: E402

mode = os.environ.get("RB_EXACT_KERNEL", "auto")
L = _native.lib()
err = C.c_char_p()


def topk(d_db, d_q, k, slot_base=0):
    n, nq = d_db.shape[0], d_q.shape[0]
    ws = L.rb_exact_topk_workspace_bytes(n, nq, k)
    d_ws = torch.empty(ws, dtype=torch.uint8, device="cuda")
    out = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    L.rb_exact_topk_device(d_db.data_ptr(), n, slot_base, d_q.data_ptr(), nq, k, out.data_ptr(), d_ws.data_ptr(), ws,
                           torch.cuda.current_stream().cuda_stream, C.byref(err))
    _native.check(err)
    e1.record()
    torch.cuda.synchronize()
    return out, e0.elapsed_time(e1)


if len(sys.argv) > 1 and sys.argv[1] == "child":
    cases = [(300, 64, 10), (5000, 130, 10), (20000, 200, 10), (4097, 129, 100), (100000, 1000, 10)]
    res = {}
    for n, nq, k in cases:
        db = synth.clustered_fingerprints(n, seed=n + 3)
        q = synth.heldout_queries(nq, seed=n + 3, n_total=n)
        q[0] = 0
        out, ms = topk(torch.from_numpy(db).cuda(), torch.from_numpy(q).cuda(), k, slot_base=7)
        res[f"{n}_{nq}_{k}"] = out.cpu().numpy()
        print(mode, n, nq, k, f"{ms:.3f} ms", flush=True)
    np.savez(f"gpurun_out/i8_check_{mode}.npz", **res)
    for n, nq in ((1_000_000, 10_000), (10_000_000, 10_000)):
        centres = synth.centre_table(0, n)
        d_c = torch.from_numpy(centres).cuda()
        d_db = torch.empty((n, 128), dtype=torch.uint8, device="cuda")
        L.rb_synth_fingerprints_device(d_db.data_ptr(), 0, n, 0, d_c.data_ptr(), len(centres), 1,
                                       torch.cuda.current_stream().cuda_stream, C.byref(err))
        d_q = d_db[torch.arange(nq, device="cuda") * 97].contiguous()
        for _ in range(3):
            out, ms = topk(d_db, d_q, 10)
        print(mode, f"n={n} q={nq}: {ms:.2f} ms {n * nq / ms / 1e6:.1f} G pairs/s", flush=True)
        np.save(f"gpurun_out/i8_big_{mode}_{n}.npy", out.cpu().numpy())
else:
    os.makedirs("gpurun_out", exist_ok=True)
    for m in ("popc", "i8"):
        env = dict(os.environ, RB_EXACT_KERNEL=m)
        r = subprocess.run([sys.executable, __file__, "child"], env=env, timeout=240)
        print(m, "exit", r.returncode, flush=True)
    a, b = np.load("gpurun_out/i8_check_popc.npz"), np.load("gpurun_out/i8_check_i8.npz")
    for key in a.files:
        same = np.array_equal(a[key], b[key])
        print(key, "identical" if same else f"DIFFERENT ({(a[key] != b[key]).mean():.4f} of entries)")
    for n in (1_000_000, 10_000_000):
        x, y = np.load(f"gpurun_out/i8_big_popc_{n}.npy"), np.load(f"gpurun_out/i8_big_i8_{n}.npy")
        print(n, "identical" if np.array_equal(x, y) else "DIFFERENT")
