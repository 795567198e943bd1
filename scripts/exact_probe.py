#!/usr/bin/env python3
"""Time the exact top-10 kernel(s) on device-generated rows: python scripts/exact_probe.py ROWS QUERIES [REPS]."""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from rad_b200 import _native, synth  # noqa: E402

n, nq = int(sys.argv[1]), int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
L = _native.lib()
err = C.c_char_p()
centres = synth.centre_table(0, n)
d_c = torch.from_numpy(centres).cuda()
d_db = torch.empty((n, 128), dtype=torch.uint8, device="cuda")
s = torch.cuda.current_stream().cuda_stream
L.rb_synth_fingerprints_device(d_db.data_ptr(), 0, n, 0, d_c.data_ptr(), len(centres), 1, s, C.byref(err))
d_q = d_db[torch.arange(nq, device="cuda") * (n // nq)].contiguous()
ws = L.rb_exact_topk_workspace_bytes(n, nq, 10)
d_ws = torch.empty(ws, dtype=torch.uint8, device="cuda")
out = torch.empty((nq, 10), dtype=torch.int64, device="cuda")
for _ in range(reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    L.rb_exact_topk_device(d_db.data_ptr(), n, 0, d_q.data_ptr(), nq, 10, out.data_ptr(), d_ws.data_ptr(), ws, s,
                           C.byref(err))
    _native.check(err)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"n={n} q={nq}: {ms:.2f} ms {n * nq / ms / 1e6:.1f} G pairs/s", flush=True)
