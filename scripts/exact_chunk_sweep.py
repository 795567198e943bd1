#!/usr/bin/env python3
"""Sweep the database chunk count of the exact top-k launch (RB_EXACT_CHUNKS):
python scripts/exact_chunk_sweep.py ROWS QUERIES auto c1 c2 ..."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, ".")
from rad_b200 import _native, synth  # noqa: E402

n, nq = int(sys.argv[1]), int(sys.argv[2])
L = _native.lib()
err = C.c_char_p()
centres = synth.centre_table(0, n)
d_c = torch.from_numpy(centres).cuda()
d_db = torch.empty((n, 128), dtype=torch.uint8, device="cuda")
s = torch.cuda.current_stream().cuda_stream
L.rb_synth_fingerprints_device(d_db.data_ptr(), 0, n, 0, d_c.data_ptr(), len(centres), 1, s, C.byref(err))
d_q = d_db[torch.arange(nq, device="cuda") * (n // nq)].contiguous()
out = torch.empty((nq, 10), dtype=torch.int64, device="cuda")
for c in sys.argv[3:]:
    if c != "auto":
        os.environ["RB_EXACT_CHUNKS"] = c
    else:
        os.environ.pop("RB_EXACT_CHUNKS", None)
    ws = L.rb_exact_topk_workspace_bytes(n, nq, 10)
    d_ws = torch.empty(ws, dtype=torch.uint8, device="cuda")
    best = 1e9
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        L.rb_exact_topk_device(d_db.data_ptr(), n, 0, d_q.data_ptr(), nq, 10, out.data_ptr(), d_ws.data_ptr(), ws, s,
                               C.byref(err))
        _native.check(err)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print(f"chunks={c}: {best:.2f} ms {n * nq / best / 1e6:.1f} G pairs/s", flush=True)
