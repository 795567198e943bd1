#!/usr/bin/env python3
"""The case ncu captures for search_kernel (profiles/r2_ncu_search_summary.txt): the bench's 10 M-row graph, 10 k queries
drawn from the set, ef 64 and 128, two launches each.   python scripts/search_case.py [rows]"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rad_b200 import _native, synth  # noqa: E402
from rad_b200.index import Index  # noqa: E402

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
L, err = _native.lib(), C.c_char_p()
centres = synth.centre_table(0, rows)
d_c = torch.from_numpy(centres).cuda()
d_db = torch.empty((rows, 128), dtype=torch.uint8, device="cuda")
L.rb_synth_fingerprints_device(d_db.data_ptr(), 0, rows, 0, d_c.data_ptr(), len(centres), 1,
                               torch.cuda.current_stream().cuda_stream, C.byref(err))
_native.check(err)
torch.cuda.synchronize()
ix = Index(ndim=1024, connectivity=8, expansion_add=400, expansion_search=64, lazy_host_sync=True)
L.rb_add_device(ix._handle(), None, d_db.data_ptr(), rows, C.byref(err))
_native.check(err)
nq = 10000
g = torch.Generator(device="cuda")
g.manual_seed(1234)
d_q = d_db[torch.randperm(rows, device="cuda", generator=g)[:nq]].contiguous()
sl = torch.empty((nq, 10), dtype=torch.int32, device="cuda")
ds = torch.empty((nq, 10), dtype=torch.float32, device="cuda")
ct = torch.empty((nq,), dtype=torch.int32, device="cuda")
for ef in (64, 128):
    ix.expansion_search = ef
    for _ in range(2):
        L.rb_search_device(ix._handle(), d_q.data_ptr(), nq, 10, 0, sl.data_ptr(), ds.data_ptr(), ct.data_ptr(), C.byref(err))
        _native.check(err)
    c = ix.last_counters
    print(f"ef={ef}: {c['device_ms']:.3f} ms, {c['computed_distances'] / nq:.1f} distances and "
          f"{c['visited_members'] / nq:.1f} expansions per query")
