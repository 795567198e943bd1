#!/usr/bin/env python3
"""One-off, full-size parity check of BASELINE.json configs[1]: build the SAME 10 M rows with the unmodified
reference (oracle/_ref, all host threads) and with rad_b200 on the GPU, and compare recall@10 at ef 64 / 128 against
the exact answers.  Also the reference's full-size build rate on this box.  Writes gpurun_out/full_size_reference.json.

usage: python scripts/full_size_reference.py [ROWS] [QUERIES]"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from oracle import oracle as O  # noqa: E402
from rad_b200 import _native, synth  # noqa: E402
from rad_b200.index import Index  # noqa: E402
from rad_b200.sharded import ShardedExactSearch, unpack_topk_keys  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
K, M, EF_ADD = 10, 8, 400
L = _native.lib()
err = C.c_char_p()
centres = synth.centre_table(0, n)
d_c = torch.from_numpy(centres).cuda()
d_db = torch.empty((n, 128), dtype=torch.uint8, device="cuda")
stream = torch.cuda.current_stream().cuda_stream
L.rb_synth_fingerprints_device(d_db.data_ptr(), 0, n, 0, d_c.data_ptr(), len(centres), 1, stream, C.byref(err))
_native.check(err)
torch.cuda.synchronize()
pick = torch.arange(nq, device="cuda") * (n // nq)
d_q = d_db[pick].contiguous()
q = d_q.cpu().numpy()
truth = ShardedExactSearch(d_db, 0).search(d_q, K)
_, exact_d, _ = unpack_topk_keys(truth.cpu().numpy())
kth = exact_d[:, K - 1:K]


def recall(d):
    return float((np.asarray(d)[:, :K] <= kth).mean())


out = {"rows": n, "queries": nq, "connectivity": M, "expansion_add": EF_ADD}
gpu = Index(ndim=1024, connectivity=M, expansion_add=EF_ADD)
t = time.perf_counter()
L.rb_add_device(gpu._handle(), None, d_db.data_ptr(), n, C.byref(err))
_native.check(err)
out["gpu_build_s"] = time.perf_counter() - t
out["gpu_graph"] = {}
for ef in (64, 128):
    gpu.expansion_search = ef
    out["gpu_graph"][f"ef{ef}"] = recall(gpu.search(q, K).distances)
print(json.dumps(out), flush=True)

db = d_db.cpu().numpy()
del d_db
ui = O.ref_index_module()
threads = os.cpu_count() or 1
ref = ui.Index(ndim=1024, dtype="b1", metric="tanimoto", connectivity=M, expansion_add=EF_ADD, expansion_search=64)
t = time.perf_counter()
ref.add(np.arange(n, dtype=np.uint64), db, threads=threads)
out["reference_build_s"] = time.perf_counter() - t
out["reference_build_vectors_per_s"] = n / out["reference_build_s"]
out["reference_threads"] = threads
out["reference_graph"] = {}
for ef in (64, 128):
    ref.expansion_search = ef
    out["reference_graph"][f"ef{ef}"] = recall(ref.search(q, K, threads=threads).distances)
print(json.dumps(out), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/full_size_reference.json", "w"), indent=1)
