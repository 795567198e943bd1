#!/usr/bin/env python3
"""A/B timing of two BUILDS of librad_b200.so on one GPU box (development aid, not the bench).

    python scripts/ab_build.py ROWS libA.so libB.so [...]        # each library in its own process, twice, interleaved

Prints per library: device ms of the build, the phase times (insert_search / tile / select / sort / reverse), the work
counters, a checksum of the adjacency (equal checksums = the same graph, bit for bit), and ef-search QPS at ef 64."""
import ctypes as C
import hashlib
import json
import os
import subprocess
import sys


def child(rows: int):
    import numpy as np
    import torch
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from rad_b200 import _native, synth
    from rad_b200.index import Index
    L = _native.lib()
    err = C.c_char_p()
    centres = synth.centre_table(0, rows)
    d_c = torch.from_numpy(centres).cuda()
    d_db = torch.empty((rows, 128), dtype=torch.uint8, device="cuda")
    L.rb_synth_fingerprints_device(d_db.data_ptr(), 0, rows, 0, d_c.data_ptr(), len(centres), 1,
                                   torch.cuda.current_stream().cuda_stream, C.byref(err))
    _native.check(err)
    torch.cuda.synchronize()
    out = {}
    for rep in range(2):
        ix = Index(ndim=1024, connectivity=8, expansion_add=400, expansion_search=64, lazy_host_sync=True)
        L.rb_add_device(ix._handle(), None, d_db.data_ptr(), rows, C.byref(err))
        _native.check(err)
        c = ix.last_counters
        out[f"build{rep}"] = {k: round(c[k], 1) for k in ("device_ms", "insert_ms", "tile_ms", "select_ms", "sort_ms",
                                                          "reverse_ms")}
        out["distances_per_insert"] = c["computed_distances"] / rows
        if rep == 0:
            del ix
    ix.sync_host()
    a0, au = ix._debug_adjacency()
    out["graph_sha"] = hashlib.sha1(np.ascontiguousarray(a0).tobytes() + np.ascontiguousarray(au).tobytes()).hexdigest()[:16]
    nq = 100000
    d_q = d_db[torch.randperm(rows, device="cuda", generator=torch.Generator("cuda").manual_seed(1))[:nq]].contiguous()
    sl = torch.empty((nq, 10), dtype=torch.int32, device="cuda")
    ds = torch.empty((nq, 10), dtype=torch.float32, device="cuda")
    ct = torch.empty((nq,), dtype=torch.int32, device="cuda")
    for ef in (64, 128):
        ix.expansion_search = ef
        best = 1e9
        for _ in range(3):
            L.rb_search_device(ix._handle(), d_q.data_ptr(), nq, 10, 0, sl.data_ptr(), ds.data_ptr(), ct.data_ptr(),
                               C.byref(err))
            _native.check(err)
            best = min(best, ix.last_counters["device_ms"])
        out[f"search_ef{ef}_ms"] = round(best, 3)
    out["search_sha"] = hashlib.sha1(sl.cpu().numpy().tobytes()).hexdigest()[:16]
    print("AB " + json.dumps(out), flush=True)


if __name__ == "__main__":
    if sys.argv[1] == "--child":
        child(int(sys.argv[2]))
        sys.exit(0)
    rows, libs = int(sys.argv[1]), sys.argv[2:]
    for rnd in range(2):
        for lib in libs:
            env = dict(os.environ, RAD_B200_LIB=os.path.abspath(lib))
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", str(rows)], env=env,
                               capture_output=True, text=True)
            line = [ln for ln in r.stdout.splitlines() if ln.startswith("AB ")]
            print(f"round {rnd} {os.path.basename(lib)}: {line[0][3:] if line else 'FAILED ' + r.stderr[-800:]}", flush=True)
