import ctypes as C, os, sys
sys.path.insert(0, ".")
import torch
from rad_b200 import _native, synth
from rad_b200.index import Index
L = _native.lib(); err = C.c_char_p()
rows = int(sys.argv[1])
centres = synth.centre_table(0, rows); d_c = torch.from_numpy(centres).cuda()
d_db = torch.empty((rows, 128), dtype=torch.uint8, device="cuda")
L.rb_synth_fingerprints_device(d_db.data_ptr(), 0, rows, 0, d_c.data_ptr(), len(centres), 1, torch.cuda.current_stream().cuda_stream, C.byref(err))
torch.cuda.synchronize()
ix = Index(ndim=1024, connectivity=8, expansion_add=400, expansion_search=64, lazy_host_sync=True)
L.rb_add_device(ix._handle(), None, d_db.data_ptr(), rows, C.byref(err)); _native.check(err)
print(ix.last_counters)
