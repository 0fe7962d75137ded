"""One split plan + one key split + one payload split, for ncu: python profiles/tools/r02_split_one.py [n] [ctas_per_sm]"""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from ray_memory_management_b200 import _lib  # noqa: E402

lib = _lib.lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 30_000_000
per_sm = int(sys.argv[2]) if len(sys.argv) > 2 else 1
nparts = 8
g = torch.Generator(device="cuda").manual_seed(2)
keys = torch.randint(0, 256, (n * 10,), dtype=torch.uint8, device="cuda", generator=g)
payload = torch.randint(0, 256, (n * 90,), dtype=torch.uint8, device="cuda", generator=g)
bnd = torch.zeros((nparts - 1, 10), dtype=torch.uint8, device="cuda")
for i in range(nparts - 1):
    bnd[i, 0] = (i + 1) * 256 // nparts
dest = torch.empty(n, dtype=torch.uint8, device="cuda")
offs = torch.empty(int(lib.ds_split_table_words(n, nparts)), dtype=torch.int32, device="cuda")
counts = torch.empty(nparts, dtype=torch.int64, device="cuda")
s = torch.cuda.current_stream().cuda_stream
lib.ds_debug_set_split_ctas(per_sm)
for _ in range(2):
    _lib.call("ds_split_plan", _lib.DS_KEY_FSB, keys.data_ptr(), 10, None, n, bnd.data_ptr(), nparts - 1, 0,
              dest.data_ptr(), offs.data_ptr(), counts.data_ptr(), s)
    c = counts.cpu().tolist()
    for col, W in ((keys, 10), (payload, 90)):
        outs = [torch.empty(int(x) * W, dtype=torch.uint8, device="cuda") for x in c]
        arr = (ctypes.c_void_p * nparts)(*[o.data_ptr() for o in outs])
        _lib.call("ds_split_rows", col.data_ptr(), W, n, dest.data_ptr(), offs.data_ptr(), nparts, arr, 0, 0, s)
        torch.cuda.synchronize()
print("ok")
