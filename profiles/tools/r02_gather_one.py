"""One gather launch per variant, for ncu: python profiles/tools/r02_gather_one.py [n]"""
import sys

import torch

sys.path.insert(0, ".")
from ray_memory_management_b200 import _lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 30_000_000
W = 90
g = torch.Generator(device="cuda").manual_seed(1)
src = torch.randint(0, 256, (n * W,), dtype=torch.uint8, device="cuda", generator=g)
idx = torch.randperm(n, device="cuda", generator=g).to(torch.int32)
out = torch.empty(n * W, dtype=torch.uint8, device="cuda")
s = torch.cuda.current_stream().cuda_stream
for mode, rows, stages in ((0, 0, 0), (2, 0, 0), (7, 64, 2)):  # auto (= cp.async variant), old granule kernel, bulk variant
    _lib.lib.ds_debug_set_gather_mode(mode)
    _lib.lib.ds_debug_set_gather_bulk(rows, stages, 0)
    for _ in range(2):
        _lib.call("ds_gather_rows", src.data_ptr(), W, n, idx.data_ptr(), n, out.data_ptr(), s)
    torch.cuda.synchronize()
print("ok")
