"""A/B on one box: ds_gather_rows (single-source specialisation of gather_bulk_kernel) against
ds_gather_rows_multi with ONE source (the general kernel: source table in shared memory), same
data, runs interleaved.  100 M x 90-byte rows; CUDA events; prints every sample.

    python profiles/tools/r02_gather_ab.py
"""
import ctypes
import sys

import torch

sys.path.insert(0, ".")
from ray_memory_management_b200 import _lib  # noqa: E402


def main(n=100_000_000, W=90, reps=7):
    g = torch.Generator(device="cuda").manual_seed(1)
    src = torch.randint(0, 256, (n * W,), dtype=torch.uint8, device="cuda", generator=g)
    idx = torch.randperm(n, device="cuda", generator=g).to(torch.int32)
    out = torch.empty(n * W, dtype=torch.uint8, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    srcs = (ctypes.c_void_p * 1)(src.data_ptr())
    rows = (ctypes.c_uint64 * 1)(n)

    def single():
        _lib.call("ds_gather_rows", src.data_ptr(), W, n, idx.data_ptr(), n, out.data_ptr(), s)

    def multi():
        _lib.call("ds_gather_rows_multi", srcs, rows, 1, W, idx.data_ptr(), n, out.data_ptr(), s)

    single()
    ref = out.clone()
    out.zero_()
    multi()
    print("equal:", bool(torch.equal(out, ref)), flush=True)
    t = {"single": [], "multi(1 source)": []}
    for _ in range(reps):
        for name, fn in (("single", single), ("multi(1 source)", multi)):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            t[name].append(a.elapsed_time(b))
    for name, v in t.items():
        v2 = sorted(v)
        print(f"{name}: median {v2[len(v2) // 2]:.3f} ms  min {v2[0]:.3f}  all {[round(x, 3) for x in v]}", flush=True)


if __name__ == "__main__":
    main()
