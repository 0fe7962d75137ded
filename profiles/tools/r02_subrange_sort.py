"""Key sort of one rank's share of a range-partitioned dataset (10-byte keys whose leading bytes
cover 1/parts of the key space), with and without the hybrid plan on a sub-range (KeyXform).

    python profiles/tools/r02_subrange_sort.py [rows] [parts]

Prints milliseconds per ds_sort_indices_fsb call (median of 5 after 2 warm-ups, CUDA events)
and the plan the call ended with (ds_debug_last_sort_plan).
"""
import sys

import torch

sys.path.insert(0, ".")
from ray_memory_management_b200 import _lib  # noqa: E402

lib = _lib.lib


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
    parts = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    g = torch.Generator(device="cuda").manual_seed(3)
    s = torch.cuda.current_stream().cuda_stream
    idx = torch.empty(n, dtype=torch.int32, device="cuda")
    out = torch.empty(n * 10, dtype=torch.uint8, device="cuda")
    ws_holder = {}

    def alloc(nb):
        if ws_holder.get("n", -1) < nb:
            ws_holder["t"] = torch.empty(nb, dtype=torch.uint8, device="cuda")
            ws_holder["n"] = nb
        return ws_holder["t"]

    for part in (None, 3 % parts, parts - 1):
        keys = torch.randint(0, 256, (n, 10), dtype=torch.uint8, device="cuda", generator=g)
        if part is not None:
            lo, hi = part * 65536 // parts, (part + 1) * 65536 // parts
            top = torch.randint(lo, hi, (n,), device="cuda", generator=g)
            keys[:, 0] = (top >> 8).to(torch.uint8)
            keys[:, 1] = (top & 255).to(torch.uint8)
            del top
        keys = keys.reshape(-1)
        for sub in (1, 0):
            lib.ds_debug_set_subrange(sub)

            def run():
                _lib.call_ws("ds_sort_indices_fsb", alloc, keys.data_ptr(), 10, n, 0, idx.data_ptr(),
                             out.data_ptr(), 0, stream=s)

            ms = timed(run)
            what = "whole key space" if part is None else f"share {part} of {parts}"
            print(f"rows={n} keys={what} subrange={'on' if sub else 'off'}: {ms:.3f} ms  "
                  f"plan={lib.ds_debug_last_sort_plan()}", flush=True)
        lib.ds_debug_set_subrange(1)
        # sortedness of the key column that came out (10-byte rows compared as big-endian)
        v = out.view(n, 10)
        hi8 = torch.zeros(n, dtype=torch.int64, device="cuda")
        for b in range(8):
            hi8 = (hi8 << 8) | v[:, b].to(torch.int64)
        hi8 ^= -(2 ** 63)  # unsigned order
        lo2 = (v[:, 8].to(torch.int32) << 8) | v[:, 9].to(torch.int32)
        ok = bool(((hi8[1:] > hi8[:-1]) | ((hi8[1:] == hi8[:-1]) & (lo2[1:] >= lo2[:-1]))).all())
        perm_ok = bool(torch.equal(torch.sort(idx.to(torch.int64))[0], torch.arange(n, device="cuda")))
        print(f"  sorted={ok} permutation={perm_ok}", flush=True)
        del hi8, lo2, keys


if __name__ == "__main__":
    main()
