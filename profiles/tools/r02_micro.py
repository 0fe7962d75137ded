"""Micro-benchmarks of the round-2 kernels on one B200 (CUDA events, inputs >> L2).

    python profiles/tools/r02_micro.py gather   # ds_gather_rows, 100 M x 90-byte rows, every mode / shape
    python profiles/tools/r02_micro.py split    # ds_split_plan + ds_split_rows, all destinations local

Prints one line per variant: milliseconds (median of 5 after 2 warm-ups), algorithmic GB/s.
"""
import ctypes
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


def bench_gather(n=100_000_000, W=90):
    g = torch.Generator(device="cuda").manual_seed(1)
    src = torch.randint(0, 256, (n * W,), dtype=torch.uint8, device="cuda", generator=g)
    idx = torch.randperm(n, device="cuda", generator=g).to(torch.int32)
    out = torch.empty(n * W, dtype=torch.uint8, device="cuda")
    ref = None
    s = torch.cuda.current_stream().cuda_stream
    alg = n * (4 + 2 * W)

    def run():
        _lib.call("ds_gather_rows", src.data_ptr(), W, n, idx.data_ptr(), n, out.data_ptr(), s)

    variants = [(2, 0, 0, 0, 0)]
    for mode in (8, 7):
        for rows, stages in ((64, 2), (64, 3), (64, 4), (96, 2), (96, 3), (128, 2), (128, 3), (128, 4), (192, 2),
                             (256, 2), (32, 4)):
            variants.append((mode, rows, stages, 0, 0))
    variants += [(7, 128, 2, 0, 64), (7, 128, 2, 0, 32), (2, 0, 0, 0, 32)]
    for mode, rows, stages, ctas, l2 in variants:
        if l2:
            print("  (cudaLimitMaxL2FetchGranularity", l2, "was", lib.ds_debug_set_l2_fetch_granularity(l2), ")")
        lib.ds_debug_set_gather_mode(mode)
        lib.ds_debug_set_gather_bulk(rows, stages, ctas)
        out.zero_()
        try:
            ms = timed(run)
        except Exception as e:  # noqa: BLE001
            print(f"gather mode={mode} rows={rows} stages={stages} ctas={ctas}: FAILED {e}", flush=True)
            continue
        if ref is None:
            ref = out.clone()
            ok = True
        else:
            ok = bool(torch.equal(out, ref))
        print(f"gather mode={mode} rows={rows} stages={stages} ctas={ctas}: {ms:.3f} ms  "
              f"{alg / ms / 1e6:.0f} GB/s algorithmic  equal={ok}", flush=True)
    lib.ds_debug_set_gather_mode(0)
    lib.ds_debug_set_gather_bulk(0, 0, 0)


def bench_split(n=100_000_000, nparts=8):
    g = torch.Generator(device="cuda").manual_seed(2)
    keys = torch.randint(0, 256, (n * 10,), dtype=torch.uint8, device="cuda", generator=g)
    payload = torch.randint(0, 256, (n * 90,), dtype=torch.uint8, device="cuda", generator=g)
    bnd = torch.zeros((nparts - 1, 10), dtype=torch.uint8, device="cuda")
    for i in range(nparts - 1):
        bnd[i, 0] = (i + 1) * 256 // nparts
    ntiles = (n + 255) // 256
    dest = torch.empty(n, dtype=torch.uint8, device="cuda")
    offs = torch.empty(int(lib.ds_split_table_words(n, nparts)), dtype=torch.int32, device="cuda")
    counts = torch.empty(nparts, dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream

    def plan():
        _lib.call("ds_split_plan", _lib.DS_KEY_FSB, keys.data_ptr(), 10, None, n, bnd.data_ptr(), nparts - 1, 0,
                  dest.data_ptr(), offs.data_ptr(), counts.data_ptr(), s)

    ms = timed(plan)
    print(f"split plan (10-byte keys, {nparts} destinations): {ms:.3f} ms  {n * 11 / ms / 1e6:.0f} GB/s", flush=True)
    c = counts.cpu().tolist()
    print("counts", c, flush=True)
    for name, col, W in (("key", keys, 10), ("payload", payload, 90)):
        outs = [torch.empty(int(x) * W, dtype=torch.uint8, device="cuda") for x in c]
        arr = (ctypes.c_void_p * nparts)(*[o.data_ptr() for o in outs])
        for per_sm in (1, 2, 3):
            lib.ds_debug_set_split_ctas(per_sm)
            for max_ctas in (0, 74, 37):
                if max_ctas and per_sm > 1:
                    continue

                def rows():
                    _lib.call("ds_split_rows", col.data_ptr(), W, n, dest.data_ptr(), offs.data_ptr(), nparts, arr,
                              0, max_ctas, s)

                ms = timed(rows)
                print(f"split rows {name} W={W} ctas/SM={per_sm} max_ctas={max_ctas}: {ms:.3f} ms  "
                      f"{n * (1 + 2 * W) / ms / 1e6:.0f} GB/s", flush=True)
        # check against a torch reference: rows of destination d in input order
        d0 = (dest == 3).nonzero().squeeze(1)[:1000]
        want = col.view(n, W)[d0]
        assert torch.equal(outs[3].view(-1, W)[:1000], want), name
    lib.ds_debug_set_split_ctas(0)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what in ("gather", "all"):
        bench_gather()
    if what in ("split", "all"):
        bench_split()
