"""Kernel timeline of one bench step (torch.profiler / CUPTI): start, duration and the idle gap
before each kernel, to see where the step's time outside kernels goes."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench  # noqa: E402
import ray_memory_management_b200 as ds  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
torch.cuda.set_device(0)
block = bench.make_records(n, seed=1)


def step():
    res, _ = ds.sort_impl(ds.BlockList.from_blocks([block]), False, "key")
    return res.get_blocks()[0]


for _ in range(3):
    out = step()
    del out
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(2):
        out = step()
        del out
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
t0 = evs[0].time_range.start
prev_end = t0
total_gap = 0.0
for e in evs:
    gap = e.time_range.start - prev_end
    total_gap += max(gap, 0)
    print(f"{(e.time_range.start - t0) / 1e3:9.3f} ms  dur {e.time_range.elapsed_us() / 1e3:8.3f} ms  gap {gap:8.1f} us  {e.name[:70]}")
    prev_end = max(prev_end, e.time_range.end)
print(f"total span {(prev_end - t0) / 1e3:.3f} ms, idle {total_gap / 1e3:.3f} ms over 2 steps")
