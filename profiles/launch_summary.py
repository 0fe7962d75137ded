#!/usr/bin/env python
"""ncu launch list (--metrics gpu__time_duration.sum --csv) -> markdown table of kernel shares.
    python profiles/launch_summary.py gpurun_out/launches.csv "title" > profiles/rNN_launches_summary.md
"""
import collections
import csv
import sys

path, title = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else sys.argv[1])
rows = [r for r in csv.reader(open(path)) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot = collections.defaultdict(float)
cnt = collections.Counter()
for r in rows[1:]:
    v = float(r[vi].replace(",", ""))
    v = {"ns": v * 1e-6, "us": v * 1e-3, "usecond": v * 1e-3, "ms": v, "msecond": v, "nsecond": v * 1e-6}.get(r[ui], v * 1e-6)
    tot[r[ki]] += v
    cnt[r[ki]] += 1
total = sum(tot.values())
print(f"# {title}\n")
print("`ncu --metrics gpu__time_duration.sum --clock-control none`; per-launch times are cold-cache and")
print("serialised: compare SHARES with bench.py's CUDA-event figures, not absolute times.")
print("`at::` kernels are torch's input generation and the untimed sanity check, not the sort.\n")
print("| share | launches | avg ms | kernel |\n|---|---|---|---|")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    print(f"| {100 * v / total:.1f}% | {cnt[k]} | {v / cnt[k]:.3f} | `{k[:110]}` |")
ours = {k: v for k, v in tot.items() if not any(s in k for s in ("at::", "nccl", "cub::", "thrust::"))}
so = sum(ours.values())
print(f"\nOur kernels only ({so:.2f} ms over the captured launches):\n")
print("| share of ours | launches | avg ms | kernel |\n|---|---|---|---|")
for k, v in sorted(ours.items(), key=lambda kv: -kv[1]):
    print(f"| {100 * v / so:.1f}% | {cnt[k]} | {v / cnt[k]:.3f} | `{k[:110]}` |")
