#!/usr/bin/env python
"""Summarise an .ncu-rep: key raw metrics, stall mix, and stall samples by code region.
    python profiles/ncu_summary.py gpurun_out/prof.ncu-rep [chunk]
"""
import collections, csv, subprocess, sys

rep = sys.argv[1]
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 50
skip = sys.argv[3] if len(sys.argv) > 3 else "0"   # which launch of the report
sel = ["--launch-skip", skip, "--launch-count", "1"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"] + sel, capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts.sum",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.max", "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"]
for r in rows[2:3]:
    d = dict(zip(hdr, r))
    print(d["Kernel Name"][:100])
    for k in keys:
        if k in d:
            print(f"  {k} = {d[k]} {units[hdr.index(k)]}")
    st = [(float(d[h]), h.split("stalled_")[1].split("_per")[0]) for h in hdr if "average_warps_issue_stalled" in h and d[h]]
    print("  stalls/issue:", ", ".join(f"{n}={v:.2f}" for v, n in sorted(st, reverse=True)[:8]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + sel, capture_output=True, text=True).stdout
rows = [r for r in csv.reader(src.splitlines()) if len(r) > 5 and r[0].startswith("0x")]
# first kernel only: addresses restart when the listing repeats
first = []
seen = set()
for r in rows:
    if r[0] in seen:
        break
    seen.add(r[0]); first.append(r)
tot = sum(int(r[2]) for r in first) or 1
print(f"  source: {len(first)} SASS instrs, {tot} samples")
for s in range(0, len(first), chunk):
    c = first[s:s + chunk]
    share = sum(int(r[2]) for r in c) / tot * 100
    ops = collections.Counter(r[1].split()[0].split(".")[0] if not r[1].strip().startswith("@") else r[1].split()[1].split(".")[0] for r in c)
    top = max(c, key=lambda r: int(r[2]))
    print(f"  {s:5d} {share:5.1f}%  {dict(ops.most_common(4))}  hot: {top[1].strip()[:50]} ({int(top[2])/tot*100:.1f}%)")
