#!/usr/bin/env python
"""BASELINE.json configs[2] and [3] across the GPUs of one box (torchrun, one rank per GPU).

  torchrun --nproc-per-node 8 profiles/run_c3_c4.py shuffle [rows_per_gpu]   # 125 M -> 100 GB at 8 GPUs
  torchrun --nproc-per-node N profiles/run_c3_c4.py groupby [total_rows]     # 2 G rows, 1 M keys (strong scaling)

CUDA events, max over ranks, 3 timed steps after 2 warm-ups; validation is not timed.
"""
import collections
import json
import os
import sys

import pyarrow as pa
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import make_records  # noqa: E402


def timed(fn, steps=3, warmup=2):
    for _ in range(warmup):
        r = fn()
        del r
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        r = fn()
        del r
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    import ray_memory_management_b200 as ds
    from ray_memory_management_b200 import distributed as D
    from ray_memory_management_b200 import graysort

    what = sys.argv[1]
    if what == "shuffle":
        n = int(sys.argv[2]) if len(sys.argv) > 2 else 125_000_000
        blk = make_records(n, 1 + rank)
        ms = timed(lambda: D.random_shuffle_block(blk, 42))
        a = D.random_shuffle_block(blk, 42)
        b = D.random_shuffle_block(blk, 42)
        c = D.random_shuffle_block(blk, 43)
        same = a.num_rows == b.num_rows and bool((a.columns["payload"].data == b.columns["payload"].data).all())
        differs = a.num_rows != c.num_rows or not bool((a.columns["key"].data == c.columns["key"].data).all())
        vin, vout = graysort.validate(blk), graysort.validate(a)
        ids = a.columns["payload"].data.view(-1, 90)[:, :8].contiguous().view(torch.int64).view(-1)
        inplace = float((ids[: min(n, a.num_rows)] == torch.arange(min(n, a.num_rows), device="cuda")).double().mean().item())
        tot = torch.tensor([vin["checksum"] % (1 << 62), vout["checksum"] % (1 << 62), vin["records"], vout["records"],
                            int(same), int(differs)], dtype=torch.int64, device="cuda")
        dist.all_reduce(tot)
        if rank == 0:
            t = tot.cpu().tolist()
            print(json.dumps({"config": f"random_shuffle(seed=42), {world} x {n} x 100-byte records ({world * n * 100 / 1e9:.0f} GB)",
                              "ms": ms, "GB_per_s": world * n * 100 / ms / 1e6,
                              "checksum_preserved": t[0] % (1 << 62) == t[1] % (1 << 62), "records_in": t[2],
                              "records_out": t[3], "deterministic_ranks": t[4], "differs_for_other_seed_ranks": t[5],
                              "fraction_of_rows_left_in_place_rank0": inplace}), flush=True)
    else:
        total = int(sys.argv[2]) if len(sys.argv) > 2 else 2_000_000_000
        n = total // world
        g = torch.Generator(device="cuda").manual_seed(2 + rank)
        k = torch.randint(0, 1_000_000, (n,), dtype=torch.int64, device="cuda", generator=g)
        v = torch.rand(n, dtype=torch.float64, device="cuda", generator=g)
        blk = ds.DeviceArrowBlock(collections.OrderedDict(
            k=ds.DeviceColumn(pa.int64(), k.view(torch.uint8)), v=ds.DeviceColumn(pa.float64(), v.view(torch.uint8))), n)
        aggs = (ds.Sum("v"), ds.Mean("v"), ds.Count())
        ms = timed(lambda: D.groupby_block(blk, "k", aggs))
        out = D.groupby_block(blk, "k", aggs)
        keys = out.columns["k"].typed()
        ok_sorted = bool((keys[1:] > keys[:-1]).all()) if out.num_rows > 1 else True
        stats = torch.tensor([float(out.columns["count()"].typed().sum().item()), float(out.columns["sum(v)"].typed().sum().item()),
                              float(v.sum().item()), float(out.num_rows), float(ok_sorted)], dtype=torch.float64, device="cuda")
        dist.all_reduce(stats)
        edge = torch.stack([keys[0], keys[-1]]) if out.num_rows else torch.zeros(2, dtype=torch.int64, device="cuda")
        edges = torch.empty(world * 2, dtype=torch.int64, device="cuda")
        dist.all_gather_into_tensor(edges, edge)
        e = edges.cpu().tolist()
        disjoint = all(e[2 * r + 1] < e[2 * r + 2] for r in range(world - 1))
        mean_ok = bool(torch.allclose(out.columns["mean(v)"].typed(), out.columns["sum(v)"].typed() / out.columns["count()"].typed().double(), rtol=1e-12))
        if rank == 0:
            s = stats.cpu().tolist()
            print(json.dumps({"config": f"groupby(int64 key, 1 M distinct).aggregate(Sum, Mean, Count), {total} rows of float64 on {world} GPU(s)",
                              "ms": ms, "rows_per_s": total / ms * 1e3, "GB_per_s": total * 16 / ms / 1e6,
                              "count_total": s[0], "sum_rel_err": abs(s[1] - s[2]) / s[2], "groups": s[3],
                              "ranks_sorted": s[4], "key_ranges_disjoint": disjoint, "mean_is_sum_over_count": mean_ok}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
