#!/usr/bin/env python
"""Times the other BASELINE.json configs on one GPU (not bench lines; recorded in profiles/).

  python profiles/run_configs.py shuffle [rows]      # Dataset.random_shuffle of 100-byte records
  python profiles/run_configs.py groupby [rows] [distinct] [blocks]
  python profiles/run_configs.py sort_i64 [rows]     # configs[0] shape: {value:int64, key:int64}
"""
import collections
import json
import os
import sys
import time

import pyarrow as pa
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ray_memory_management_b200 as ds  # noqa: E402
from bench import make_records  # noqa: E402


def timed(fn, steps=3, warmup=2):
    for _ in range(warmup):
        r = fn()
        del r
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        r = fn()
        del r
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def col(t, arrow_type):
    return ds.DeviceColumn(arrow_type, t.view(torch.uint8).reshape(-1))


def main():
    what = sys.argv[1]
    if what == "shuffle":
        n = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000_000
        blk = make_records(n, 1)
        d = ds.Dataset.from_blocks([blk])
        ms = timed(lambda: d.random_shuffle(seed=42))
        out = d.random_shuffle(seed=42).get_blocks()[0]
        ids = out.columns["payload"].data.view(n, 90)[:, :8].contiguous().view(torch.int64).view(n)
        assert int(ids.sum().item()) == n * (n - 1) // 2
        assert not bool((ids[:1000] == torch.arange(1000, device="cuda")).all())
        print(json.dumps({"config": "random_shuffle(seed) of 100-byte records, 1 GPU, 1 block", "rows": n,
                          "ms": ms, "GB_per_s": n * 100 / ms / 1e6, "rows_per_s": n / ms * 1e3}))
    elif what == "groupby":
        n = int(sys.argv[2]) if len(sys.argv) > 2 else 250_000_000
        distinct = int(sys.argv[3]) if len(sys.argv) > 3 else 1_000_000
        nblocks = int(sys.argv[4]) if len(sys.argv) > 4 else 1
        g = torch.Generator(device="cuda").manual_seed(2)
        blocks = []
        per = n // nblocks
        for _ in range(nblocks):
            k = torch.randint(0, distinct, (per,), dtype=torch.int64, device="cuda", generator=g)
            v = torch.rand(per, dtype=torch.float64, device="cuda", generator=g)
            blocks.append(ds.DeviceArrowBlock(collections.OrderedDict(k=col(k, pa.int64()), v=col(v, pa.float64())), per))
        d = ds.Dataset.from_blocks(blocks)
        ms = timed(lambda: d.groupby("k").aggregate(ds.Sum("v"), ds.Mean("v"), ds.Count()))
        out = d.groupby("k").aggregate(ds.Sum("v"), ds.Mean("v"), ds.Count())
        cnt = sum(int(b.columns["count()"].typed().sum().item()) for b in out.get_blocks() if b.num_rows)
        assert cnt == per * nblocks
        print(json.dumps({"config": "groupby(int64 key).aggregate(Sum, Mean, Count) of float64, 1 GPU",
                          "rows": per * nblocks, "distinct": distinct, "blocks": nblocks, "ms": ms,
                          "rows_per_s": per * nblocks / ms * 1e3, "GB_per_s": per * nblocks * 16 / ms / 1e6,
                          "groups_out": out.count()}))
    elif what == "sort_i64":
        n = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
        nblocks = int(sys.argv[3]) if len(sys.argv) > 3 else 1
        g = torch.Generator(device="cuda").manual_seed(0)
        key = torch.randint(0, 2**63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)
        val = torch.arange(n, dtype=torch.int64, device="cuda")
        blk = ds.DeviceArrowBlock(collections.OrderedDict(value=col(val, pa.int64()), key=col(key, pa.int64())), n)
        per = n // nblocks
        d = ds.Dataset.from_blocks([blk.slice(i * per, (i + 1) * per if i < nblocks - 1 else n) for i in range(nblocks)])
        ms = timed(lambda: d.sort("key"))
        outb = [b for b in d.sort("key").get_blocks() if b.num_rows]
        ks = torch.cat([b.columns["key"].typed() for b in outb])
        assert ks.numel() == n and bool((ks[1:] >= ks[:-1]).all())
        print(json.dumps({"config": "range_table(n) + random int64 key, Dataset.sort(key), 1 GPU", "rows": n,
                          "blocks": nblocks,
                          "ms": ms, "rows_per_s": n / ms * 1e3, "GB_per_s": n * 16 / ms / 1e6}))


if __name__ == "__main__":
    main()
