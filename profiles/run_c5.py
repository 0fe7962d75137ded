#!/usr/bin/env python
"""BASELINE.json configs[4]: GraySort-format sort with ~62.5 GB per GPU, HBM resident.

    torchrun --nproc-per-node 8 profiles/run_c5.py [rows_per_gpu]

One timed distributed sort with ``consume=True`` (the input is released as the reference
releases its input refs with clear_input_blocks), after a small warm-up that sets up NCCL and
reserves the symmetric receive buffer.  The default multi-GPU plan (streaming split straight into
the peers' receive buffers, DS_EXCHANGE=split) has no send buffer, so input (62.5 GB) + receive
buffer (65 GB) fit; the copy-engine plan (DS_EXCHANGE=p2p) does not at this size.
The single timed run includes first-use cudaMalloc of its temporaries (no same-size warm-up
fits in HBM next to a second input).  Validation (not timed): every rank's output is sorted,
rank boundaries are ordered, and the order-independent CRC-32 checksum and the record count of
the whole dataset are unchanged.  Rank 0 prints one JSON line.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import make_records  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from ray_memory_management_b200 import distributed as D
    from ray_memory_management_b200 import graysort

    n = int(sys.argv[1]) if len(sys.argv) > 1 else 625_000_000
    for _ in range(2):  # warm-up: NCCL, symmetric memory, kernels
        w = make_records(1_000_000, 100 + rank)
        D.sort_block(w, "key", consume=True)
    del w
    D.reserve_peer_buffer(int(n * 1.04) * 100 + (1 << 20))
    blk = make_records(n, 1 + rank)
    before = graysort.validate(blk)
    torch.cuda.synchronize()
    dist.barrier()
    free0 = torch.cuda.mem_get_info()[0]
    torch.cuda.reset_peak_memory_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = D.sort_block(blk, "key", consume=True)
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    after = graysort.validate(out)
    # global invariants
    tot = torch.tensor([before["checksum"] % (1 << 62), after["checksum"] % (1 << 62), before["records"],
                        after["records"], after["unordered"]], dtype=torch.int64, device="cuda")
    dist.all_reduce(tot)
    edge = out.columns["key"].data.view(-1, 10)[[0, -1]].contiguous() if out.num_rows else torch.zeros(2, 10, dtype=torch.uint8, device="cuda")
    edges = torch.empty(world * 20, dtype=torch.uint8, device="cuda")
    dist.all_gather_into_tensor(edges, edge.view(-1))
    e = edges.cpu().numpy().reshape(world, 2, 10)
    ordered = all(bytes(e[r, 1]) <= bytes(e[r + 1, 0]) for r in range(world - 1))
    rows = torch.tensor([out.num_rows], dtype=torch.int64, device="cuda")
    allrows = torch.empty(world, dtype=torch.int64, device="cuda")
    dist.all_gather_into_tensor(allrows, rows)
    if rank == 0:
        t = tot.cpu().tolist()
        total_bytes = world * n * 100
        print(json.dumps({
            "config": f"GraySort-format sort, {world} x {n} x 100-byte records ({total_bytes / 1e9:.1f} GB), HBM resident, clear_input_blocks",
            "ms": float(ms.item()), "GB_per_s": total_bytes / float(ms.item()) / 1e6,
            "rows_per_gpu_out": allrows.cpu().tolist(),
            "checksum_preserved": t[0] % (1 << 62) == t[1] % (1 << 62), "records_in": t[2], "records_out": t[3],
            "unordered_records": t[4], "rank_boundaries_ordered": ordered,
            "free_hbm_before_gb": free0 / 1e9,
            "peak_torch_allocated_gb_rank0": torch.cuda.max_memory_allocated() / 1e9,
            "exchange": os.environ.get("DS_EXCHANGE", "split"),
        }), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
