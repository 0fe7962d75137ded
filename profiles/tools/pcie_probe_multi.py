"""Concurrent host<->device copies on 1, 2, 4, 8 GPUs of one box (torchrun, one rank per GPU):
the ceiling of bench.py's e2e number at N GPUs.  In phase k only ranks < k copy (H2D and D2H at
once, pinned memory, 4 GB each way, 3 repetitions); rank 0 prints per-GPU and aggregate GB/s per
direction for every k.

    python -m torch.distributed.run --nproc-per-node 8 profiles/tools/pcie_probe_multi.py [GB] [bind]
bind = 1: ranks bind to their GPU's NUMA node first (what bench.py does).
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
gb = float(sys.argv[1]) if len(sys.argv) > 1 else 4.0
if len(sys.argv) > 2 and sys.argv[2] == "1":
    from bench import _bind_to_gpu_numa_node

    _bind_to_gpu_numa_node(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(gb * 1e9)
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_in.fill_(1)
h_out.fill_(2)
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def both():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def timed(reps=3):
    both()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    cur = torch.cuda.current_stream()
    e0.record()
    s1.wait_stream(cur)
    s2.wait_stream(cur)
    for _ in range(reps):
        both()
    cur.wait_stream(s1)
    cur.wait_stream(s2)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


res = {"bytes_each_way_per_gpu": n, "numa_bound": len(sys.argv) > 2 and sys.argv[2] == "1", "cpus": os.cpu_count()}
k = 1
while k <= world:
    dist.barrier()
    torch.cuda.synchronize()
    ms = timed() if rank < k else 0.0
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    allt = torch.empty(world, dtype=torch.float64, device="cuda")
    dist.all_gather_into_tensor(allt, t)
    times = [x for x in allt.cpu().tolist()[:k]]
    res[f"{k}_gpus"] = {"GBs_per_direction_per_gpu_min": round(n / max(times) / 1e6, 2),
                        "GBs_per_direction_per_gpu_max": round(n / min(times) / 1e6, 2),
                        "GBs_per_direction_aggregate": round(k * n / max(times) / 1e6, 2)}
    k *= 2
if rank == 0:
    print(json.dumps(res))
dist.destroy_process_group()
