#!/usr/bin/env python
"""Headline benchmark: Dataset.sort of GraySort-format 100-byte records.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

One step = one Dataset.sort(key) over this rank's batch of synthetic records
({key: fixed_size_binary(10), payload: fixed_size_binary(90)}, BASELINE.json
configs[1]: 100 M records = 10 GB per GPU).  Rank 0 prints ONE JSON line:

  value     sort GB/s (input bytes / device time), inputs resident in HBM, whole job
  e2e       the same through the public API from pinned HOST buffers: H2D copy of the
            batch + sort + D2H copy of the sorted batch inside the timed region
  roofline  the dominant kernel of the step (the payload gather; the radix pass kernel is
            reported next to it as roofline_second_kernel): algorithmic bytes per launch /
            CUDA-event duration of each launch (events recorded inside the library around
            the launches), against MEASURED_PEAKS.json's hbm_gbs
  cpu_baseline  the CPU restatement of the reference path (oracle/, kind "port") on a
            bounded sample, all host cores

--impl reference times that CPU path alone (the reference is Python over
pyarrow/numpy; see oracle/cpu_baseline.py for why the port is what runs here).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ROW_BYTES = 100
KEY_BYTES = 10
PAYLOAD_BYTES = 90
PASS_BYTES_PER_ROW = 2 * (KEY_BYTES + 4)  # read + write (key + idx)
METRIC = "sort_throughput_100B_records"
UNIT = "GB/s"


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--rows-per-gpu", type=int, default=100_000_000)
    p.add_argument("--cpu-rows", type=int, default=4_000_000,
                   help="rows of the bounded CPU-baseline sample")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--e2e-inflight", type=int, default=2,
                   help="single-GPU e2e: steps in flight, 1..4 (>= 2: H2D of step k+1 overlaps D2H of step k)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--radix-config", type=int, default=-1, help="tuning: pass-kernel variant")
    p.add_argument("--gather-mode", type=int, default=-1, help="tuning: see csrc/gather.cu g_gather_mode")
    p.add_argument("--l2-fetch", type=int, default=-1, help="tuning: cudaLimitMaxL2FetchGranularity")
    return p.parse_args()


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "200", "-i", str(self.gpu)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax = float(f[2])
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": smax,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_baseline(rows: int, steps: int = 1, warmup: int = 0):
    """(GB/s, descriptor) of the CPU port on a bounded sample, all host cores."""
    from oracle import cpu_baseline as C

    cores = C.host_cores()
    blocks = max(cores, 2)
    tabs = C.make_graysort_blocks(rows, blocks, seed=1)
    for _ in range(warmup):
        C.time_sort(tabs, cores)
    secs = [C.time_sort(tabs, cores) for _ in range(max(steps, 1))]
    best = min(secs)
    gbs = rows * ROW_BYTES / best / 1e9
    desc = {"value": gbs, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{rows} GraySort records ({rows * ROW_BYTES / 1e9:.2f} GB) in {blocks} "
                      f"blocks, best of {len(secs)}: {best:.2f} s; oracle/datasets_oracle.py "
                      f"(pyarrow sort_indices/take + numpy searchsorted) over a {cores}-process pool"}
    return gbs, desc, secs


# ------------------------------------------------------------------------------ ours
def make_records(n: int, seed: int):
    """Synthetic GraySort batch in HBM: random 10-byte keys, payload = LE row id + filler."""
    import collections

    import pyarrow as pa
    import torch

    import ray_memory_management_b200 as ds

    g = torch.Generator(device="cuda").manual_seed(seed)
    keys = torch.randint(0, 256, (n * KEY_BYTES,), dtype=torch.uint8, device="cuda", generator=g)
    payload = torch.zeros((n, PAYLOAD_BYTES), dtype=torch.uint8, device="cuda")
    ids = torch.arange(n, dtype=torch.int64, device="cuda")
    payload[:, :8] = ids.view(torch.uint8).view(n, 8)
    del ids
    cols = collections.OrderedDict(
        key=ds.DeviceColumn(pa.binary(KEY_BYTES), keys),
        payload=ds.DeviceColumn(pa.binary(PAYLOAD_BYTES), payload.view(-1)),
    )
    return ds.DeviceArrowBlock(cols, n)


def check_sorted(block, n):
    """Cheap on-device sanity check of one output (not timed): keys non-decreasing in
    their first 8 bytes and row ids a permutation."""
    import torch

    k = block.columns["key"].data.view(n, KEY_BYTES)[:, :8].contiguous()
    # big-endian bytes -> compare as int64 after byte swap; shift to keep the sign out of it
    be = k.flip(1).contiguous().view(torch.int64).view(n)
    hi = (be >> 1) & 0x7FFFFFFFFFFFFFFF
    assert bool((hi[1:] >= hi[:-1]).all()), "output keys are not sorted"
    ids = block.columns["payload"].data.view(n, PAYLOAD_BYTES)[:, :8].contiguous().view(torch.int64).view(n)
    assert int(ids.sum().item()) == n * (n - 1) // 2, "row ids are not a permutation"


def _bind_to_gpu_numa_node(gpu_index: int) -> None:
    """One process per GPU: run on the CPU cores next to that GPU (NVML's ideal affinity), so
    that the pinned host buffers of the e2e leg are first-touched on the GPU's own NUMA node
    instead of all ranks sharing one socket's memory controllers.  Best effort."""
    try:
        import pynvml

        import torch

        pynvml.nvmlInit()
        try:  # NVML ignores CUDA_VISIBLE_DEVICES: find the device by its PCI address
            p = torch.cuda.get_device_properties(gpu_index)
            h = pynvml.nvmlDeviceGetHandleByPciBusId(
                f"{p.pci_domain_id:08x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0".encode())
        except AttributeError:
            h = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
        pynvml.nvmlDeviceSetCpuAffinity(h)
    except Exception as e:  # no NVML, or affinity not permitted in this container
        print(f"[bench] CPU affinity not set for GPU {gpu_index}: {e!r}", file=sys.stderr)


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        _bind_to_gpu_numa_node(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    import ray_memory_management_b200 as ds
    from ray_memory_management_b200 import _lib

    lib = _lib.lib
    lib.ds_debug_launch_count.restype = __import__("ctypes").c_uint64
    if args.radix_config >= 0:
        lib.ds_debug_set_radix_config(args.radix_config)
    if args.gather_mode >= 0:
        lib.ds_debug_set_gather_mode(args.gather_mode)
    if args.l2_fetch > 0:
        print("l2 fetch granularity was", lib.ds_debug_set_l2_fetch_granularity(args.l2_fetch), file=sys.stderr)

    n = args.rows_per_gpu
    block = make_records(n, seed=1 + rank)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_step():
        if world > 1:
            from ray_memory_management_b200 import distributed

            out = distributed.sort_block(block, "key", descending=False)
        else:
            bl = ds.BlockList.from_blocks([block])
            res, _ = ds.sort_impl(bl, False, "key")
            out = res.get_blocks()[0]
        return out

    # the first exchanges on a fresh box also set up NCCL and the symmetric buffers
    warmup = args.warmup if world == 1 else max(args.warmup, 5)
    for i in range(warmup):
        out = one_step()
        if i == 0 and world == 1:
            torch.cuda.synchronize()
            check_sorted(out, n)
        del out

    import ctypes

    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    launches0 = lib.ds_debug_launch_count()
    lib.ds_debug_pass_timing(1)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        out = one_step()
        del out
    ev1.record()
    barrier()
    lib.ds_debug_pass_timing(0)
    launches = int(lib.ds_debug_launch_count() - launches0)
    ms = ev0.elapsed_time(ev1)
    timed = {}
    for kind, name in ((0, "pass"), (1, "gather")):
        t_ms, t_cnt, t_bytes = ctypes.c_double(0), ctypes.c_int32(0), ctypes.c_double(0)
        lib.ds_debug_collect_timing(kind, ctypes.byref(t_ms), ctypes.byref(t_cnt), ctypes.byref(t_bytes))
        timed[name] = (t_ms.value, t_cnt.value, t_bytes.value)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    value = world * n * ROW_BYTES / (ms_per_step * 1e-3) / 1e9

    peak, peak_src = measured_peak()
    try:
        with open(os.path.join(ROOT, "profiles", "r01_kernel_traffic.json")) as f:
            traffic_per_alg_byte = json.load(f)
    except Exception:
        traffic_per_alg_byte = {}

    def roofline_of(name, label):
        t_ms, t_cnt, t_bytes = timed[name]
        if not t_cnt:
            return None
        achieved = t_bytes / (t_ms * 1e-3) / 1e9           # algorithmic bytes / CUDA-event time
        ratio = traffic_per_alg_byte.get(name, {}).get("dram_bytes_per_algorithmic_byte")
        return {
            "kernel": label, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak, "frac_of_8TBs_spec": achieved / 8000.0,
            "traffic": (t_bytes / t_cnt * ratio) if ratio else None,
            "traffic_source": "ncu dram__bytes_read.sum+dram__bytes_write.sum per launch (profiles/r01_kernel_traffic.json), "
                              "scaled to this launch's rows",
            "peak_source": peak_src, "launches_timed": t_cnt, "avg_launch_ms": t_ms / t_cnt,
            "algorithmic_bytes_per_launch": t_bytes / t_cnt,
            "share_of_step": t_ms / ms if world == 1 else None,
        }

    roof_gather = roofline_of("gather", "gather_rows_kernel<uint16> (payload gather by permutation, 4 + 2*90 B per row)")
    roof_pass = roofline_of("pass", "onesweep_pass_kernel_v3 (one 8-bit LSD pass over key80+idx32, 28 B per row)")
    # the dominant kernel is the one with the larger share of the timed region
    dominant, other = (roof_gather, roof_pass)
    if roof_gather is None or (roof_pass is not None and timed["pass"][0] > timed["gather"][0]):
        dominant, other = roof_pass, roof_gather

    if world > 1 and os.environ.get("DS_DIST_TIMING") == "1":
        from ray_memory_management_b200 import distributed as _D

        one_step()  # collective: every rank takes part, rank 0 reports
        if rank == 0:
            print("dist phases (ms):", [(k, round(v, 2)) for k, v in _D.LAST_TIMING], file=sys.stderr)

    # ---- e2e: pinned host buffers in, pinned host buffers out
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, ds, block, n, world, rank, barrier)
    del block
    # sampled from the start of the timed region to the end of the e2e region
    clocks = sampler.stop() if rank == 0 else None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "rows_per_s": world * n / (ms_per_step * 1e-3),
        "config": {
            "workload": f"GraySort-format Dataset.sort: {n} x 100-byte records per GPU "
                        f"(10-byte key fixed_size_binary(10) + 90-byte payload), "
                        f"{world * n * ROW_BYTES / 1e9:.1f} GB total, BASELINE.json configs[1] per GPU",
            "rows_per_gpu": n, "blocks_per_gpu": 1,
            "l2": f"inputs ({n * ROW_BYTES / 1e9:.1f} GB/GPU) far exceed the 126 MB L2; no flush needed",
            "parallelism": f"range-partitioned x{world}" if world > 1 else "single GPU",
        },
        "clocks": clocks,
        "gpu_launches": launches,
        "roofline": dominant,
        "roofline_second_kernel": other,
        "e2e": e2e,
    }
    if not args.no_cpu_baseline and world == 1:
        _, desc, _ = cpu_baseline(args.cpu_rows, steps=2)
        line["cpu_baseline"] = desc
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_e2e(args, ds, block, n, world, rank, barrier):
    import pyarrow as pa
    import torch
    import torch.distributed as dist

    schema = [("key", pa.binary(KEY_BYTES)), ("payload", pa.binary(PAYLOAD_BYTES))]
    hk = torch.empty(n * KEY_BYTES, dtype=torch.uint8, pin_memory=True)
    hp = torch.empty(n * PAYLOAD_BYTES, dtype=torch.uint8, pin_memory=True)
    hk.copy_(block.columns["key"].data)
    hp.copy_(block.columns["payload"].data)
    ok = torch.empty(int(n * KEY_BYTES * (1.5 if world > 1 else 1)), dtype=torch.uint8, pin_memory=True)
    op = torch.empty(int(n * PAYLOAD_BYTES * (1.5 if world > 1 else 1)), dtype=torch.uint8, pin_memory=True)
    torch.cuda.synchronize()

    def step():
        blk = ds.DeviceArrowBlock.from_host_buffers(schema, [hk, hp], n)
        if world > 1:
            from ray_memory_management_b200 import distributed

            out = distributed.sort_block(blk, "key", descending=False)
        else:
            out = ds.Dataset.from_blocks([blk]).sort("key").get_blocks()[0]
        del blk
        out.to_host_buffers([ok, op])
        return out.num_rows

    # Single GPU: several steps in flight, each on its own stream with its own pinned output
    # buffers, so step k's D2H overlaps step k+1's H2D (PCIe is full duplex) and, with three,
    # step k+2's H2D does not have to wait for step k's D2H either.  Every step still copies
    # its whole input in and its whole result out inside the timed region.
    inflight = max(1, min(args.e2e_inflight, 4)) if world == 1 else 1
    outs = [(ok, op)]
    streams = [torch.cuda.current_stream()]
    if inflight >= 2:
        for _ in range(inflight - 1):
            outs.append((torch.empty_like(ok, pin_memory=True), torch.empty_like(op, pin_memory=True)))
        streams = [torch.cuda.Stream() for _ in range(inflight)]

    def step_on(i):
        nonlocal ok, op
        ok, op = outs[i % inflight]
        if inflight == 1:
            return step()
        with torch.cuda.stream(streams[i % inflight]):
            return step()

    if world > 1 and args.e2e_inflight >= 2:
        return _run_e2e_pipelined(args, ds, schema, (hk, hp), (ok, op), n, world, barrier)

    for i in range(inflight):
        step_on(i)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    cur = torch.cuda.current_stream()
    ev0.record()
    for st in streams:
        st.wait_stream(cur)
    for i in range(args.steps):
        step_on(i)
    for st in streams:
        cur.wait_stream(st)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    per = ms / args.steps
    # spot-check what landed in host memory
    if world == 1:
        a = ok[: 8 * KEY_BYTES].view(8, KEY_BYTES)
        assert bytes(a[0].tolist()) <= bytes(a[7].tolist())
    return {"value": world * n * ROW_BYTES / (per * 1e-3) / 1e9, "unit": UNIT,
            "h2d_bytes_per_step": world * n * ROW_BYTES, "d2h_bytes_per_step": world * n * ROW_BYTES,
            "ms_per_step": per, "steps_in_flight": inflight}


def _run_e2e_pipelined(args, ds, schema, host_in, host_out, n, world, barrier):
    """N > 1: the sort itself (collectives + shared peer buffers) stays one step at a time on
    the main stream; the host copies move to two side streams, so step k+1's H2D and step k's
    D2H run beside step k+1's sort (PCIe is full duplex).  Every step still copies its whole
    input in and its whole result out inside the timed region."""
    import torch
    import torch.distributed as dist

    from ray_memory_management_b200 import distributed

    main = torch.cuda.current_stream()
    up, down = torch.cuda.Stream(), torch.cuda.Stream()
    outs = [host_out, tuple(torch.empty_like(b, pin_memory=True) for b in host_out)]
    torch.cuda.synchronize()

    def upload():
        with torch.cuda.stream(up):
            blk = ds.DeviceArrowBlock.from_host_buffers(schema, list(host_in), n)
            ev = torch.cuda.Event()
            ev.record(up)
        return blk, ev

    def run(steps):
        nxt = upload()
        pending = []  # device results whose D2H is still running
        for k in range(steps):
            blk, ev = nxt
            main.wait_event(ev)
            for c in blk.columns.values():  # allocated on `up`, consumed on `main`
                c.data.record_stream(main)
            if k + 1 < steps:
                nxt = upload()
            out = distributed.sort_block(blk, "key", descending=False)
            del blk
            done = torch.cuda.Event()
            done.record(main)
            down.wait_event(done)
            with torch.cuda.stream(down):
                out.to_host_buffers(list(outs[k % 2]))
            for c in out.columns.values():  # produced on `main`, read on `down`
                c.data.record_stream(down)
            pending.append(out)
            if len(pending) > 1:
                pending.pop(0)
        main.wait_stream(down)

    run(2)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    run(args.steps)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    per = float(t.item()) / args.steps
    # spot-check what landed in this rank's host memory (its slice of the sorted output)
    a = outs[(args.steps - 1) % 2][0][: 8 * KEY_BYTES].view(8, KEY_BYTES)
    assert bytes(a[0].tolist()) <= bytes(a[7].tolist())
    return {"value": world * n * ROW_BYTES / (per * 1e-3) / 1e9, "unit": UNIT,
            "h2d_bytes_per_step": world * n * ROW_BYTES, "d2h_bytes_per_step": world * n * ROW_BYTES,
            "ms_per_step": per, "steps_in_flight": 2,
            "overlap": "H2D of step k+1 and D2H of step k on side streams; one sort at a time"}


# ------------------------------------------------------------------------------ reference arm
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    rows = args.cpu_rows
    gbs, desc, secs = cpu_baseline(rows, steps=args.steps, warmup=min(args.warmup, 1))
    per = sum(secs) / len(secs)
    gbs_mean = rows * ROW_BYTES / per / 1e9
    desc = dict(desc, value=gbs_mean)
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs_mean, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": per * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": {"workload": f"GraySort-format Dataset.sort on host cores: bounded sample of "
                               f"{rows} x 100-byte records per step (same record format as the GPU arm)",
                   "rows_per_step": rows},
        "cpu_baseline": desc,
        "e2e": {"value": gbs_mean, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
