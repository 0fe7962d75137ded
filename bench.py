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
DEFAULT_PLAN = "split"  # multi-GPU plan of a run without --plan
UNIT = "GB/s"


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--rows-per-gpu", type=int, default=100_000_000)
    p.add_argument("--cpu-rows", type=int, default=20_000_000,
                   help="rows of the bounded CPU-baseline sample")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--e2e-inflight", type=int, default=2,
                   help="single-GPU e2e: steps in flight, 1..4 (>= 2: H2D of step k+1 overlaps D2H of step k)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--radix-config", type=int, default=-1, help="tuning: pass-kernel variant")
    p.add_argument("--gather-mode", type=int, default=-1, help="tuning: see csrc/gather.cu g_gather_mode (0 auto, 1, 2, 7, 8)")
    p.add_argument("--l2-fetch", type=int, default=-1, help="tuning: cudaLimitMaxL2FetchGranularity")
    p.add_argument("--blocks-per-gpu", type=int, default=20,
                   help="single GPU: also time the same records held as this many blocks (value_multiblock)")
    p.add_argument("--corrupt", action="store_true",
                   help="flip one byte of the first output before validating it: the run must exit non-zero")
    p.add_argument("--no-validate", action="store_true")
    p.add_argument("--plan", default="auto", choices=["auto", "split", "push"],
                   help="N > 1: one exchange round (split) or the push-based plan in rounds (push); "
                        "auto = the plan DESIGN.md records as the faster one")
    p.add_argument("--rounds", type=int, default=0, help="rounds of the push-based plan (0 = library default)")
    p.add_argument("--staged", type=int, default=-1,
                   help="push-based plan: trailing rounds that are staged locally and shipped by the copy "
                        "engines (-1 = library default: all but the first)")
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


def cpu_rows_that_fit(rows: int) -> int:
    """The CPU sample needs ~6x its size in host memory (tables, pickled map outputs, sorted
    copies in the workers); shrink it when the box is small."""
    try:
        with open("/proc/meminfo") as f:
            avail_kb = next(int(ln.split()[1]) for ln in f if ln.startswith("MemAvailable"))
        cap = int(avail_kb * 1024 / (6 * ROW_BYTES))
        return max(1_000_000, min(rows, cap))
    except Exception:
        return rows


def cpu_baseline(rows: int, steps: int = 1, warmup: int = 0):
    """(GB/s, descriptor) of the CPU port on a bounded sample, all host cores.  The worker pool
    is started before the timed region (the reference's workers are running when Dataset.sort is
    called) and serves every step."""
    from oracle import cpu_baseline as C

    rows = cpu_rows_that_fit(rows)
    cores = C.host_cores()
    blocks = max(cores, 2)
    tabs = C.make_graysort_blocks(rows, blocks, seed=1)
    pool = C.make_pool(tabs, cores) if cores > 1 else None
    try:
        for _ in range(warmup):
            C.time_sort(tabs, cores, pool=pool)
        secs = [C.time_sort(tabs, cores, pool=pool) for _ in range(max(steps, 1))]
    finally:
        if pool is not None:
            pool.close()
            pool.join()
    best = min(secs)
    gbs = rows * ROW_BYTES / best / 1e9
    desc = {"value": gbs, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{rows} GraySort records ({rows * ROW_BYTES / 1e9:.2f} GB) in {blocks} "
                      f"blocks, best of {len(secs)}: {best:.2f} s; oracle/datasets_oracle.py "
                      f"(pyarrow sort_indices/take + numpy searchsorted) over a {cores}-process pool "
                      f"started before the timed region"}
    return gbs, desc, secs, rows


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


def _wrap64(c: int) -> int:
    return c - (1 << 64) if c >= (1 << 63) else c


def validate_output(ds, block_in, out, world, rank):
    """Untimed proof that one step's output is the sort of its input, on every rank:
    valsort's invariants on device (ds_validate_records: no record below its predecessor, the
    sum of the records' CRC-32 equal to the input's, record count equal) summed over ranks,
    and the ranks' key ranges in rank order.  Returns the dict that goes into the JSON line."""
    import torch
    import torch.distributed as dist

    from ray_memory_management_b200 import graysort

    vin, vout = graysort.validate(block_in), graysort.validate(out)
    t = torch.tensor([vin["records"], _wrap64(vin["checksum"]), vout["records"], _wrap64(vout["checksum"]),
                      vout["unordered"]], dtype=torch.int64, device="cuda")
    n_out = out.num_rows
    edge = torch.zeros(2 * KEY_BYTES + 1, dtype=torch.uint8, device="cuda")
    if n_out:
        k = out.columns["key"].data
        edge[:KEY_BYTES] = k[:KEY_BYTES]
        edge[KEY_BYTES: 2 * KEY_BYTES] = k[(n_out - 1) * KEY_BYTES: n_out * KEY_BYTES]
        edge[2 * KEY_BYTES] = 1
    edges = edge[None, :]
    if world > 1:
        dist.all_reduce(t)  # int64 sums wrap, like the 64-bit checksum itself
        edges = torch.empty((world, 2 * KEY_BYTES + 1), dtype=torch.uint8, device="cuda")
        dist.all_gather_into_tensor(edges.view(-1), edge)
    rin, cin, rout, cout, unordered = (int(x) for x in t.cpu().tolist())
    e = edges.cpu().numpy()
    live = [(bytes(r[:KEY_BYTES]), bytes(r[KEY_BYTES: 2 * KEY_BYTES])) for r in e if r[2 * KEY_BYTES]]
    rank_order = all(live[i][1] <= live[i + 1][0] for i in range(len(live) - 1))
    res = {"records_in": rin, "records_out": rout, "crc_equal": cin == cout, "unordered": unordered,
           "rank_order": rank_order}
    res["ok"] = rin == rout and cin == cout and unordered == 0 and rank_order
    return res


def validate_tie_order(ds, block_in, out):
    """Single GPU, full size: the step's output against a second sort of the same keys that runs
    every radix pass (no prefix/repair, no bucket kernel): row order and sorted keys must be
    identical, and the payload must be the gather by that order -- this pins the tie order of the
    production plan (the bucket kernel at its production fill) bit for bit."""
    import torch

    from ray_memory_management_b200 import _lib
    from ray_memory_management_b200.block import _argsort, _gather

    key = block_in.columns["key"]
    idx_a, sk_a = _argsort(key, False, True)
    idx_b, sk_b = _argsort(key, False, True, options=_lib.DS_SORT_NO_HYBRID | _lib.DS_SORT_NO_PREFIX_REPAIR)
    same_idx = bool(torch.equal(idx_a, idx_b)) and bool(torch.equal(sk_a.data, sk_b.data))
    same_keys = bool(torch.equal(out.columns["key"].data, sk_b.data))
    del idx_a, sk_a, sk_b
    pay = _gather(block_in.columns["payload"], idx_b)
    same_payload = bool(torch.equal(out.columns["payload"].data, pay.data))
    return same_idx and same_keys and same_payload


def validate_miniature(ds, world, rank, sort_fn):
    """Every N: a ragged, tie-heavy miniature (4096 distinct keys) whose expected output -- the
    stable sort of the ranks' inputs in rank order -- every rank recomputes with torch.sort(stable)
    from the seeds; the rows of all ranks' outputs, in rank order, must match it exactly."""
    import collections

    import pyarrow as pa
    import torch
    import torch.distributed as dist

    def make(r):
        m = 120_000 + 1_000 * r
        g = torch.Generator(device="cuda").manual_seed(1000 + r)
        low = torch.randint(0, 4096, (m,), dtype=torch.int64, device="cuda", generator=g)
        keys = torch.full((m, KEY_BYTES), 3, dtype=torch.uint8, device="cuda")
        keys[:, 8] = (low >> 8).to(torch.uint8)
        keys[:, 9] = (low & 255).to(torch.uint8)
        ids = torch.arange(m, dtype=torch.int64, device="cuda") + (r << 32)
        payload = torch.zeros((m, PAYLOAD_BYTES), dtype=torch.uint8, device="cuda")
        payload[:, :8] = ids.view(torch.uint8).view(m, 8)
        return low, ids, keys, payload

    lows, idss = [], []
    for r in range(world):
        low, ids, keys, payload = make(r)
        lows.append(low)
        idss.append(ids)
        if r == rank:
            mine = ds.DeviceArrowBlock(collections.OrderedDict(
                key=ds.DeviceColumn(pa.binary(KEY_BYTES), keys.view(-1)),
                payload=ds.DeviceColumn(pa.binary(PAYLOAD_BYTES), payload.view(-1))), keys.shape[0])
    order = torch.sort(torch.cat(lows), stable=True).indices
    want = torch.cat(idss)[order]
    out = sort_fn(mine)
    got = out.columns["payload"].data.view(-1, PAYLOAD_BYTES)[:, :8].contiguous().view(torch.int64).view(-1)
    if world > 1:
        cap = int(want.numel())
        pad = torch.full((cap + 1,), -1, dtype=torch.int64, device="cuda")
        pad[0] = got.numel()
        pad[1: 1 + got.numel()] = got
        allp = torch.empty((world, cap + 1), dtype=torch.int64, device="cuda")
        dist.all_gather_into_tensor(allp.view(-1), pad)
        got = torch.cat([allp[r, 1: 1 + int(allp[r, 0])] for r in range(world)])
    return got.numel() == want.numel() and bool(torch.equal(got, want))


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

    plan = args.plan if args.plan != "auto" else DEFAULT_PLAN
    rounds = args.rounds or None

    def dist_sort(blk):
        """the rank's output blocks, in key order"""
        from ray_memory_management_b200 import distributed

        if plan == "push":
            return distributed.sort_block_push(blk, "key", descending=False, rounds=rounds,
                                               staged=args.staged if args.staged >= 0 else None)[0]
        return [distributed.sort_block(blk, "key", descending=False, rounds=1)]

    def one_step():
        if world > 1:
            out = dist_sort(block)
        else:
            bl = ds.BlockList.from_blocks([block])
            res, _ = ds.sort_impl(bl, False, "key")
            out = res.get_blocks()[0]
        return out

    def one_block(res):
        """a step's result as ONE block (validation only, untimed)"""
        if isinstance(res, list):
            return res[0] if len(res) == 1 else ds.DeviceArrowBlock.concat(res)
        return res

    def sort_one(blk):
        if world > 1:
            return one_block(dist_sort(blk))
        return ds.Dataset.from_blocks([blk]).sort("key").get_blocks()[0]

    # the first exchanges on a fresh box also set up NCCL and the symmetric buffers
    warmup = args.warmup if world == 1 else max(args.warmup, 5)
    validated = None
    for i in range(warmup):
        out = one_step()
        if i == 0 and not args.no_validate:
            out = one_block(out)
            torch.cuda.synchronize()
            if args.corrupt and rank == world - 1:
                out.columns["payload"].data[out.num_rows * PAYLOAD_BYTES // 2] ^= 1
            validated = validate_output(ds, block, out, world, rank)
            if world == 1:
                validated["idx_equal_all_passes"] = validate_tie_order(ds, block, out)
                validated["ok"] = validated["ok"] and validated["idx_equal_all_passes"]
            del out
            validated["miniature_bit_exact"] = validate_miniature(ds, world, rank, sort_one)
            validated["ok"] = validated["ok"] and validated["miniature_bit_exact"]
            if not validated["ok"]:
                if rank == 0:
                    print(json.dumps({"metric": METRIC, "validated": validated, "error": "output failed validation"}),
                          flush=True)
                sys.exit(3)
            continue
        del out

    import ctypes

    if world > 1:
        from ray_memory_management_b200 import distributed as _D

        _D.EXCHANGE_LOG = []
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    launches0 = lib.ds_debug_launch_count()
    lib.ds_debug_pass_timing(1)
    step_ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    # Objects that exist now (modules, the input, cached buffers) leave the collector's sight: a
    # full collection over them pauses one rank's host thread for tens of ms, and with a host
    # read-back inside the multi-GPU step every rank waits for it.  The collector stays on.
    import gc

    gc.collect()
    gc.freeze()
    host_t = [time.perf_counter()]
    step_ev[0].record()
    for k in range(args.steps):
        out = one_step()
        del out
        step_ev[k + 1].record()
        host_t.append(time.perf_counter())
    barrier()
    gc.unfreeze()
    host_ms = sorted((b - a) * 1e3 for a, b in zip(host_t, host_t[1:]))
    lib.ds_debug_pass_timing(0)
    launches = int(lib.ds_debug_launch_count() - launches0)
    ms = step_ev[0].elapsed_time(step_ev[-1])
    step_ms = [step_ev[k].elapsed_time(step_ev[k + 1]) for k in range(args.steps)]
    timed = {}
    for kind, name in ((0, "pass"), (1, "gather")):
        t_ms, t_cnt, t_bytes = ctypes.c_double(0), ctypes.c_int32(0), ctypes.c_double(0)
        lib.ds_debug_collect_timing(kind, ctypes.byref(t_ms), ctypes.byref(t_cnt), ctypes.byref(t_bytes))
        timed[name] = (t_ms.value, t_cnt.value, t_bytes.value)
    if world > 1:
        t = torch.tensor([ms] + step_ms, dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0].item())
        step_ms = [float(x) for x in t[1:].cpu().tolist()]
    ms_per_step = ms / args.steps
    value = world * n * ROW_BYTES / (ms_per_step * 1e-3) / 1e9
    ss = sorted(step_ms)
    step_stats = {"min": ss[0], "median": ss[len(ss) // 2], "max": ss[-1],
                  "host_issue_ms_rank0": {"median": host_ms[len(host_ms) // 2], "max": host_ms[-1]}}

    peak, peak_src = measured_peak()
    try:
        with open(os.path.join(ROOT, "profiles", "r02_kernel_traffic.json")) as f:
            traffic_per_alg_byte = json.load(f)
    except Exception:
        traffic_per_alg_byte = {}

    def roofline_of(name, label):
        t_ms, t_cnt, t_bytes = timed[name]
        if not t_cnt:
            return None
        achieved = t_bytes / (t_ms * 1e-3) / 1e9           # algorithmic bytes / CUDA-event time
        ent = traffic_per_alg_byte.get(name, {})
        ratio = ent.get("dram_bytes_per_algorithmic_byte")
        return {
            "kernel": label, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak, "frac_of_8TBs_spec": achieved / 8000.0,
            "traffic": (t_bytes / t_cnt * ratio) if ratio else None,
            "traffic_source": (f"ncu dram__bytes_read.sum+dram__bytes_write.sum of {ent.get('kernel')} at git "
                               f"{ent.get('git')} (profiles/r02_kernel_traffic.json), scaled to this launch's rows")
            if ratio else None,
            "peak_source": peak_src, "launches_timed": t_cnt, "avg_launch_ms": t_ms / t_cnt,
            "algorithmic_bytes_per_launch": t_bytes / t_cnt,
            "share_of_step": t_ms / ms if world == 1 else None,
        }

    roof_gather = roofline_of("gather", "payload gather by permutation (ds_gather_rows, 4 + 2*90 B per row)")
    roof_pass = roofline_of("pass", "onesweep_pass_kernel_v3 (one 8-bit LSD pass over key80+idx32, 28 B per row)")
    # the dominant kernel is the one with the larger share of the timed region
    dominant, other = (roof_gather, roof_pass)
    if roof_gather is None or (roof_pass is not None and timed["pass"][0] > timed["gather"][0]):
        dominant, other = roof_pass, roof_gather

    # ---- N > 1: the exchange against NVLink, from CUDA events inside sort_block
    roofline_exchange = None
    phases = None
    if world > 1:
        log = _D.exchange_log()
        _D.EXCHANGE_LOG = None
        if log:
            k = len(log)
            mean = {f: sum(e[f] for e in log) / k for f in log[0]}
            # a transfer is as slow as its slowest participant: max over ranks of the durations,
            # min over ranks of the byte counts (conservative numerator)
            t = torch.tensor([mean["ms_keys"], mean["ms_payload"], mean["ms_exposed_wait"], mean["ms_first_to_last"],
                              -mean["bytes_out"], -mean["bytes_in"], -mean["key_bytes_out"]],
                             dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            mk, mp_, mw, mt, bo, bi, kbo = (float(x) for x in t.cpu().tolist())
            bo, bi, kbo = -bo, -bi, -kbo
            achieved = (bo - kbo) / (mp_ * 1e-3) / 1e9 if mp_ > 0 else None
            roofline_exchange = {
                "bound": "nvlink", "unit": "GB/s per direction per GPU", "peak": 900.0,
                "peak_measured_peer_copy": 770.0,
                "bytes_out_per_gpu": bo, "bytes_in_per_gpu": bi, "key_bytes_out_per_gpu": kbo,
                "ms_keys_incl_barriers": mk, "ms_payload": mp_, "ms_exposed_wait_after_key_sort": mw,
                "ms_total_first_key_to_last_payload": mt,
                "overlap_hidden_ms": max(mp_ - mw, 0.0),
                "achieved": achieved, "frac": achieved / 900.0 if achieved else None,
                "frac_of_measured_peer_copy": achieved / 770.0 if achieved else None,
                "achieved_whole_exchange": bo / (mt * 1e-3) / 1e9 if mt > 0 else None,
                "how": "payload bytes this GPU stored into peers' memory / CUDA-event duration of the payload "
                       "split kernels (ds_split_rows, bulk stores over NVLink) on their stream; mean over the "
                       "timed steps, max over ranks; the whole-exchange figure divides all remote bytes by the "
                       "span from the first key byte to the last payload byte, key sort included",
                "steps": k,
            }
        # phases of one extra (untimed) step, and the same with the payload sent AFTER the key sort
        _D.set_phase_timing(True)
        one_step()
        phases = {"concurrent": [(a, round(b, 3)) for a, b in _D.LAST_TIMING]}
        _D.PAYLOAD_AFTER_SORT = True
        one_step()
        one_step()
        phases["payload_after_key_sort"] = [(a, round(b, 3)) for a, b in _D.LAST_TIMING]
        _D.PAYLOAD_AFTER_SORT = False
        _D.set_phase_timing(False)

    # ---- single GPU: the same records held as blocks_per_gpu blocks (sample_boundaries, the
    # partition cuts and the merge are then inside the timed region)
    multiblock = None
    if world == 1 and args.blocks_per_gpu > 1:
        nb = args.blocks_per_gpu
        cuts = [n * i // nb for i in range(nb + 1)]
        blocks = [block.slice(cuts[i], cuts[i + 1]) for i in range(nb)]

        def multi_step():
            res, _ = ds.sort_impl(ds.BlockList.from_blocks(blocks), False, "key")
            return res.get_blocks()

        for _ in range(3):
            del_me = multi_step()
        rows_out = sum(b.num_rows for b in del_me)
        assert rows_out == n, (rows_out, n)
        del del_me
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            del_me = multi_step()
            del del_me
        e1.record()
        torch.cuda.synchronize()
        mb_ms = e0.elapsed_time(e1) / args.steps
        multiblock = {"value": n * ROW_BYTES / (mb_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": mb_ms,
                      "blocks_per_gpu": nb, "output_blocks": nb,
                      "what": f"the same {n} records as {nb} input blocks of {n // nb} rows "
                              "(python/ray/data/context.py:19 targets 512 MiB blocks): splitter sampling, "
                              "range cuts and the merge run inside the timed region"}
        del blocks

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
            "exchange": os.environ.get("DS_EXCHANGE", "split") if world > 1 else None,
            "plan": (plan + (f" x{rounds}" if rounds and plan == "push" else "")) if world > 1 else None,
            "l2": f"inputs ({n * ROW_BYTES / 1e9:.1f} GB/GPU) far exceed the 126 MB L2; no flush needed",
            "parallelism": f"range-partitioned x{world}" if world > 1 else "single GPU",
            "host": "gc.freeze() before the timed steps (collector left on)",
        },
        "clocks": clocks,
        "gpu_launches": launches,
        "roofline": dominant,
        "roofline_second_kernel": other,
        "e2e": e2e,
        "validated": validated,
        "step_ms": step_stats,
        "value_multiblock": multiblock,
        "roofline_exchange": roofline_exchange,
        "dist_phases_ms": phases,
    }
    if not args.no_cpu_baseline and world == 1:
        _, desc, _, _ = cpu_baseline(args.cpu_rows, steps=2, warmup=1)
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

            out = distributed.sort_block(blk, "key", descending=False, rounds=1)
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
            out = distributed.sort_block(blk, "key", descending=False, rounds=1)
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
    # same warm-up count as the GPU arm, bounded so that the run ends within minutes
    warm = min(args.warmup, 3)
    gbs, desc, secs, rows = cpu_baseline(args.cpu_rows, steps=args.steps, warmup=warm)
    per = sum(secs) / len(secs)
    gbs_mean = rows * ROW_BYTES / per / 1e9
    desc = dict(desc, value=gbs_mean)
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs_mean, "unit": UNIT, "n_gpus": world,
        "steps": args.steps, "warmup": warm, "ms_per_step": per * 1e3,
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
