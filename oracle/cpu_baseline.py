"""TEST/BENCH INFRASTRUCTURE ONLY -- times the CPU restatement of the reference path.

Used by bench.py for the ``cpu_baseline`` object and for ``--impl reference``.  The
reference itself is Python over pyarrow/numpy and needs /root/reference to run, which
does not exist on the GPU box, so what is timed is oracle/datasets_oracle.py (kind
"port"): the same pyarrow / numpy calls the reference makes, with the reference's
task parallelism (one map task per block, one reduce task per output block;
python/ray/data/_internal/shuffle.py:81-112) approximated by a multiprocessing pool
with one worker per host core.  Blocks cross process boundaries pickled, as they cross
the object store in the reference.
"""
import multiprocessing as mp
import os
import random
import time

import numpy as np
import pyarrow as pa

from oracle import datasets_oracle as O

_G = {}


def _ser(t: pa.Table) -> bytes:
    # Arrow IPC writes only the sliced range (what the reference's _copy_table,
    # _internal/arrow_block.py:590-605, is there to guarantee for plasma)
    sink = pa.BufferOutputStream()
    with pa.ipc.new_stream(sink, t.schema) as w:
        w.write_table(t)
    return sink.getvalue().to_pybytes()


def _de(b: bytes) -> pa.Table:
    return pa.ipc.open_stream(b).read_all()


def _map_task(args):
    i, desc, ser, boundaries = args
    parts = O.sort_and_partition(_G["tables"][i], boundaries, "key", desc)
    return [_ser(p) for p in parts] if ser else parts


def _reduce_task(args):
    parts, desc, ser = args
    if ser:
        parts = [_de(p) for p in parts]
    return O.merge_sorted_blocks(parts, "key", desc).num_rows


def make_pool(tables, procs: int):
    """The worker processes, started BEFORE any timed region (the reference's Ray workers are
    running when Dataset.sort is called).  fork: the workers inherit the input tables, which
    stands for blocks that are "already materialised in the object store"."""
    _G["tables"] = tables
    return mp.get_context("fork").Pool(procs)


def make_graysort_blocks(n_rows: int, n_blocks: int, seed: int = 1):
    t = O.graysort_table(n_rows, seed)
    cuts = np.linspace(0, n_rows, n_blocks + 1).astype(int)
    return [t.slice(a, b - a) for a, b in zip(cuts[:-1], cuts[1:])]


def time_sort(tables, procs: int, descending: bool = False, pool=None) -> float:
    """Seconds for one Dataset.sort of `tables` (sample_boundaries + map + reduce).  With
    ``procs`` > 1 the map and reduce tasks run in ``pool`` (make_pool; created here, outside
    the timed region, when none is given); map outputs come back pickled."""
    r = len(tables)
    own = None
    if procs > 1 and pool is None:
        own = pool = make_pool(tables, procs)
    _G["tables"] = tables
    t0 = time.perf_counter()
    random.seed(0)
    boundaries = O.sample_boundaries(tables, "key", r)
    if procs <= 1:
        maps = [_map_task((i, descending, False, boundaries)) for i in range(r)]
        for j in range(r):
            _reduce_task(([maps[i][j] for i in range(r)], descending, False))
    else:
        maps = pool.map(_map_task, [(i, descending, True, boundaries) for i in range(r)], chunksize=1)
        pool.map(_reduce_task, [([maps[i][j] for i in range(r)], descending, True) for j in range(r)], chunksize=1)
    dt = time.perf_counter() - t0
    if own is not None:
        own.close()
        own.join()
    return dt


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:  # pragma: no cover
        return os.cpu_count() or 1
