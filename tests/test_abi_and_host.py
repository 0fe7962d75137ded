"""CPU: the C-ABI library loads and exports every symbol include/ds_shuffle.h declares
(no compute calls), and the host-side logic that needs no GPU."""
import ctypes
import os
import re

import numpy as np
import pyarrow as pa
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    with open(os.path.join(ROOT, "include", "ds_shuffle.h")) as f:
        src = f.read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ds_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from ray_memory_management_b200 import _lib

    names = _declared_symbols()
    assert len(names) >= 19
    for n in names:
        assert hasattr(_lib.lib, n), f"{n} declared in ds_shuffle.h but not exported"
    assert sorted(_lib.EXPORTED) == names
    assert _lib.lib.ds_version() >= 100
    assert isinstance(_lib.lib.ds_last_error(), bytes)


def test_workspace_size_queries_need_no_gpu():
    from ray_memory_management_b200 import _lib

    need = ctypes.c_size_t(0)
    n = 1_000_000
    assert _lib.lib.ds_sort_indices_fsb(None, 10, n, 0, None, None, 0, None, ctypes.byref(need), None) == 0
    # two buffers of hi (8 B) and lo (2 B), one spare idx (4 B), + look-back + control
    assert 24 * n <= need.value <= 26 * n
    assert _lib.lib.ds_sort_indices_i64(None, n, 0, None, None, 0, None, ctypes.byref(need), None) == 0
    assert 20 * n <= need.value <= 22 * n
    assert _lib.lib.ds_sort_indices_fsb(None, 12, n, 0, None, None, 0, None, ctypes.byref(need), None) == _lib.DS_ERR_UNSUPPORTED
    assert b"width" in _lib.lib.ds_last_error()
    assert _lib.lib.ds_sort_indices_i64(None, 1 << 30, 0, None, None, 0, None, ctypes.byref(need), None) == _lib.DS_ERR_TOO_LARGE


def test_subrange_key_transform_arithmetic():
    """host arithmetic of the hybrid sort plan on a sub-range of the key space (KeyXform in
    csrc/radix_sort.cu): for any sampled [min, max] the 28-bit G = (hi' * R) >> 64 of
    hi' = (hi - base) << s is monotone in hi, stays below 2^28, reaches the top of that range,
    and (hi' >> s) + base gives the key back.  Checked with Python integers."""
    import random

    from ray_memory_management_b200 import _lib

    f = _lib.lib.ds_debug_key_xform
    f.argtypes = [ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64, ctypes.POINTER(ctypes.c_uint64)]
    f.restype = ctypes.c_int32
    out = (ctypes.c_uint64 * 4)()
    M = (1 << 64) - 1
    rnd = random.Random(5)
    used = 0
    for _ in range(3000):
        bits = rnd.choice([1, 9, 17, 33, 47, 62, 63, 64])
        mn = rnd.getrandbits(64)
        mx = min(M, mn + rnd.getrandbits(bits))
        n = rnd.choice([1, 1000, 100_000_000])
        on = f(mn, mx, n, out)
        base, span, R, s = (int(v) for v in out)
        if not on:
            # nothing sampled apart, the range is the whole key space, or too few distinct values
            margin = (mx - mn) // 64 + 1
            lo = mn - margin if mn > margin else 0
            hi = mx + margin if mx < M - margin else M
            assert mx <= mn or hi - lo == M or hi - lo < n // 4
            continue
        used += 1
        assert base <= mn and mx <= base + span - 1 <= M
        assert (span << s) >> 63 == 1 and (span << s) <= M   # hi' fills 64 bits, no overflow
        prev = -1
        for d0 in sorted([0, span - 1, mn - base, mx - base] + [rnd.randrange(span) for _ in range(16)]):
            hp = d0 << s
            g = (hp * R) >> 64
            assert prev <= g < (1 << 28)
            prev = g
            assert (hp >> s) + base == base + d0
        if span >= 1 << 30:
            assert prev >= (1 << 28) - 2                       # the top of the range maps to the last bucket
    assert used > 1000
    assert f(5, 5, 10, out) == 0 and f(0, M, 10, out) == 0
    assert f(5000, 6000, 300_000, out) == 0                    # 1000 values for 300 k rows


def test_nearest_rank_matches_oracle_and_golden(golden):
    from oracle import datasets_oracle as O
    from ray_memory_management_b200 import sort

    for ns in (1, 2, 3, 7, 10, 33, 40, 80, 100, 1001):
        for r in (1, 2, 3, 4, 5, 8, 16):
            assert sort.nearest_rank_positions(ns, r) == list(O.nearest_rank_indices(ns, r))
    g = golden("boundaries")
    for nsamp, r in g["q_cases"]:
        got = sort.boundaries_from_samples(pa.array(g[f"q_{nsamp}_{r}_samples"]), int(r))
        assert np.array_equal(np.array(got, dtype=np.int64), g[f"q_{nsamp}_{r}_bounds"])
    assert sort.boundaries_from_samples(pa.array([], pa.int64()), 4) == [None] * 3


def test_boundaries_from_fsb_samples_are_memcmp_ordered():
    from oracle import datasets_oracle as O
    from ray_memory_management_b200 import sort

    rng = np.random.default_rng(0)
    rows = rng.integers(0, 256, (80, 10), dtype=np.uint8)
    rows[:40, :9] = 0  # includes embedded / trailing NUL bytes
    got = sort.boundaries_from_samples(O.fsb_from_numpy(rows), 8)
    srt = sorted(bytes(r) for r in rows)
    assert got == [srt[i] for i in O.nearest_rank_indices(80, 8)]


def test_agg_names_and_shuffle_op_surface():
    import ray_memory_management_b200 as ds
    from ray_memory_management_b200.block import resolve_agg_names

    aggs = (ds.Sum("v"), ds.Sum("v"), ds.Count(), ds.Count(), ds.Mean("v"))
    assert resolve_agg_names(aggs) == ["sum(v)", "sum(v)_2", "count()", "count()_2", "mean(v)"]
    op = ds.ShuffleOp(map_args=[1], reduce_args=[2])
    assert op._map_args == [1] and op._reduce_args == [2]
    with pytest.raises(NotImplementedError):
        ds.ShuffleOp.map(0, None, 1)
    import inspect

    sig = inspect.signature(ds.DeviceShufflePlan.execute)
    assert list(sig.parameters) == ["self", "input_blocks", "output_num_blocks", "clear_input_blocks",
                                    "map_ray_remote_args", "reduce_ray_remote_args"]
