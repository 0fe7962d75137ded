"""GPU parity: Dataset.sort / sort_and_partition / merge_sorted_blocks through the C ABI,
against the oracle (seeded inputs) and the golden vectors recorded from the reference."""
import ctypes
import random

import numpy as np
import pyarrow as pa
import pyarrow.compute  # noqa: F401
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def ds():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ray_memory_management_b200 as m

    return m


from oracle import datasets_oracle as O  # noqa: E402


def _tabs(cuts, **cols):
    return [pa.table({k: v[a:b] for k, v in cols.items()}) for a, b in cuts]


def _argsort_c(ds, col_np, kind, descending, width=0, options=0, want_sorted=True):
    """Direct C-ABI call of ds_sort_indices_* on a numpy column."""
    lib = ds._lib
    n = len(col_np) if kind != "fsb" else col_np.shape[0]
    dev = torch.from_numpy(np.ascontiguousarray(col_np).view(np.uint8).reshape(-1).copy()).cuda()
    idx = torch.empty(n, dtype=torch.int32, device="cuda")
    out = torch.empty(dev.numel(), dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    optr = out.data_ptr() if want_sorted else None
    if kind == "fsb":
        args = (dev.data_ptr(), width, n, int(descending), idx.data_ptr(), optr, options)
    else:
        args = (dev.data_ptr(), n, int(descending), idx.data_ptr(), optr, options)
    ws = lib.call_ws(f"ds_sort_indices_{kind}", lambda b: torch.empty(b, dtype=torch.uint8, device="cuda"),
                     *args, stream=stream)
    torch.cuda.synchronize()
    del ws
    return idx.cpu().numpy().view(np.uint32), out.cpu().numpy()


@pytest.mark.parametrize("name", ["unique", "ties"])
@pytest.mark.parametrize("desc", [False, True])
def test_sort_int64_golden(ds, golden, name, desc):
    g = golden("sort_int64")
    keys = g[f"{name}_keys"]
    tabs = _tabs(g["block_cuts"], value=np.arange(len(keys), dtype=np.int64), key=keys)
    random.seed(0)
    out = ds.Dataset.from_arrow(tabs).sort("key", descending=desc)
    assert out.num_blocks() == len(tabs)
    got = out.to_arrow_concat()
    assert np.array_equal(got["value"].to_numpy(), g[f"{name}_{'desc' if desc else 'asc'}_value_order"])
    assert got.schema == tabs[0].schema
    # a key value never spans two output blocks (sort.py:118-119 of the reference)
    seen = set()
    for t in out.to_arrow():
        ks = set(t["key"].to_pylist()) if t.num_rows else set()
        assert not (ks & seen)
        seen |= ks


@pytest.mark.parametrize("name", ["uniform", "ties"])
@pytest.mark.parametrize("desc", [False, True])
def test_sort_graysort_golden(ds, golden, name, desc):
    g = golden("sort_graysort")
    keys = g[f"{name}_keys"]
    base = O.graysort_table(len(keys), seed=11)
    t = pa.table({"key": O.fsb_from_numpy(keys), "payload": base["payload"]})
    tabs = [t.slice(a, b - a) for a, b in g["block_cuts"]]
    got = ds.Dataset.from_arrow(tabs).sort("key", descending=desc).to_arrow_concat()
    rowid = O.fsb_to_numpy(got["payload"])[:, :8].copy().view("<u8").ravel()
    assert np.array_equal(rowid, g[f"{name}_{'desc' if desc else 'asc'}_rowid_order"])
    assert got.equals(O.global_stable_sort(tabs, "key", desc))


@pytest.mark.parametrize("desc", [False, True])
def test_sort_and_partition_golden(ds, golden, desc):
    g = golden("partition")
    keys = g["keys"]
    t = pa.table({"key": keys, "value": np.arange(len(keys), dtype=np.int64)})
    tag = "desc" if desc else "asc"
    key = [("key", "descending" if desc else "ascending")]
    parts = ds.BlockAccessor.for_block(t).sort_and_partition(list(g[f"{tag}_boundaries"]), key, desc)
    assert [p.num_rows for p in parts] == list(g[f"{tag}_sizes"])
    cat = pa.concat_tables([p.to_arrow() for p in parts])
    assert np.array_equal(cat["value"].to_numpy(), g[f"{tag}_value_order"])
    # the unsorted stable partition used by the device plan obeys the same rule
    parts_u = ds.BlockAccessor.for_block(t).partition_unsorted(list(g[f"{tag}_boundaries"]), key, desc)
    assert [p.num_rows for p in parts_u] == list(g[f"{tag}_sizes"])
    for pu, ps in zip(parts_u, parts):
        if pu.num_rows:
            v = pu.to_arrow()["value"].to_numpy()
            assert np.all(np.diff(v) > 0)  # input order kept inside a partition
            assert np.array_equal(np.sort(v), np.sort(ps.to_arrow()["value"].to_numpy()))


def test_merge_sorted_blocks_golden(ds, golden):
    g = golden("partition")
    keys = g["keys"]
    t = pa.table({"key": keys, "value": np.arange(len(keys), dtype=np.int64)})
    key = [("key", "ascending")]
    blocks = [
        ds.BlockAccessor.for_block(t.slice(a, b - a)).sort_and_partition([], key, False)[0]
        for a, b in g["merge_cuts"]
    ]
    merged, meta = ds.DeviceArrowBlockAccessor.merge_sorted_blocks(blocks, key, False)
    assert np.array_equal(merged.to_arrow()["value"].to_numpy(), g["merge_value_order"])
    assert meta.num_rows == len(keys) and meta.size_bytes == 16 * len(keys)


def test_empty_block_contracts(ds):
    # test_sort.py:110-127 of the reference
    empty = ds.DeviceArrowBlock.empty()
    acc = ds.BlockAccessor.for_block(empty)
    assert acc.sample(10, "A").num_rows == 0
    parts = acc.sort_and_partition([1, 5, 10], "A", descending=False)
    assert len(parts) == 4 and all(p.num_rows == 0 for p in parts)
    assert acc.merge_sorted_blocks([empty], "A", False)[0].num_rows == 0
    # empty dataset: boundaries are [None]*(R-1) and the sort has no rows
    e = pa.table({"value": pa.array([], pa.int64())})
    assert ds.sample_boundaries([ds.DeviceArrowBlock.from_arrow(e)] * 3, [("value", "ascending")], 3) == [None, None]
    out = ds.Dataset.from_arrow([e, e, e]).sort("value")
    assert out.count() == 0
    assert ds.Dataset.from_arrow([]).sort("value").count() == 0


def test_sort_with_empty_and_ragged_blocks(ds):
    rng = np.random.default_rng(3)
    sizes = [0, 1, 4097, 0, 33, 70000]
    tabs, base = [], 0
    for s in sizes:
        tabs.append(pa.table({"key": rng.integers(-50, 50, s, dtype=np.int64),
                              "value": np.arange(base, base + s, dtype=np.int64)}))
        base += s
    for desc in (False, True):
        got = ds.Dataset.from_arrow(tabs).sort("key", descending=desc).to_arrow_concat()
        assert got.equals(O.global_stable_sort(tabs, "key", desc))


def test_reference_test_sort_arrow(ds):
    # test_sort.py:55-97 of the reference (int key; the string column is carried as
    # fixed_size_binary(3) payload and also used as a memcmp key)
    for num_items, parallelism in ((100, 1), (1000, 4)):
        a = list(reversed(range(num_items)))
        b = [f"{x:03}".encode() for x in range(num_items)]
        shard = int(np.ceil(num_items / parallelism))
        tabs = []
        for off in range(0, num_items, shard):
            tabs.append(pa.table({"a": pa.array(a[off:off + shard], pa.int64()),
                                  "b": pa.array(b[off:off + shard], pa.binary(3))}))
        d = ds.Dataset.from_arrow(tabs)

        def rows(x):
            return [(r["a"], r["b"]) for r in x.iter_rows()]

        assert rows(d.sort(key="a")) == list(zip(reversed(a), reversed(b)))
        assert len([n for n in d.sort(key="a")._block_num_rows() if n > 0]) == parallelism
        assert rows(d.sort(key="b")) == list(zip(a, b))
        assert rows(d.sort(key="a", descending=True)) == list(zip(a, b))


def test_same_key_to_same_block(ds):
    # test_sort.py:37-52 of the reference
    tabs = [pa.table({"x": np.ones(1, dtype=np.int64)}) for _ in range(100)]
    out = ds.Dataset.from_arrow(tabs).sort("x")
    nr = out._block_num_rows()
    assert len(nr) == 100 and sum(1 for x in nr if x > 0) == 1 and sum(nr) == 100


@pytest.mark.parametrize("kind", ["i64", "u64", "f64"])
@pytest.mark.parametrize("desc", [False, True])
@pytest.mark.parametrize("n", [1, 31, 4096, 4097, 100_003])
def test_argsort_c_abi_64bit(ds, kind, desc, n):
    rng = np.random.default_rng(n)
    if kind == "i64":
        col = rng.integers(-(2**63), 2**63 - 1, n, dtype=np.int64)
        col[::7] = col[0]
    elif kind == "u64":
        col = rng.integers(0, 2**64 - 1, n, dtype=np.uint64)
        col[::5] = 2**63
    else:
        col = rng.standard_normal(n) * 1e3
        col[::11] = 0.0
        col[1::11] = -0.0
        if n > 40:
            col[3], col[17], col[29], col[5] = np.nan, np.inf, -np.inf, -np.nan
    idx, srt = _argsort_c(ds, col, kind, desc)
    order = "descending" if desc else "ascending"
    want = pa.compute.sort_indices(pa.table({"k": col}), sort_keys=[("k", order)]).to_numpy()
    assert np.array_equal(idx.astype(np.int64), want)
    assert np.array_equal(srt.view(col.dtype), col[want], equal_nan=True)
    # the no-host-sync variant gives the same answer
    idx2, _ = _argsort_c(ds, col, kind, desc, options=ds._lib.DS_SORT_NO_PASS_SKIP, want_sorted=False)
    assert np.array_equal(idx, idx2)


@pytest.mark.parametrize("width", [1, 2, 3, 7, 8, 9, 10])
@pytest.mark.parametrize("desc", [False, True])
def test_argsort_c_abi_fsb(ds, width, desc):
    n = 50_001
    rng = np.random.default_rng(width)
    col = rng.integers(0, 256, (n, width), dtype=np.uint8)
    col[:, : max(0, width - 2)] &= 1  # plenty of ties in the leading bytes
    idx, srt = _argsort_c(ds, col, "fsb", desc, width=width)
    arr = O.fsb_from_numpy(col)
    order = "descending" if desc else "ascending"
    want = pa.compute.sort_indices(pa.table({"k": arr}), sort_keys=[("k", order)]).to_numpy()
    assert np.array_equal(idx.astype(np.int64), want)
    assert np.array_equal(srt.reshape(n, width), col[want])


def test_pass_skipping_constant_bytes(ds):
    # keys in [0, 2^20): five of the eight digit positions are constant and skipped
    rng = np.random.default_rng(0)
    col = rng.integers(0, 1 << 20, 300_000, dtype=np.int64)
    idx, srt = _argsort_c(ds, col, "i64", False)
    want = np.argsort(col, kind="stable")
    assert np.array_equal(idx.astype(np.int64), want)
    # all-equal column: every pass is skipped
    col = np.full(10_000, -7, dtype=np.int64)
    idx, srt = _argsort_c(ds, col, "i64", True)
    assert np.array_equal(idx, np.arange(10_000, dtype=np.uint32))
    assert np.array_equal(srt.view(np.int64), col)


def test_errors(ds):
    # (nulls are supported since the validity masks were added: see tests/test_gpu_nulls.py)
    lib = ds._lib
    need = ctypes.c_size_t(0)
    st = lib.lib.ds_sort_indices_fsb(None, 11, 10, 0, None, None, 0, None, ctypes.byref(need), None)
    assert st == lib.DS_ERR_UNSUPPORTED and b"width" in lib.lib.ds_last_error()
    st = lib.lib.ds_sort_indices_i64(None, 1 << 31, 0, None, None, 0, None, ctypes.byref(need), None)
    assert st == lib.DS_ERR_TOO_LARGE
    t = pa.table({"a": np.arange(4, dtype=np.int64), "b": np.arange(4, dtype=np.int64)})
    with pytest.raises(NotImplementedError):
        ds.BlockAccessor.for_block(t).sort_and_partition([], [("a", "ascending"), ("b", "ascending")], False)
    with pytest.raises(ValueError):
        ds.sample_boundaries([ds.DeviceArrowBlock.from_arrow(t)], [("a", "ascending"), ("b", "ascending")], 2)


def test_large_graysort_properties(ds):
    """4M records (400 MB): sortedness, permutation and stability checked on device
    without the oracle's O(n log n) host sort of the payload."""
    n = 4_000_000
    g = torch.Generator(device="cuda").manual_seed(1)
    keys = torch.randint(0, 256, (n, 10), dtype=torch.uint8, device="cuda", generator=g)
    keys[:, :6] = 0  # 4 random bytes over 4M rows -> many ties
    keys[:, 6] &= 0x0F
    payload = torch.zeros((n, 90), dtype=torch.uint8, device="cuda")
    payload[:, :8] = torch.arange(n, device="cuda", dtype=torch.int64).view(torch.uint8).view(n, 8)
    import collections

    blk = ds.DeviceArrowBlock(collections.OrderedDict(
        key=ds.DeviceColumn(pa.binary(10), keys.reshape(-1)),
        payload=ds.DeviceColumn(pa.binary(90), payload.reshape(-1))), n)
    out = ds.Dataset.from_blocks([blk]).sort("key").get_blocks()[0]
    k = out.columns["key"].data.view(n, 10).to(torch.int64)
    composite = torch.zeros(n, dtype=torch.int64, device="cuda")
    for j in range(6, 10):
        composite = composite * 256 + k[:, j]
    rowid = out.columns["payload"].data.view(n, 90)[:, :8].contiguous().view(torch.int64).view(n)
    assert bool((composite[1:] >= composite[:-1]).all())
    tie = composite[1:] == composite[:-1]
    assert bool((rowid[1:][tie] > rowid[:-1][tie]).all())  # stable
    assert bool((torch.sort(rowid).values == torch.arange(n, device="cuda")).all())
    # key bytes travelled with their rows
    assert bool((keys[rowid] == out.columns["key"].data.view(n, 10)).all())


@pytest.mark.parametrize("cfg", [17, 18, 19, 20, 21, 25, 28])
def test_every_pass_kernel_variant(ds, cfg):
    """all tile shapes / matching and exchange variants of the pass kernel give the same bit-exact order."""
    lib = ds._lib.lib
    old = lib.ds_debug_set_radix_config(cfg)
    try:
        n = 70_001
        rng = np.random.default_rng(cfg)
        col = rng.integers(0, 256, (n, 10), dtype=np.uint8)
        col[:, :7] &= 1
        for desc in (False, True):
            idx, srt = _argsort_c(ds, col, "fsb", desc, width=10)
            order = "descending" if desc else "ascending"
            want = pa.compute.sort_indices(pa.table({"k": O.fsb_from_numpy(col)}), sort_keys=[("k", order)]).to_numpy()
            assert np.array_equal(idx.astype(np.int64), want)
            assert np.array_equal(srt.reshape(n, 10), col[want])
        k64 = rng.integers(-1000, 1000, n, dtype=np.int64)
        idx, _ = _argsort_c(ds, k64, "i64", False, want_sorted=False)
        assert np.array_equal(idx.astype(np.int64), np.argsort(k64, kind="stable"))
        # unaligned sorted_out pointer takes the byte-store decode path
    finally:
        lib.ds_debug_set_radix_config(old)


@pytest.mark.parametrize("width", [1, 2, 3, 8, 10, 16, 17, 18, 24, 33, 90, 100, 250, 1022])
@pytest.mark.parametrize("start_row", [0, 1, 3])
@pytest.mark.parametrize("mode", [0, 1, 2, 7, 8])
def test_gather_rows_c_abi(ds, width, start_row, mode):
    """every granule path of ds_gather_rows and both asynchronous-copy kernels (rows that are not a
    multiple of 16 bytes), with the source column starting at odd byte offsets."""
    lib = ds._lib
    rng = np.random.default_rng(width * 7 + start_row)
    n_src, n_out = 10_007, 12_345
    full = rng.integers(0, 256, (n_src + start_row, width), dtype=np.uint8)
    dev = torch.from_numpy(full.reshape(-1).copy()).cuda()
    src = dev[start_row * width:]
    idx = rng.integers(0, n_src, n_out, dtype=np.int64)
    idx[:3] = [n_src - 1, 0, n_src - 1]  # first / last source rows: guarded edge vectors
    idx_dev = torch.from_numpy(idx.astype(np.int32)).cuda()
    out = torch.zeros(n_out * width + 16, dtype=torch.uint8, device="cuda")
    old = lib.lib.ds_debug_set_gather_mode(mode)
    try:
        lib.call("ds_gather_rows", src.data_ptr(), width, n_src, idx_dev.data_ptr(), n_out,
                 out.data_ptr(), torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
    finally:
        lib.lib.ds_debug_set_gather_mode(old)
    got = out.cpu().numpy()
    assert np.array_equal(got[: n_out * width].reshape(n_out, width), full[start_row:][idx])
    assert not got[n_out * width:].any()  # nothing written past the output


def test_gather_variants_agree(ds):
    lib = ds._lib
    rng = np.random.default_rng(0)
    n = 300_001
    src = torch.from_numpy(rng.integers(0, 256, n * 90, dtype=np.uint8)).cuda()
    idx = torch.from_numpy(rng.permutation(n).astype(np.int32)).cuda()
    outs = []
    for mode in (1, 2, 7, 8):
        old = lib.lib.ds_debug_set_gather_mode(mode)
        out = torch.empty(n * 90, dtype=torch.uint8, device="cuda")
        lib.call("ds_gather_rows", src.data_ptr(), 90, n, idx.data_ptr(), n, out.data_ptr(),
                 torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        lib.lib.ds_debug_set_gather_mode(old)
        outs.append(out)
    for o in outs[1:]:
        assert bool((outs[0] == o).all())
    assert bool((outs[0].view(n, 90) == src.view(n, 90)[idx.long()]).all())


def _arrow_order(col, kind, desc, width=0):
    order = "descending" if desc else "ascending"
    arr = O.fsb_from_numpy(col) if kind == "fsb" else pa.array(col)
    return pa.compute.sort_indices(pa.table({"k": arr}), sort_keys=[("k", order)]).to_numpy()


@pytest.mark.parametrize("desc", [False, True])
def test_prefix_repair_mode_matches_full_sort(ds, desc):
    """uniform keys take the prefix-sort + run-repair path; the result must equal both the
    all-passes mode and Arrow, including planted short runs, duplicates (stability) and a
    planted run that is too long for the repair kernel (verified fallback)."""
    L = ds._lib
    n = 400_000
    rng = np.random.default_rng(12)
    col = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    # short runs sharing 6 leading bytes, in random low-byte order, with exact duplicates
    for s in range(0, 3000, 10):
        rows = rng.choice(n, 7, replace=False)
        col[rows, :6] = col[rows[0], :6]
        col[rows[3]] = col[rows[1]]
    want = _arrow_order(col, "fsb", desc)
    a, sa = _argsort_c(ds, col, "fsb", desc, width=10)
    b, sb = _argsort_c(ds, col, "fsb", desc, width=10, options=L.DS_SORT_NO_PREFIX_REPAIR)
    assert np.array_equal(a.astype(np.int64), want) and np.array_equal(b, a)
    assert np.array_equal(sa, sb) and np.array_equal(sa.reshape(n, 10), col[want])
    # one run of 500 rows with equal 8-byte prefix: estimate still says "prefix mode"
    rows = rng.choice(n, 500, replace=False)
    col[rows, :8] = 77
    want = _arrow_order(col, "fsb", desc)
    a, sa = _argsort_c(ds, col, "fsb", desc, width=10)
    assert np.array_equal(a.astype(np.int64), want)
    assert np.array_equal(sa.reshape(n, 10), col[want])
    # 64-bit keys take the same path
    k = rng.integers(-(2**63), 2**63 - 1, n, dtype=np.int64)
    k[::1000] = k[0]
    a, _ = _argsort_c(ds, k, "i64", desc, want_sorted=False)
    assert np.array_equal(a.astype(np.int64), _arrow_order(k, "i64", desc))


@pytest.mark.parametrize("n", [65_536, 66_001, 400_003])
@pytest.mark.parametrize("desc", [False, True])
def test_fused_decode_and_listed_repair_equal_separate_kernels(ds, n, desc):
    """prefix mode on 10-byte keys: the fused decode + run-head list + sparse repair pair
    against the separate repair and decode kernels and against Arrow, at sizes that leave a
    ragged last block / a ragged last 4-row group, with runs that straddle block boundaries."""
    L = ds._lib
    rng = np.random.default_rng(n)
    col = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    for s in range(400):                      # runs of 2..9 rows sharing their first 7 bytes
        rows = rng.choice(n, int(rng.integers(2, 10)), replace=False)
        col[rows, :7] = col[rows[0], :7]
    col[rng.choice(n, 12, replace=False)] = col[1]    # 13 equal keys (stability inside a run)
    col[n - 3:, :7] = col[n - 4, :7]                   # a run that ends at the last row
    col[1021:1027, :7] = col[1020, :7]                 # rows next to each other that tie on the prefix
    want = _arrow_order(col, "fsb", desc)
    res = {}
    for fused in (1, 0):
        old = L.lib.ds_debug_set_fused_decode(fused)
        try:
            res[fused] = _argsort_c(ds, col, "fsb", desc, width=10)
        finally:
            L.lib.ds_debug_set_fused_decode(old)
    for fused in (1, 0):
        a, sa = res[fused]
        assert np.array_equal(a.astype(np.int64), want), fused
        assert np.array_equal(sa.reshape(n, 10), col[want]), fused


@pytest.mark.parametrize("n", [70_001, 400_003, 3_000_000])
@pytest.mark.parametrize("desc", [False, True])
def test_hybrid_bucket_plan_equals_arrow(ds, n, desc):
    """10-byte keys through the hybrid plan (two global passes, then one CTA per 16-bit bucket:
    two shared-memory passes + run repair + decode), forced for small n so that buckets are
    tiny, empty, or unaligned in the output; against Arrow and against the all-global plan."""
    L = ds._lib
    rng = np.random.default_rng(n + desc)
    col = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    for s in range(300):                      # ties on the leading 32 bits, and exact duplicates
        rows = rng.choice(n, int(rng.integers(2, 10)), replace=False)
        col[rows, :6] = col[rows[0], :6]
        col[rows[-1]] = col[rows[0]]
    col[rng.choice(n, 1500, replace=False), :2] = (5, 9)        # one well-filled bucket
    col[col[:, 0] == 200, 0] = 201                               # 256 empty buckets in a row
    want = _arrow_order(col, "fsb", desc)
    old = L.lib.ds_debug_set_hybrid(2)
    try:
        L.lib.ds_debug_launch_count.restype = __import__("ctypes").c_uint64
        launches0 = L.lib.ds_debug_launch_count()
        a, sa = _argsort_c(ds, col, "fsb", desc, width=10)
        launches = L.lib.ds_debug_launch_count() - launches0
    finally:
        L.lib.ds_debug_set_hybrid(old)
    b, sb = _argsort_c(ds, col, "fsb", desc, width=10, options=L.DS_SORT_NO_HYBRID)
    assert np.array_equal(a.astype(np.int64), want) and np.array_equal(a, b)
    assert np.array_equal(sa, sb) and np.array_equal(sa.reshape(n, 10), col[want])
    assert launches <= 12, launches            # sample, scan, encode, scan, 2 x (pass [+ ragged]), bounds, local


@pytest.mark.parametrize("desc", [False, True])
def test_hybrid_bucket_plan_64bit_keys(ds, desc):
    """the same plan for keys that are one 64-bit word: int64, uint64, float64 (finite, from
    random bit patterns so that the leading bytes spread) and the Philox permutation keys."""
    L = ds._lib
    n = 300_007
    rng = np.random.default_rng(21 + desc)
    bits = rng.integers(-(2**63), 2**63 - 1, n, dtype=np.int64)
    for g in range(300):                                   # groups of equal keys: stability
        rows = rng.choice(n, int(rng.integers(2, 9)), replace=False)
        bits[rows] = bits[rows[0]]
    f = bits.view(np.float64).copy()
    bad = ~np.isfinite(f)
    f[bad] = np.arange(bad.sum()) + 0.5                    # distinct, so no long run of equal keys
    cases = {"i64": bits, "u64": bits.view(np.uint64), "f64": f}
    stream = torch.cuda.current_stream().cuda_stream
    old = L.lib.ds_debug_set_hybrid(2)
    try:
        for kind, col in cases.items():
            a, sa = _argsort_c(ds, col, kind, desc)
            b, sb = _argsort_c(ds, col, kind, desc, options=L.DS_SORT_NO_HYBRID)
            assert np.array_equal(a.astype(np.int64), _arrow_order(col, kind, desc)), kind
            assert np.array_equal(a, b) and np.array_equal(sa, sb), kind
            assert np.array_equal(sa.view(col.dtype), col[a.astype(np.int64)]), kind
        perms = []
        for mode in (2, 0):
            L.lib.ds_debug_set_hybrid(mode)
            perm = torch.empty(n, dtype=torch.int32, device="cuda")
            ws = L.call_ws("ds_random_permutation", lambda nb: torch.empty(nb, dtype=torch.uint8, device="cuda"),
                           1234, 7, n, perm.data_ptr(), stream=stream)
            torch.cuda.synchronize()
            del ws
            perms.append(perm.cpu().numpy())
        assert np.array_equal(perms[0], perms[1])
        assert np.array_equal(np.sort(perms[0]), np.arange(n))
    finally:
        L.lib.ds_debug_set_hybrid(old)


def test_hybrid_bucket_plan_falls_back(ds):
    """a bucket that does not fit a CTA, or a long run inside a bucket, sends the call back to
    the all-global plan; the result is the same."""
    L = ds._lib
    n = 500_000
    rng = np.random.default_rng(3)
    for kind in ("bucket", "run"):
        col = rng.integers(0, 256, (n, 10), dtype=np.uint8)
        rows = rng.choice(n, 3000 if kind == "bucket" else 40, replace=False)
        col[rows, :2 if kind == "bucket" else 5] = 77
        want = _arrow_order(col, "fsb", False)
        old = L.lib.ds_debug_set_hybrid(2)
        try:
            a, sa = _argsort_c(ds, col, "fsb", False, width=10)
        finally:
            L.lib.ds_debug_set_hybrid(old)
        assert np.array_equal(a.astype(np.int64), want), kind
        assert np.array_equal(sa.reshape(n, 10), col[want]), kind


def _subrange_fsb(rng, n, lo16, hi16):
    """10-byte keys whose leading 16 bits lie in [lo16, hi16): one rank's share of a
    range-partitioned dataset; the other bytes are random."""
    col = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    top = rng.integers(lo16, hi16, n)
    col[:, 0] = top >> 8
    col[:, 1] = top & 255
    return col


@pytest.mark.parametrize("desc", [False, True])
@pytest.mark.parametrize("lo16,hi16", [(0x2000, 0x4000), (0xE000, 0x10000), (0, 0x1555), (0x7123, 0x7124), (0x5555, 0xAAAB)])
def test_hybrid_plan_on_a_subrange_of_the_key_space(ds, desc, lo16, hi16):
    """keys that fill only part of the key space (any part, not a power of two wide) still take
    the hybrid plan: the bucket digits come from the keys' position inside the sampled range
    (KeyXform in radix_sort.cu).  Same order as Arrow and as the all-global plan, with ties on
    the leading bytes and exact duplicates in the data."""
    L = ds._lib
    n = 300_011
    rng = np.random.default_rng(lo16 + desc)
    col = _subrange_fsb(rng, n, lo16, hi16)
    for s in range(300):
        rows = rng.choice(n, int(rng.integers(2, 10)), replace=False)
        col[rows, :7] = col[rows[0], :7]
        col[rows[-1]] = col[rows[0]]
    want = _arrow_order(col, "fsb", desc)
    old = L.lib.ds_debug_set_hybrid(2)
    try:
        a, sa = _argsort_c(ds, col, "fsb", desc, width=10)
        plan = L.lib.ds_debug_last_sort_plan()
        off = L.lib.ds_debug_set_subrange(0)
        c, sc = _argsort_c(ds, col, "fsb", desc, width=10)
        plan_off = L.lib.ds_debug_last_sort_plan()
        L.lib.ds_debug_set_subrange(off)
    finally:
        L.lib.ds_debug_set_hybrid(old)
    b, sb = _argsort_c(ds, col, "fsb", desc, width=10, options=L.DS_SORT_NO_HYBRID)
    assert plan == 3 and plan_off in (0, 1), (plan, plan_off)
    assert np.array_equal(a.astype(np.int64), want) and np.array_equal(a, b) and np.array_equal(a, c)
    assert np.array_equal(sa, sb) and np.array_equal(sa, sc) and np.array_equal(sa.reshape(n, 10), col[want])


@pytest.mark.parametrize("desc", [False, True])
def test_hybrid_plan_on_a_subrange_64bit_keys(ds, desc):
    """the same for one-word keys: ids clustered in a window, time stamps, negative ranges that
    cross zero, doubles in a narrow interval."""
    L = ds._lib
    n = 250_003
    rng = np.random.default_rng(77 + desc)
    cases = {
        "i64": [rng.integers(10**15, 10**15 + 3 * 10**12, n, dtype=np.int64),
                rng.integers(-(10**9), 10**9, n, dtype=np.int64)],
        "u64": [rng.integers(2**63 - 10**13, 2**63 + 10**13, n, dtype=np.uint64)],
        "f64": [rng.uniform(1.0e6, 1.0e6 + 1.0, n)],   # ~2^33 distinct values in range
    }
    old = L.lib.ds_debug_set_hybrid(2)
    try:
        for kind, cols in cases.items():
            for col in cols:
                for g in range(200):
                    rows = rng.choice(n, int(rng.integers(2, 9)), replace=False)
                    col[rows] = col[rows[0]]
                a, sa = _argsort_c(ds, col, kind, desc)
                plan = L.lib.ds_debug_last_sort_plan()
                b, sb = _argsort_c(ds, col, kind, desc, options=L.DS_SORT_NO_HYBRID)
                assert plan == 3, (kind, plan)
                assert np.array_equal(a.astype(np.int64), _arrow_order(col, kind, desc)), kind
                assert np.array_equal(a, b) and np.array_equal(sa, sb), kind
                assert np.array_equal(sa.view(col.dtype), col[a.astype(np.int64)]), kind
    finally:
        L.lib.ds_debug_set_hybrid(old)


def test_hybrid_plan_on_a_subrange_falls_back(ds):
    """keys outside the sampled range (the sample takes every n/2^20-th row), a sub-range filled
    unevenly, and a range with too few distinct values: the plan is refused up front or the call
    falls back after the range check; the order is Arrow's either way."""
    L = ds._lib
    rng = np.random.default_rng(9)
    old = L.lib.ds_debug_set_hybrid(2)
    try:
        # 1. outliers the sample cannot see (n > 2^21: stride 2, odd rows are never sampled)
        n = (1 << 21) + 5
        col = _subrange_fsb(rng, n, 0x4000, 0x6000)
        col[1, :2] = (0xFF, 0xFF)
        col[12345, :2] = (0, 0)
        a, sa = _argsort_c(ds, col, "fsb", False, width=10)
        assert L.lib.ds_debug_last_sort_plan() != 3
        want = _arrow_order(col, "fsb", False)
        assert np.array_equal(a.astype(np.int64), want) and np.array_equal(sa.reshape(n, 10), col[want])
        # 2. a range filled unevenly (half of the rows in 1/64 of it): refused by the second sample
        n = 300_000
        col = _subrange_fsb(rng, n, 0x4000, 0x8000)
        col[: n // 2, 0] = 0x40
        col[: n // 2, 1] = rng.integers(0, 256, n // 2)
        a, sa = _argsort_c(ds, col, "fsb", True, width=10)
        assert L.lib.ds_debug_last_sort_plan() != 3
        want = _arrow_order(col, "fsb", True)
        assert np.array_equal(a.astype(np.int64), want) and np.array_equal(sa.reshape(n, 10), col[want])
        # 3. 64-bit keys with 1000 distinct values: G collides, refused
        k = rng.integers(5000, 6000, n, dtype=np.int64)
        a, sa = _argsort_c(ds, k, "i64", False)
        assert L.lib.ds_debug_last_sort_plan() != 3
        assert np.array_equal(a.astype(np.int64), _arrow_order(k, "i64", False))
        assert np.array_equal(sa.view(np.int64), np.sort(k))
    finally:
        L.lib.ds_debug_set_hybrid(old)


@pytest.mark.parametrize("width", [1, 8, 10, 90, 100])
def test_scatter_rows_to_parts_c_abi(ds, width):
    """the fused map-side slicing kernel against numpy (local destinations; over NVLink it is
    exercised by tests/test_gpu_distributed.py with DS_EXCHANGE=fused)."""
    import ctypes

    lib = ds._lib
    rng = np.random.default_rng(width)
    n, nparts = 50_003, 5
    src = rng.integers(0, 256, (n, width), dtype=np.uint8)
    bucket = rng.integers(0, nparts, n)
    bucket[bucket == 3] = 2  # an empty partition
    perm = np.argsort(bucket, kind="stable").astype(np.int32)
    counts = np.bincount(bucket, minlength=nparts)
    offs = np.concatenate([[0], np.cumsum(counts)])
    dsrc = torch.from_numpy(src.reshape(-1).copy()).cuda()
    dperm = torch.from_numpy(perm).cuda()
    outs = [torch.zeros(int(c) * width + 32, dtype=torch.uint8, device="cuda") for c in counts]
    bases = (ctypes.c_void_p * nparts)(*[o.data_ptr() + 16 for o in outs])
    offs_c = (ctypes.c_uint64 * (nparts + 1))(*[int(x) for x in offs])
    lib.call("ds_scatter_rows_to_parts", dsrc.data_ptr(), width, n, dperm.data_ptr(), n, nparts, offs_c,
             bases, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    for d in range(nparts):
        got = outs[d].cpu().numpy()
        want = src[perm[offs[d]:offs[d + 1]]].reshape(-1)
        assert np.array_equal(got[16:16 + len(want)], want)
        assert not got[:16].any() and not got[16 + len(want):].any()
