"""GPU parity for variable-length (string / binary) keys and columns, and for groupby keys
that are not int64.  Checked against tests/golden/varlen.npz (recorded from the reference's
sort_impl / groupby op), against pyarrow's sort_indices / take (the primitives the reference
calls: transform_pyarrow.py:12-16,46-60) and against Python/numpy for the C ABI calls."""
import ctypes

import numpy as np
import pandas as pd
import pyarrow as pa
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import datasets_oracle as O  # noqa: E402


@pytest.fixture(scope="module")
def m():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ray_memory_management_b200 as mod

    return mod


def _unpack_rows(offsets, data):
    raw = data.tobytes()
    return [raw[offsets[i]:offsets[i + 1]] for i in range(len(offsets) - 1)]


def _random_rows(rng, n, max_len=20, alphabet=b"ab\x00\xffz", null_frac=0.0):
    alpha = np.frombuffer(alphabet, dtype=np.uint8)
    rows = []
    for _ in range(n):
        r = alpha[rng.integers(0, len(alpha), int(rng.integers(0, max_len + 1)))].tobytes()
        if rng.random() < 0.3:
            r = b"prefix--prefix--" + r[:3]
        rows.append(None if rng.random() < null_frac else r)
    return rows


# ------------------------------------------------------------------ import / export
@pytest.mark.parametrize("typ", [pa.binary(), pa.string(), pa.large_binary(), pa.large_string()])
def test_round_trip_types_nulls_slices(m, typ):
    is_str = pa.types.is_string(typ) or pa.types.is_large_string(typ)
    vals = ["", "a", None, "héllo wörld", "ab", "ab\x00", None, "x" * 300, ""]
    if not is_str:
        vals = [None if v is None else v.encode() for v in vals]
    t = pa.table({"s": pa.array(vals, typ), "i": np.arange(len(vals), dtype=np.int64)})
    blk = m.DeviceArrowBlock.from_arrow(t)
    assert blk.to_arrow().equals(t) and blk.schema == t.schema
    assert blk.nbytes >= sum(len(v if isinstance(v, bytes) else v.encode()) for v in vals if v is not None)
    sl = t.slice(2, 5)                                    # arr.offset != 0 on import
    assert m.DeviceArrowBlock.from_arrow(sl).to_arrow().equals(sl)
    assert blk.slice(3, 8).to_arrow().equals(t.slice(3, 5))   # device slice = view of the bytes
    assert blk.slice(4, 4).to_arrow().equals(t.slice(4, 0))
    cat = m.DeviceArrowBlock.concat([blk.slice(5, 9), blk.slice(0, 2), blk])
    assert cat.to_arrow().equals(pa.concat_tables([t.slice(5, 4), t.slice(0, 2), t]).combine_chunks())
    empty = pa.table({"s": pa.array([], typ)})
    assert m.DeviceArrowBlock.from_arrow(empty).to_arrow().equals(empty)
    acc = m.BlockAccessor.for_block(blk)
    idx = [8, 0, 3, 3, 2, 7]
    assert acc.take(idx).to_arrow().equals(t.take(idx))
    assert acc.slice(1, 6, copy=True).to_arrow().equals(t.slice(1, 5))


# ------------------------------------------------------------------ sort
@pytest.mark.parametrize("fuse", [True, False])
@pytest.mark.parametrize("desc", [False, True])
def test_sort_binary_golden(m, golden, desc, fuse):
    g = golden("varlen")
    rows = _unpack_rows(g["bin_offsets"], g["bin_data"])
    tabs = [pa.table({"key": pa.array(rows[a:b], pa.binary()), "row": np.arange(a, b, dtype=np.int64)})
            for a, b in g["block_cuts"]]
    ctx = m.DatasetContext.get_current()
    old, ctx.fuse_device_blocks = ctx.fuse_device_blocks, fuse
    try:
        out = m.Dataset.from_arrow(tabs).sort("key", descending=desc)
    finally:
        ctx.fuse_device_blocks = old
    assert out.num_blocks() == len(tabs)
    got = out.to_arrow_concat()
    assert np.array_equal(got["row"].to_numpy(), g[f"plain_{'desc' if desc else 'asc'}_row_order"])
    assert got["key"].to_pylist() == [rows[i] for i in got["row"].to_numpy()]
    # a key value never spans two output blocks (sort.py:118-119)
    seen = set()
    for t in out.to_arrow():
        ks = set(t["key"].to_pylist()) if t.num_rows else set()
        assert not (ks & seen)
        seen |= ks


def test_sort_string_golden(m, golden):
    g = golden("varlen")
    words = [r.decode() for r in _unpack_rows(g["str_offsets"], g["str_data"])]
    tabs = [pa.table({"key": pa.array(words[a:b], pa.string()), "row": np.arange(a, b, dtype=np.int64),
                      "v": g["v"][a:b]}) for a, b in g["block_cuts"]]
    got = m.Dataset.from_arrow(tabs).sort("key").to_arrow_concat()
    assert np.array_equal(got["row"].to_numpy(), g["str_asc_row_order"])
    assert got.schema.field("key").type == pa.string()


@pytest.mark.parametrize("typ", [pa.binary(), pa.large_string()])
@pytest.mark.parametrize("desc", [False, True])
def test_sort_with_null_keys_vs_arrow(m, typ, desc):
    rng = np.random.default_rng(5)
    n = 30_000
    if pa.types.is_large_string(typ):
        rows = [None if r is None else r.decode("latin-1") for r in _random_rows(rng, n, 12, b"abcxyz", 0.1)]
    else:
        rows = _random_rows(rng, n, 27, null_frac=0.1)
    cuts = [0, 1, 9000, 9000, 20_000, n]                       # a 1-row and an empty block
    tabs = [pa.table({"key": pa.array(rows[a:b], typ), "row": np.arange(a, b, dtype=np.int64)})
            for a, b in zip(cuts, cuts[1:])]
    want = O.global_stable_sort(tabs, "key", desc)
    got = m.Dataset.from_arrow(tabs).sort("key", descending=desc).to_arrow_concat()
    assert np.array_equal(got["row"].to_numpy(), want["row"].to_numpy())
    assert got["key"].to_pylist() == want["key"].to_pylist()      # nulls last in both directions


def test_sort_edge_cases(m):
    # all-empty strings, a single row, all-equal long keys, keys differing only past byte 64
    for rows in ([b""] * 50, [b"only"], [b"k" * 100] * 40,
                 [b"p" * 64 + bytes([i % 7]) for i in range(500)], [b"", b"\x00", b"\x00\x00", b""]):
        t = pa.table({"key": pa.array(rows, pa.binary()), "row": np.arange(len(rows), dtype=np.int64)})
        for desc in (False, True):
            want = O.sort_table(t, "key", desc)
            got = m.Dataset.from_arrow([t]).sort("key", descending=desc).to_arrow_concat()
            assert got.equals(want), (rows[:3], desc)


# ------------------------------------------------------------------ shuffle / repartition with string payload
def test_shuffle_and_repartition_carry_strings(m):
    rng = np.random.default_rng(2)
    n = 20_000
    words = [("w%d-" % i) + "x" * int(rng.integers(0, 9)) for i in range(n)]
    cuts = [0, 5000, 12_000, n]
    tabs = [pa.table({"id": np.arange(a, b, dtype=np.int64), "s": pa.array(words[a:b], pa.string())})
            for a, b in zip(cuts, cuts[1:])]
    a = m.Dataset.from_arrow(tabs).random_shuffle(seed=4).to_arrow_concat()
    b = m.Dataset.from_arrow(tabs).random_shuffle(seed=4).to_arrow_concat()
    assert a.equals(b)
    ids = a["id"].to_numpy()
    assert np.array_equal(np.sort(ids), np.arange(n)) and not np.array_equal(ids, np.arange(n))
    assert a["s"].to_pylist() == [words[i] for i in ids]          # strings travel with their rows
    rp = m.Dataset.from_arrow(tabs).repartition(5)
    want = O.repartition_dataset(tabs, 5)
    assert [t.num_rows for t in rp.to_arrow()] == [t.num_rows for t in want]
    assert rp.to_arrow_concat().equals(pa.concat_tables(want).combine_chunks())


# ------------------------------------------------------------------ groupby on non-int64 keys
def test_groupby_string_key_golden(m, golden):
    g = golden("varlen")
    words = [r.decode() for r in _unpack_rows(g["str_offsets"], g["str_data"])]
    tabs = [pa.table({"key": pa.array(words[a:b], pa.string()), "row": np.arange(a, b, dtype=np.int64),
                      "v": g["v"][a:b]}) for a, b in g["block_cuts"]]
    out = m.Dataset.from_arrow(tabs).groupby("key").aggregate(
        m.Count(), m.Sum("v"), m.Min("v"), m.Mean("v")).to_arrow_concat()
    assert out.column_names == list(g["gb_names"])
    assert out["key"].to_pylist() == [r.decode() for r in _unpack_rows(g["gb_key_offsets"], g["gb_key_data"])]
    for name in out.column_names[1:]:
        assert np.allclose(out[name].to_numpy(), g[f"gb::{name}"], rtol=1e-12, atol=0), name


@pytest.mark.parametrize("num_parts", [1, 6])
def test_groupby_fsb_f64_and_binary_keys_vs_pandas(m, num_parts):
    rng = np.random.default_rng(8)
    n = 12_000
    kid = rng.integers(0, 300, n)
    v = rng.normal(size=n)
    variants = {
        "fsb": pa.array([b"k%04d" % k for k in kid], pa.binary(5)),
        "f64": pa.array(kid.astype(np.float64) / 4 - 10.0),
        "bin": pa.array([b"key-" + bytes([65 + k % 26]) * (k % 11) + b"%d" % k for k in kid], pa.binary()),
    }
    for tag, keys in variants.items():
        t = pa.table({"k": keys, "v": v})
        ds = m.from_arrow(t).repartition(num_parts)
        got = ds.groupby("k").aggregate(m.Count(), m.Sum("v"), m.Max("v")).to_arrow_concat()
        df = pd.DataFrame({"k": keys.to_pylist(), "v": v})
        exp = df.groupby("k", sort=True).agg(c=("v", "size"), s=("v", "sum"), mx=("v", "max")).reset_index()
        assert got["k"].to_pylist() == exp.k.tolist(), tag
        assert got["count()"].to_pylist() == exp.c.tolist(), tag
        assert np.allclose(got["sum(v)"].to_numpy(), exp.s.to_numpy(), rtol=1e-9, atol=1e-9), tag
        assert np.array_equal(got["max(v)"].to_numpy(), exp.mx.to_numpy()), tag
    # map_groups sees whole groups of a string key, in key order
    t = pa.table({"k": pa.array(["b", "a", "b", "", "a", "b"]), "v": np.arange(6, dtype=np.int64)})
    out = m.from_arrow(t).groupby("k").map_groups(
        lambda g: pd.DataFrame({"k": [g.k.iloc[0]], "vs": [",".join(map(str, g.v))]})).to_arrow_concat()
    assert out.to_pydict() == {"k": ["", "a", "b"], "vs": ["3", "1,4", "0,2,5"]}


# ------------------------------------------------------------------ C ABI, called directly
def _dev(a):
    return torch.from_numpy(np.array(a, copy=True)).to("cuda")


def test_cabi_take_chunks_searchsorted_group_ids():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from ray_memory_management_b200 import _lib

    rng = np.random.default_rng(3)
    n = 5000
    rows = [r for r in _random_rows(rng, n, 30)]
    lens = np.array([len(r) for r in rows], dtype=np.int64)
    offs = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lens, out=offs[1:])
    offs += 17                                              # offsets need not start at 0
    data = np.frombuffer(b"\xee" * 17 + b"".join(rows), dtype=np.uint8)
    d_off, d_data = _dev(offs), _dev(data)
    stream = torch.cuda.current_stream().cuda_stream
    alloc = lambda nbytes: torch.empty(nbytes, dtype=torch.uint8, device="cuda")  # noqa: E731

    # take
    idx = rng.integers(0, n, 3000).astype(np.uint32)
    d_idx = _dev(idx.view(np.int32))
    out_off = torch.empty(len(idx) + 1, dtype=torch.int64, device="cuda")
    _lib.call_ws("ds_varlen_take_offsets", alloc, d_off.data_ptr(), n, d_idx.data_ptr(), len(idx),
                 out_off.data_ptr(), stream=stream)
    want_off = np.concatenate([[0], np.cumsum(lens[idx])])
    assert np.array_equal(out_off.cpu().numpy(), want_off)
    out = torch.full((int(want_off[-1]) + 8,), 0xAB, dtype=torch.uint8, device="cuda")  # 8 guard bytes
    _lib.call("ds_varlen_take_bytes", d_off.data_ptr(), d_data.data_ptr(), d_idx.data_ptr(), len(idx),
              out_off.data_ptr(), out.data_ptr(), stream)
    got = out.cpu().numpy()
    assert got[: want_off[-1]].tobytes() == b"".join(rows[i] for i in idx)
    assert (got[want_off[-1]:] == 0xAB).all()

    # key parts
    keys = torch.empty(n, dtype=torch.int64, device="cuda")
    mx = torch.empty(1, dtype=torch.int64, device="cuda")
    _lib.call("ds_varlen_key_lengths", d_off.data_ptr(), None, n, keys.data_ptr(), mx.data_ptr(), stream)
    assert np.array_equal(keys.cpu().numpy(), lens) and int(mx.item()) == lens.max()
    for byte0 in (0, 8, 24):
        _lib.call("ds_varlen_key_chunk", d_off.data_ptr(), d_data.data_ptr(), None, n, byte0, keys.data_ptr(), stream)
        want = np.array([int.from_bytes(r[byte0:byte0 + 8].ljust(8, b"\0"), "big") for r in rows], dtype=np.uint64)
        assert np.array_equal(keys.cpu().numpy().view(np.uint64), want)

    # searchsorted on the sorted rows, both directions
    bounds = [b"", b"a", b"ab\x00", b"prefix--prefix--", b"\xff\xff\xff", rows[10], rows[11]]
    b_off = np.concatenate([[0], np.cumsum([len(b) for b in bounds])]).astype(np.int64)
    d_boff, d_bdata = _dev(b_off), _dev(np.frombuffer(b"".join(bounds), dtype=np.uint8))
    res = torch.empty(len(bounds), dtype=torch.int64, device="cuda")
    for desc in (False, True):
        srt = sorted(rows, reverse=desc)
        s_off = np.concatenate([[0], np.cumsum([len(r) for r in srt])]).astype(np.int64)
        ds_off, ds_data = _dev(s_off), _dev(np.frombuffer(b"".join(srt), dtype=np.uint8))
        _lib.call("ds_searchsorted_varlen", ds_off.data_ptr(), ds_data.data_ptr(), n, d_boff.data_ptr(),
                  d_bdata.data_ptr(), len(bounds), int(desc), res.data_ptr(), stream)
        want = [sum(1 for r in srt if (r > b if desc else r < b)) for b in bounds]
        assert res.cpu().tolist() == want
        # group ids of the sorted rows
        gid = torch.empty(n, dtype=torch.int64, device="cuda")
        _lib.call_ws("ds_group_ids_varlen", alloc, ds_off.data_ptr(), ds_data.data_ptr(), n, gid.data_ptr(), stream=stream)
        want_gid = np.concatenate([[0], np.cumsum([srt[i] != srt[i - 1] for i in range(1, n)])])
        assert np.array_equal(gid.cpu().numpy(), want_gid)

    # fixed-width group ids: odd width (bytewise), 8-byte words, float64 with -0.0 / NaN
    fsb = np.sort(rng.integers(0, 40, 4000)).astype(np.uint8)
    rows3 = np.stack([fsb // 8, fsb % 8, np.zeros_like(fsb)], axis=1)
    gid = torch.empty(4000, dtype=torch.int64, device="cuda")
    d_rows3 = _dev(rows3)  # keep the inputs referenced: the workspace is allocated inside the call
    _lib.call_ws("ds_group_ids_fixed", alloc, d_rows3.data_ptr(), 3, 0, 4000, gid.data_ptr(), stream=stream)
    assert np.array_equal(gid.cpu().numpy(), np.concatenate([[0], np.cumsum(fsb[1:] != fsb[:-1])]))
    i64 = np.sort(rng.integers(-5, 5, 3000))
    gid = torch.empty(3000, dtype=torch.int64, device="cuda")
    d_i64 = _dev(i64)
    _lib.call_ws("ds_group_ids_fixed", alloc, d_i64.data_ptr(), 8, 0, 3000, gid.data_ptr(), stream=stream)
    assert np.array_equal(gid.cpu().numpy(), np.concatenate([[0], np.cumsum(i64[1:] != i64[:-1])]))
    f = np.array([-1.0, -0.0, 0.0, 0.0, 2.5, 2.5, np.nan, np.nan])
    gid = torch.empty(8, dtype=torch.int64, device="cuda")
    d_f = _dev(f)
    _lib.call_ws("ds_group_ids_fixed", alloc, d_f.data_ptr(), 8, 1, 8, gid.data_ptr(), stream=stream)
    assert gid.cpu().tolist() == [0, 1, 1, 1, 2, 2, 3, 4]
    need = ctypes.c_size_t(0)
    assert _lib.lib.ds_group_ids_fixed(None, 0, 0, 8, None, None, ctypes.byref(need), None) == _lib.DS_ERR_INVALID_ARG
