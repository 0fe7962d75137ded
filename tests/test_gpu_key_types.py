"""GPU parity for key and value columns of the numeric types that reach the kernels through a
64-bit stand-in (ds_widen_keys): narrow and unsigned integers, float32/float16, temporal types.
Oracle: pyarrow's sort_indices via oracle.global_stable_sort (the primitive the reference sorts
with, transform_pyarrow.py:12-16) and pandas groupby."""
import numpy as np
import pandas as pd
import pyarrow as pa
import pyarrow.compute  # noqa: F401
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


def _keys(typ, rng, n):
    T = pa.types
    if T.is_floating(typ):
        a = rng.normal(size=n).astype(np.float32 if T.is_float32(typ) else np.float16)
        a[rng.integers(0, n, n // 50)] = np.nan
        a[rng.integers(0, n, n // 50)] = 0.0
        a[rng.integers(0, n, n // 50)] = -0.0
        return pa.array(a, type=typ)
    if T.is_uint64(typ):
        a = rng.integers(0, 2**64, n, dtype=np.uint64)
        a[: n // 4] = rng.integers(2**63 - 50, 2**63 + 50, n // 4, dtype=np.uint64)  # around the sign bit
        return pa.array(a, type=typ)
    if T.is_temporal(typ):
        if typ.bit_width == 32:
            a = rng.integers(0, 80_000 if T.is_time32(typ) else 40_000, n).astype(np.int32)
            return pa.array(a).cast(typ)
        a = rng.integers(-10**9, 10**9, n).astype(np.int64)
        if T.is_date64(typ):
            a = (a % 40_000) * 86_400_000
        if T.is_time64(typ):
            a = np.abs(a)
        return pa.array(a).cast(typ)
    info = np.iinfo(typ.to_pandas_dtype())
    a = rng.integers(max(info.min, -300), min(info.max, 300) + 1, n).astype(typ.to_pandas_dtype())
    a[:4] = [info.min, info.max, info.min, info.max]
    return pa.array(a, type=typ)


TYPES = [pa.int8(), pa.int16(), pa.int32(), pa.uint8(), pa.uint16(), pa.uint32(), pa.uint64(), pa.float32(),
         pa.timestamp("us"), pa.date32(), pa.date64(), pa.duration("ms"), pa.time32("s"), pa.time64("ns")]


@pytest.mark.parametrize("typ", TYPES, ids=str)
@pytest.mark.parametrize("desc", [False, True])
def test_sort_key_types(m, typ, desc):
    rng = np.random.default_rng(abs(hash(str(typ))) % 1000)
    n = 20_000
    keys = _keys(typ, rng, n)
    mask = rng.random(n) < 0.05
    keys = pa.compute.if_else(pa.array(mask), pa.scalar(None, typ), keys)
    cuts = [0, 7000, 7000, 15_000, n]
    tabs = [pa.table({"key": keys[a:b], "row": np.arange(a, b, dtype=np.int64)}) for a, b in zip(cuts, cuts[1:])]
    want = O.global_stable_sort(tabs, "key", desc)
    for fuse in (True, False):
        ctx = m.DatasetContext.get_current()
        old, ctx.fuse_device_blocks = ctx.fuse_device_blocks, fuse
        try:
            got = m.Dataset.from_arrow(tabs).sort("key", descending=desc).to_arrow_concat()
        finally:
            ctx.fuse_device_blocks = old
        assert got.schema == want.schema
        assert np.array_equal(got["row"].to_numpy(), want["row"].to_numpy()), (typ, desc, fuse)
        assert got["key"].combine_chunks().equals(want["key"].combine_chunks()) or pa.types.is_floating(typ)


def test_sort_float16_vs_numpy(m):
    # pyarrow has no sort kernel for halffloat, so neither has the reference; numpy is the oracle
    rng = np.random.default_rng(4)
    a = rng.normal(size=5000).astype(np.float16)
    t = pa.table({"key": pa.array(a, type=pa.float16()), "row": np.arange(5000, dtype=np.int64)})
    got = m.Dataset.from_arrow([t]).sort("key").to_arrow_concat()
    assert np.array_equal(got["row"].to_numpy(), np.argsort(a, kind="stable"))


@pytest.mark.parametrize("typ", [pa.int32(), pa.uint64(), pa.float32(), pa.timestamp("us"), pa.uint8()], ids=str)
def test_groupby_key_and_value_types(m, typ):
    rng = np.random.default_rng(12)
    n = 15_000
    base = _keys(typ, rng, 64)
    if pa.types.is_floating(typ):
        base = pa.array(np.unique(np.nan_to_num(base.to_numpy()) + 0.0), type=typ)  # no NaN / -0.0 keys
    keys = base.take(pa.array(rng.integers(0, len(base), n)))
    vi = rng.integers(-1000, 1000, n).astype(np.int32)
    vf = rng.normal(size=n).astype(np.float32)
    t = pa.table({"k": keys, "vi": vi, "vf": vf})
    got = m.from_arrow(t).repartition(5).groupby("k").aggregate(
        m.Count(), m.Sum("vi"), m.Min("vi"), m.Max("vf"), m.Mean("vf"), m.Std("vi")).to_arrow_concat()
    df = pd.DataFrame({"k": keys.cast(pa.int64()).to_numpy() if pa.types.is_temporal(typ) else keys.to_numpy(),
                       "vi": vi.astype(np.int64), "vf": vf.astype(np.float64)})
    exp = df.groupby("k", sort=True).agg(c=("vi", "size"), s=("vi", "sum"), mn=("vi", "min"), mx=("vf", "max"),
                                          me=("vf", "mean"), sd=("vi", "std")).reset_index()
    gk = got["k"].cast(pa.int64()).to_numpy() if pa.types.is_temporal(typ) else got["k"].to_numpy()
    assert got.schema.field("k").type == typ
    assert np.array_equal(gk, exp.k.to_numpy())
    assert got["count()"].to_pylist() == exp.c.tolist()
    assert got["sum(vi)"].to_pylist() == exp.s.tolist() and got.schema.field("sum(vi)").type == pa.int64()
    assert got["min(vi)"].to_pylist() == exp.mn.tolist()
    assert np.array_equal(got["max(vf)"].to_numpy(), exp.mx.to_numpy())
    assert np.allclose(got["mean(vf)"].to_numpy(), exp.me.to_numpy(), rtol=1e-9, atol=1e-12)
    assert np.allclose(got["std(vi)"].to_numpy(), exp.sd.fillna(0.0).to_numpy(), rtol=1e-9, atol=1e-9)
    # global forms on narrow value columns
    ds = m.from_arrow(t)
    assert ds.sum("vi") == int(vi.astype(np.int64).sum()) and ds.min("vi") == int(vi.min())
    assert ds.max("vf") == float(vf.max())
