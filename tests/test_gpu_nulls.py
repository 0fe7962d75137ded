"""GPU parity for nulls: nulls-last sort and the null policy of the
aggregations (python/ray/data/_internal/null_aggregate.py), against vectors recorded from
the reference (tests/golden/nulls.npz) and the reference tests' known answers
(python/ray/data/tests/test_dataset.py:3108-3151)."""
import numpy as np
import pyarrow as pa
import pyarrow.compute as pac
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def ds():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ray_memory_management_b200 as m

    return m


def _tabs(g):
    return [
        pa.table({"k": g["k"][a:b], "vf": pa.array(g["vf"][a:b], mask=g["mf"][a:b]),
                  "vi": pa.array(g["vi"][a:b], mask=g["mi"][a:b])})
        for a, b in g["block_cuts"]
    ]


@pytest.mark.parametrize("ignore", [True, False])
def test_groupby_null_policy_golden(ds, golden, ignore):
    g = golden("nulls")
    aggs = (ds.Sum("vf", ignore), ds.Mean("vf", ignore), ds.Min("vi", ignore), ds.Max("vi", ignore),
            ds.Sum("vi", ignore), ds.Mean("vi", ignore), ds.Count(), ds.AbsMax("vi", ignore),
            ds.Std("vf", ignore_nulls=ignore), ds.Std("vi", ddof=0, ignore_nulls=ignore))
    out = ds.Dataset.from_arrow(_tabs(g)).groupby("k").aggregate(*aggs).to_arrow_concat()
    tag = "ign" if ignore else "strict"
    assert out.column_names == list(g[f"{tag}_names"])
    for name in out.column_names:
        col = out[name].combine_chunks()
        valid = np.array(col.is_valid())
        assert np.array_equal(valid, g[f"{tag}::{name}::valid"]), name
        got = np.array(col.fill_null(0).to_numpy(zero_copy_only=False), dtype=np.float64)
        assert np.allclose(got[valid], g[f"{tag}::{name}::values"][valid], rtol=1e-6, atol=0), name


@pytest.mark.parametrize("desc", [False, True])
def test_sort_nulls_last_golden(ds, golden, desc):
    g = golden("nulls")
    n = len(g["k"])
    ks = pa.array(g["vi"], mask=g["mi"])
    tabs = [pa.table({"key": ks[a:b], "row": np.arange(a, b, dtype=np.int64)}) for a, b in g["block_cuts"]]
    out = ds.Dataset.from_arrow(tabs).sort("key", descending=desc).to_arrow_concat()
    assert np.array_equal(out["row"].to_numpy(), g[f"sort_{'desc' if desc else 'asc'}_row_order"])
    full = pa.concat_tables(tabs)
    want = full.take(pac.sort_indices(full, sort_keys=[("key", "descending" if desc else "ascending")]))
    assert out.equals(want)
    # accessor-level: null keys land in the last partition of sort_and_partition
    parts = ds.BlockAccessor.for_block(full).sort_and_partition(
        [0, 20] if not desc else [20, 0], [("key", "descending" if desc else "ascending")], desc)
    assert sum(p.num_rows for p in parts) == n
    assert parts[-1].to_arrow()["key"].null_count == int(g["mi"].sum())
    assert all(p.to_arrow()["key"].null_count == 0 for p in parts[:-1])


def test_fsb_key_and_payload_nulls_roundtrip(ds):
    keys = [b"bb", None, b"aa", b"cc", None, b"aa"]
    pay = [1.5, None, 2.5, None, 3.5, 4.5]
    t = pa.table({"k": pa.array(keys, pa.binary(2)), "p": pa.array(pay, pa.float64())})
    blk = ds.DeviceArrowBlock.from_arrow(t)
    assert blk.to_arrow().equals(t)  # validity survives H2D / D2H
    out = ds.Dataset.from_arrow([t.slice(0, 3), t.slice(3)]).sort("k").to_arrow_concat()
    assert out.equals(t.take(pac.sort_indices(t, sort_keys=[("k", "ascending")])))
    shuf = ds.Dataset.from_arrow([t]).random_shuffle(seed=3).to_arrow_concat()
    assert sorted(map(str, shuf.to_pylist())) == sorted(map(str, t.to_pylist()))


def test_reference_sum_with_nulls_known_answers(ds):
    # test_dataset.py:3108-3151 of the reference
    xs = np.arange(100)
    np.random.default_rng(1).shuffle(xs)
    A = pa.array(list(xs % 3) + [0], pa.int64())
    B = pa.array([int(x) for x in xs] + [None], pa.int64())
    t = pa.table({"A": A, "B": B})
    d = ds.Dataset.from_arrow([t.slice(0, 40), t.slice(40, 30), t.slice(70)])
    assert [r for r in d.groupby("A").sum("B").iter_rows()] == [
        {"A": 0, "sum(B)": 1683}, {"A": 1, "sum(B)": 1617}, {"A": 2, "sum(B)": 1650}]
    assert [r for r in d.groupby("A").sum("B", ignore_nulls=False).iter_rows()] == [
        {"A": 0, "sum(B)": None}, {"A": 1, "sum(B)": 1617}, {"A": 2, "sum(B)": 1650}]
    alln = pa.table({"A": pa.array(xs % 3, pa.int64()), "B": pa.array([None] * 100, pa.int64())})
    d = ds.Dataset.from_arrow([alln.slice(0, 50), alln.slice(50)])
    assert [r for r in d.groupby("A").sum("B").iter_rows()] == [
        {"A": 0, "sum(B)": None}, {"A": 1, "sum(B)": None}, {"A": 2, "sum(B)": None}]
    assert [r for r in d.groupby("A").mean("B").iter_rows()] == [
        {"A": 0, "mean(B)": None}, {"A": 1, "mean(B)": None}, {"A": 2, "mean(B)": None}]
    # Count counts rows, nulls included (aggregate.py:104-106)
    assert [r["count()"] for r in d.groupby("A").count().iter_rows()] == [34, 33, 33]


def test_std_matches_pandas_like_reference_test(ds):
    # test_dataset.py:3542-3599 of the reference compares Std with pandas (ddof 1 and 0)
    import pandas as pd

    rng = np.random.default_rng(4)
    n = 40_000
    df = pd.DataFrame({"A": rng.integers(0, 37, n), "B": rng.standard_normal(n) * 100 + 1e6})
    tabs = [pa.Table.from_pandas(df.iloc[i::5], preserve_index=False) for i in range(5)]
    d = ds.Dataset.from_arrow(tabs)
    for ddof in (1, 0):
        got = d.groupby("A").std("B", ddof=ddof).to_arrow_concat().to_pandas()
        want = df.groupby("A")["B"].std(ddof=ddof).reset_index()
        assert np.array_equal(got["A"].to_numpy(), want["A"].to_numpy())
        assert np.allclose(got["std(B)"].to_numpy(), want["B"].to_numpy(), rtol=1e-6, atol=0)
    # single-row groups give 0.0 (aggregate.py:273-274)
    t = pa.table({"A": np.arange(5, dtype=np.int64), "B": np.arange(5, dtype=np.float64)})
    assert ds.Dataset.from_arrow([t]).groupby("A").std("B").to_arrow_concat()["std(B)"].to_pylist() == [0.0] * 5


def test_slice_copy_and_take_keep_validity(ds):
    t = pa.table({"a": pa.array([1, None, 3, None, 5], pa.int64())})
    acc = ds.BlockAccessor.for_block(t)
    assert acc.slice(1, 4, copy=True).to_arrow().equals(t.slice(1, 3))
    assert acc.slice(1, 4, copy=False).to_arrow().equals(t.slice(1, 3))
    assert acc.take([4, 1, 0]).to_arrow().equals(t.take([4, 1, 0]))
    import torch

    with pytest.raises(NotImplementedError):
        acc.to_block().to_host_buffers([torch.empty(40, dtype=torch.uint8)])
