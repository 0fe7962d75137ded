"""GPU tests of the aggregation call shapes around groupby().aggregate(): ``on`` as a column,
a list of columns or None, the global forms ``Dataset.sum/min/max/mean/std`` and the error
behaviour -- the reference's test_dataset.py:2984-3016 (errors), :3144-3196 (global sum),
:3653-3686 (multicolumn), :3688-3726 (bad on), :3729-3781 (multi agg) restated for tables."""
import math
import random

import numpy as np
import pandas as pd
import pyarrow as pa
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def m():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ray_memory_management_b200 as mod

    return mod


def _xs():
    xs = list(range(100))
    random.Random(1234).shuffle(xs)
    return xs


@pytest.mark.parametrize("num_parts", [1, 30])
def test_groupby_multicolumn(m, num_parts):
    xs = _xs()
    df = pd.DataFrame({"A": [x % 3 for x in xs], "B": xs, "C": [2 * x for x in xs]})
    want = [
        {"A": 0, "mean(B)": 49.5, "mean(C)": 99.0},
        {"A": 1, "mean(B)": 49.0, "mean(C)": 98.0},
        {"A": 2, "mean(B)": 50.0, "mean(C)": 100.0},
    ]
    agg = m.from_pandas(df).repartition(num_parts).groupby("A").mean(["B", "C"])
    assert agg.count() == 3
    assert list(agg.sort("A").iter_rows()) == want
    # unspecified column: every column except the groupby key
    agg = m.from_pandas(df).repartition(num_parts).groupby("A").mean()
    assert list(agg.sort("A").iter_rows()) == want
    df = pd.DataFrame({"A": xs, "B": [2 * x for x in xs]})
    row = m.from_pandas(df).repartition(num_parts).mean(["A", "B"])
    assert row["mean(A)"] == df["A"].mean() and row["mean(B)"] == df["B"].mean()
    row = m.from_pandas(df).repartition(num_parts).max()          # on=None: all columns
    assert row == {"max(A)": 99, "max(B)": 198}


@pytest.mark.parametrize("num_parts", [1, 30])
def test_groupby_multi_agg(m, num_parts):
    xs = _xs()
    df = pd.DataFrame({"A": [x % 3 for x in xs], "B": xs})
    agg = m.from_pandas(df).repartition(num_parts).groupby("A").aggregate(
        m.Count(), m.Sum("B"), m.Min("B"), m.Max("B"), m.Mean("B"), m.Std("B"))
    assert agg.count() == 3
    got = agg.sort("A").to_pandas()
    exp = df.groupby("A")["B"]
    np.testing.assert_array_equal(got["count()"].to_numpy(), [34, 33, 33])
    for name in ["sum", "min", "max", "mean", "std"]:
        res, want = got[f"{name}(B)"].to_numpy(), getattr(exp, name)().to_numpy()
        if name == "std":
            np.testing.assert_array_almost_equal(res, want)
        else:
            np.testing.assert_array_equal(res, want)
    df = pd.DataFrame({"A": xs})
    row = m.from_pandas(df).repartition(num_parts).aggregate(
        m.Sum("A"), m.Min("A"), m.Max("A"), m.Mean("A"), m.Std("A"))
    for name in ["sum", "min", "max", "mean", "std"]:
        want = getattr(df["A"], name)()
        assert math.isclose(row[f"{name}(A)"], want) if name == "std" else row[f"{name}(A)"] == want


@pytest.mark.parametrize("num_parts", [1, 30])
def test_global_aggregations_with_nulls(m, num_parts):
    xs = _xs()
    ds = m.from_arrow(pa.table({"A": pa.array(xs, pa.int64())})).repartition(num_parts)
    assert ds.sum("A") == 4950 and ds.min("A") == 0 and ds.max("A") == 99 and ds.mean("A") == 49.5
    assert math.isclose(ds.std("A"), pd.Series(xs).std())
    assert math.isclose(ds.std("A", ddof=0), pd.Series(xs).std(ddof=0))
    assert m.Dataset.from_arrow([]).sum("value") is None                 # empty dataset
    nan_ds = m.from_arrow(pa.table({"A": pa.array(xs + [None], pa.int64())})).repartition(num_parts)
    assert nan_ds.sum("A") == 4950 and nan_ds.mean("A") == 49.5
    assert nan_ds.sum("A", ignore_nulls=False) is None
    assert nan_ds.std("A", ignore_nulls=False) is None
    all_null = m.from_arrow(pa.table({"A": pa.array([None] * 100, pa.int64())})).repartition(num_parts)
    assert all_null.sum("A") is None and all_null.sum("A", ignore_nulls=False) is None
    assert all_null.max("A") is None and all_null.mean("A") is None


def test_bad_on_and_key_errors(m):
    xs = list(range(100))
    df = pd.DataFrame({"A": [x % 3 for x in xs], "B": xs, "C": [2 * x for x in xs]})
    ds = m.from_pandas(df)
    for target in (ds.groupby("A"), ds):
        with pytest.raises(TypeError):
            target.mean(5)
        with pytest.raises(TypeError):
            target.mean([5])
        with pytest.raises(ValueError):
            target.mean([])
        with pytest.raises(ValueError):
            target.mean("D")
        with pytest.raises(ValueError):
            target.mean(["B", "D"])
    t = m.range_table(100)
    assert t.groupby(None).count().take() == [{"count()": 100}]
    with pytest.raises(ValueError):
        t.groupby(lambda x: x % 2).count()
    with pytest.raises(ValueError):
        t.groupby("foo").count()
    assert t.aggregate(m.Max("value")) == {"max(value)": 99}
    with pytest.raises(ValueError):
        t.aggregate(m.Max())
    with pytest.raises(ValueError):
        t.aggregate(m.Max(lambda x: x))
    with pytest.raises(ValueError):
        t.aggregate(m.Max("bad_field"))
    with pytest.raises(ValueError):
        t.aggregate()
    assert m.range_table(10).schema().names == ["value"] and m.Dataset.from_arrow([]).schema() is None
    assert m.range_table(3).take_all() == [{"value": 0}, {"value": 1}, {"value": 2}]
    with pytest.raises(ValueError):
        m.range_table(30).take_all(limit=10)
