"""GPU tests of the immediate callers of the hot path: map_groups,
DatasetPipeline windows/epochs, the loaders and Dataset.stats().  Expected values are the
reference's own test expectations (test_dataset.py:3941-4065, test_dataset_pipeline.py:
188-382,601-622, test_stats.py:58-133) restated for table datasets, or pandas as oracle."""
import re

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


def _values(rows):
    return [r["value"] for r in rows]


# ------------------------------------------------------------------------- map_groups
@pytest.mark.parametrize("num_parts", [1, 2, 30])
def test_map_groups_none_groupkey(m, num_parts):
    ds = m.range_table(100, parallelism=num_parts)
    out = ds.groupby(None).map_groups(lambda df: pd.DataFrame({"value": [df.value.min() + df.value.max()]}))
    assert out.count() == 1
    assert _values(out.take()) == [99]


@pytest.mark.parametrize("num_parts", [1, 2, 30])
def test_map_groups_returning_empty_result(m, num_parts):
    t = pa.table({"k": np.arange(100, dtype=np.int64) % 3, "value": np.arange(100, dtype=np.int64)})
    ds = m.from_arrow(t).repartition(num_parts)
    out = ds.groupby("k").map_groups(lambda df: df.iloc[:0])
    assert out.count() == 0
    assert out.take() == []


def test_map_groups_empty_dataset(m):
    out = m.Dataset.from_arrow([]).groupby("k").map_groups(lambda df: df)
    assert out.count() == 0


@pytest.mark.parametrize("num_parts", [1, 2, 3, 30])
def test_map_groups_for_pandas(m, num_parts):
    df = pd.DataFrame({"A": np.array([1, 1, 2], dtype=np.int64), "B": [1, 1, 3], "C": [4, 6, 5]})
    grouped = m.from_pandas(df).repartition(num_parts).groupby("A")
    mapped = grouped.map_groups(
        lambda g: g.apply(lambda col: col / g[col.name].sum() if col.name in ["B", "C"] else col)
    )
    assert mapped.count() == 3
    expected = pd.DataFrame({"A": [1, 1, 2], "B": [0.5, 0.5, 1.0], "C": [0.4, 0.6, 1.0]})
    assert mapped.to_arrow_concat().to_pandas().equals(expected)


@pytest.mark.parametrize("num_parts", [1, 2, 3, 30])
def test_map_groups_for_arrow(m, num_parts):
    at = pa.table({"A": np.array([1, 1, 2], dtype=np.int64), "B": [1, 1, 3], "C": [4, 6, 5]})
    grouped = m.from_arrow(at).repartition(num_parts).groupby("A")

    def normalize(t: pa.Table):
        import pyarrow.compute as pc

        r = t.select(["A"])
        r = r.append_column("B", pc.divide(t.column("B"), pc.sum(t.column("B")).cast(pa.float64())))
        r = r.append_column("C", pc.divide(t.column("C"), pc.sum(t.column("C")).cast(pa.float64())))
        return r

    mapped = grouped.map_groups(normalize, batch_format="pyarrow")
    assert mapped.count() == 3
    expected = pa.table({"A": np.array([1, 1, 2], dtype=np.int64), "B": [0.5, 0.5, 1.0], "C": [0.4, 0.6, 1.0]})
    assert mapped.to_arrow_concat().equals(expected)


@pytest.mark.parametrize("num_parts", [1, 7])
def test_map_groups_median_vs_pandas(m, num_parts):
    rng = np.random.default_rng(11)
    n = 20_000
    df = pd.DataFrame({"k": rng.integers(-50, 50, n), "v": rng.normal(size=n)})
    cuts = [n * i // num_parts for i in range(num_parts + 1)]
    tabs = [pa.Table.from_pandas(df.iloc[a:b], preserve_index=False) for a, b in zip(cuts, cuts[1:])]
    seen = []

    def fn(g):
        seen.append(int(g.k.iloc[0]))
        assert (g.k == g.k.iloc[0]).all()
        return pd.DataFrame({"k": [g.k.iloc[0]], "med": [g.v.median()], "n": [len(g)]})

    out = m.from_arrow(tabs).groupby("k").map_groups(fn).to_arrow_concat().to_pandas()
    exp = df.groupby("k").agg(med=("v", "median"), n=("v", "size")).reset_index()
    assert seen == sorted(set(df.k))            # every group exactly once, in key order
    assert out.k.tolist() == exp.k.tolist() and out.n.tolist() == exp.n.tolist()
    assert np.allclose(out.med, exp.med, rtol=0, atol=0)
    # within a group rows keep their input order (the sort is stable)
    first = m.from_arrow(tabs).groupby("k").map_groups(lambda g: g.iloc[:1]).to_arrow_concat().to_pandas()
    assert np.array_equal(first.v.to_numpy(), df.groupby("k", sort=True).v.first().to_numpy())


def test_map_groups_fixed_size_binary_key_and_errors(m):
    keys = [b"bb", b"aa", b"bb", b"cc", b"aa", b"bb"]
    t = pa.table({"k": pa.array(keys, pa.binary(2)), "v": np.arange(6, dtype=np.int64)})
    out = m.from_arrow(t).groupby("k").map_groups(
        lambda g: pa.table({"k": g.column("k").slice(0, 1), "s": [pa.compute.sum(g.column("v")).as_py()]}),
        batch_format="pyarrow",
    ).to_arrow_concat()
    assert out["k"].to_pylist() == [b"aa", b"bb", b"cc"] and out["s"].to_pylist() == [5, 7, 3]
    with pytest.raises(TypeError):              # test_dataset.py:3956-3966
        m.from_arrow(t).groupby("k").map_groups(lambda g: None)
    with pytest.raises(ValueError):
        m.from_arrow(t).groupby("k").map_groups(lambda g: g, batch_format="numpy")


# ------------------------------------------------------------------------- pipelines
def test_window_and_rewindow(m):
    ds = m.range_table(10, parallelism=10)
    pipe = ds.window(blocks_per_window=1)
    assert str(pipe) == "DatasetPipeline(num_windows=10, num_stages=1)"
    pipe = pipe.rewindow(blocks_per_window=3)
    assert str(pipe) == "DatasetPipeline(num_windows=None, num_stages=1)"
    got = [_values(d.take()) for d in pipe.iter_datasets()]
    assert got == [[0, 1, 2], [3, 4, 5], [6, 7, 8], [9]]
    with pytest.raises(RuntimeError):
        list(pipe.iter_datasets())          # "Pipeline cannot be read multiple times."
    with pytest.raises(ValueError):
        ds.window(blocks_per_window=1, bytes_per_window=8)


def test_window_by_bytes(m):
    ds = m.range_table(16, parallelism=16)   # 8 bytes per block
    sizes = [d.num_blocks() for d in ds.window(bytes_per_window=24).iter_datasets()]
    assert sizes == [3, 3, 3, 3, 3, 1]
    sizes = [d.num_blocks() for d in ds.window(bytes_per_window=1).iter_datasets()]
    assert sizes == [1] * 16                 # at least one block per window


def test_repeat_and_epochs(m):
    ds = m.range_table(5, parallelism=5)
    pipe = ds.window(blocks_per_window=1).repeat(2)
    assert str(pipe) == "DatasetPipeline(num_windows=10, num_stages=1)"
    assert _values(pipe.take()) == list(range(5)) * 2
    pipe = ds.window(blocks_per_window=1).repeat()
    assert str(pipe) == "DatasetPipeline(num_windows=inf, num_stages=1)"
    assert len(pipe.take(99)) == 99
    with pytest.raises(ValueError):
        ds.repeat().repeat()
    got = [_values(p.take()) for p in m.range_table(3, parallelism=3).window(blocks_per_window=2).repeat(3).iter_epochs()]
    assert got == [[0, 1, 2]] * 3
    got = [_values(p.take()) for p in m.range_table(3, parallelism=3).window(blocks_per_window=2).repeat(3).iter_epochs(2)]
    assert got == [[0, 1, 2]] * 2
    got = [_values(p.take()) for p in m.range_table(5).repeat(2).rewindow(blocks_per_window=2).iter_epochs()]
    assert got == [[0, 1, 2, 3, 4]] * 2


def test_from_iterable_and_foreach_window(m):
    pipe = m.DatasetPipeline.from_iterable([lambda: m.range_table(3), lambda: m.range_table(2)])
    assert _values(pipe.take()) == [0, 1, 2, 0, 1]
    pipe = m.DatasetPipeline.from_iterable([lambda: m.range_table(3), lambda: m.range_table(2)])
    assert pipe.foreach_window(lambda d: d.sort("value", descending=True)).count() == 5


def test_sort_each_window(m):
    pipe = m.range_table(12, parallelism=12).window(blocks_per_window=3).sort_each_window("value")
    assert _values(pipe.take()) == list(range(12))
    pipe = m.range_table(12, parallelism=12).window(blocks_per_window=3).sort_each_window("value", descending=True)
    assert _values(pipe.take()) == [2, 1, 0, 5, 4, 3, 8, 7, 6, 11, 10, 9]


def test_random_shuffle_each_window(m):
    # per-epoch shuffle: every epoch is a permutation of the data; a fixed seed repeats it,
    # no seed gives a fresh order per epoch (dataset_pipeline.py:809-819 -> Dataset.random_shuffle)
    n = 1000
    base = m.range_table(n, parallelism=4)
    epochs = [_values(p.take(n)) for p in base.repeat(3).random_shuffle_each_window(seed=3).iter_epochs()]
    assert all(sorted(e) == list(range(n)) for e in epochs)
    assert epochs[0] == epochs[1] == epochs[2] and epochs[0] != list(range(n))
    epochs = [_values(p.take(n)) for p in base.repeat(2).random_shuffle_each_window().iter_epochs()]
    assert all(sorted(e) == list(range(n)) for e in epochs) and epochs[0] != epochs[1]
    # windows are shuffled independently: values never leave their window
    wins = [_values(d.take(n)) for d in base.window(blocks_per_window=2).random_shuffle_each_window(seed=1).iter_datasets()]
    assert [sorted(w) for w in wins] == [list(range(0, 500)), list(range(500, 1000))]
    rp = base.window(blocks_per_window=2).repartition_each_window(5, shuffle=True)
    assert [d.num_blocks() for d in rp.iter_datasets()] == [5, 5]


# ------------------------------------------------------------------------- stats
def _canonicalize(stats: str) -> str:
    s = re.sub(r"[0-9\.]+(ms|us|s)", "T", stats)      # test_stats.py:12-23
    s = re.sub(r"[0-9]+(\.[0-9]+)?", "N", s)
    return s.replace("\t", "    ")


_SUBSTAGE = """    Substage N {name}: N/N blocks executed
    * Remote wall time: T min, T max, T mean, T total
    * Remote cpu time: T min, T max, T mean, T total
    * Peak heap memory usage (MiB): N min, N max, N mean
    * Output num rows: N min, N max, N mean, N total
    * Output size bytes: N min, N max, N mean, N total
    * Tasks per node: N min, N max, N mean; N nodes used
"""


@pytest.mark.parametrize("fuse", [True, False])
def test_stats_shuffle_text(m, fuse):
    ctx = m.DatasetContext.get_current()
    old = ctx.fuse_device_blocks
    ctx.fuse_device_blocks = fuse
    try:
        ds = m.range_table(1000, parallelism=10).random_shuffle().repartition(1, shuffle=True)
        stats = _canonicalize(ds.stats())
    finally:
        ctx.fuse_device_blocks = old
    expected = (
        "Stage N from_blocks: N/N blocks split from parent in T\n"
        "* Output num rows: N min, N max, N mean, N total\n"
        "* Output size bytes: N min, N max, N mean, N total\n"
        "\n"
        "Stage N random_shuffle: executed in T\n\n"
        + _SUBSTAGE.format(name="random_shuffle_map") + "\n"
        + _SUBSTAGE.format(name="random_shuffle_reduce") + "\n"
        + "Stage N repartition: executed in T\n\n"
        + _SUBSTAGE.format(name="repartition_map") + "\n"
        + _SUBSTAGE.format(name="repartition_reduce")
    )
    assert stats == expected, stats


def test_stats_sort_and_aggregate(m):
    ds = m.range_table(1000, parallelism=10).sort("value")
    stats = ds.stats()
    assert "sort_map" in stats and "sort_reduce" in stats      # test_stats.py:128-133
    assert list(ds._stats.stages) == ["sort_map", "sort_reduce"]
    assert len(ds._stats.stages["sort_map"]) == 10 and len(ds._stats.stages["sort_reduce"]) == 10
    assert sum(md.num_rows for md in ds._stats.stages["sort_reduce"]) == 1000
    agg = ds.groupby("value").count()
    assert "aggregate_map" in agg.stats() and "aggregate_reduce" in agg.stats()
    assert agg.stats().index("sort_map") < agg.stats().index("aggregate_map")   # parents first


def test_stats_wall_time_is_device_time(m):
    """'Remote wall time' of a sort is the DEVICE time of its launches (CUDA events recorded by
    the BlockExecStats builder), not the host time it took to enqueue them: it agrees with a
    CUDA-event pair around the whole sort, and NVTX ranges do not change results."""
    import torch

    from ray_memory_management_b200 import _nvtx

    n = 4_000_000
    rng = np.random.default_rng(9)
    t = pa.table({"k": rng.integers(0, 2**62, n, dtype=np.int64), "v": np.arange(n, dtype=np.int64)})
    blk = m.DeviceArrowBlock.from_arrow(t)
    for _ in range(2):  # warm-up: workspace allocations
        m.Dataset.from_blocks([blk]).sort("k")
    torch.cuda.synchronize()
    nvtx_on = _nvtx.enable(True)  # ranges around every stage phase and ds_* call (no-op without NVTX)
    try:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ds = m.Dataset.from_blocks([blk]).sort("k")
        e1.record()
        torch.cuda.synchronize()
    finally:
        _nvtx.enable(False)
    assert isinstance(nvtx_on, bool)
    outer_s = e0.elapsed_time(e1) / 1e3
    metas = ds._stats.stages["sort_reduce"]
    dev_s = max(md.exec_stats.wall_time_s for md in metas)
    assert dev_s > 0
    assert 0.3 * outer_s <= dev_s <= 1.05 * outer_s, (dev_s, outer_s)
    assert all(md.exec_stats.host_time_s is not None for md in metas)
    assert np.array_equal(ds.to_arrow_concat()["k"].to_numpy(), np.sort(t["k"].to_numpy()))


# ------------------------------------------------------------------------- loaders
def test_read_parquet_and_ipc_round_trip(m, tmp_path):
    import pyarrow.parquet as pq

    rng = np.random.default_rng(5)
    n = 10_000
    t = pa.table({
        "key": pa.array([bytes(r) for r in rng.integers(0, 256, (n, 10), dtype=np.uint8)], pa.binary(10)),
        "v": rng.normal(size=n),
        "i": rng.integers(-2**40, 2**40, n),
    })
    pq.write_table(t, tmp_path / "a.parquet", row_group_size=3000)
    ds = m.read_parquet(str(tmp_path / "a.parquet"))
    assert ds._block_num_rows() == [3000, 3000, 3000, 1000]
    assert ds.to_arrow_concat().equals(t)
    assert m.read_parquet(str(tmp_path), columns=["i"], parallelism=3)._block_num_rows() == [3333, 3333, 3334]
    out = ds.sort("key")
    order = np.argsort(np.array(t["key"].to_pylist(), dtype="S10"), kind="stable")
    assert out.to_arrow_concat().equals(t.take(order))
    paths = m.write_arrow_ipc(out, str(tmp_path / "ipc"))
    assert len(paths) == out.num_blocks()
    back = m.read_arrow_ipc(str(tmp_path / "ipc"))
    assert back._block_num_rows() == out._block_num_rows()
    assert back.to_arrow_concat().equals(t.take(order))
    with pytest.raises(FileNotFoundError):
        m.read_parquet(str(tmp_path / "missing.parquet"))
