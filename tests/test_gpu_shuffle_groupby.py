"""GPU parity: random_shuffle (permutation + per-seed determinism, the reference's
contract in test_dataset.py:4510-4566) and groupby().aggregate() against the oracle and
the golden vectors recorded from the reference."""
import numpy as np
import pyarrow as pa
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

from oracle import datasets_oracle as O  # noqa: E402


@pytest.fixture(scope="module")
def ds():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import ray_memory_management_b200 as m

    return m


def _range_tabs(cuts):
    return [pa.table({"value": np.arange(a, b, dtype=np.int64)}) for a, b in cuts]


def test_random_shuffle_contract(ds, golden):
    g = golden("shuffle")
    tabs = _range_tabs(g["block_cuts"])
    n = int(g["block_cuts"][-1][1])
    r1 = ds.Dataset.from_arrow(tabs).random_shuffle(seed=42)
    r2 = ds.Dataset.from_arrow(tabs).random_shuffle(seed=42)
    r3 = ds.Dataset.from_arrow(tabs).random_shuffle(seed=43)
    v1, v2, v3 = (r.to_arrow_concat()["value"].to_numpy() for r in (r1, r2, r3))
    assert np.array_equal(v1, v2)                      # deterministic per seed
    assert not np.array_equal(v1, v3)                  # differs across seeds
    assert np.array_equal(np.sort(v1), np.arange(n))   # a permutation (multiset-equal)
    # same output block sizes as the reference for the same plan shape
    assert r1._block_num_rows() == list(g["seed42_sizes"])
    # unseeded shuffles differ run to run (test_dataset.py:4521-4527)
    u1 = ds.Dataset.from_arrow(tabs).random_shuffle().to_arrow_concat()["value"].to_numpy()
    u2 = ds.Dataset.from_arrow(tabs).random_shuffle().to_arrow_concat()["value"].to_numpy()
    assert not np.array_equal(u1, u2)


def test_random_shuffle_ragged_num_blocks(ds, golden):
    g = golden("shuffle")
    tabs = _range_tabs(g["ragged_cuts"])
    out = ds.Dataset.from_arrow(tabs).random_shuffle(seed=5, num_blocks=3)
    assert out._block_num_rows() == list(g["ragged_sizes"])
    vals = out.to_arrow_concat()["value"].to_numpy()
    assert np.array_equal(np.sort(vals), np.sort(g["ragged_values"]))
    assert ds.Dataset.from_arrow([]).random_shuffle(seed=1).count() == 0


def test_random_shuffle_mixing_quality(ds):
    # test_dataset.py:4569-4614 of the reference: shuffled neighbours are far apart
    n = 100_000
    tabs = _range_tabs([(i * n // 10, (i + 1) * n // 10) for i in range(10)])
    v = ds.Dataset.from_arrow(tabs).random_shuffle(seed=7).to_arrow_concat()["value"].to_numpy()
    assert np.array_equal(np.sort(v), np.arange(n))
    assert np.mean(np.abs(np.diff(v))) > n / 4          # ~n/3 for a uniform permutation
    # each output block draws from every input block
    for blk in np.array_split(v, 10):
        assert len(np.unique(blk // (n // 10))) == 10
    # positions are uniform: mean displacement ~ n/3
    assert abs(np.mean(np.abs(v - np.arange(n))) - n / 3) < n / 30


def test_random_permutation_c_abi(ds):
    lib = ds._lib
    for n in (1, 2, 1000, 4097, 300_000):
        outs = []
        for seed, sub in ((1, 0), (1, 0), (2, 0), (1, 1)):
            perm = torch.empty(n, dtype=torch.int32, device="cuda")
            ws = lib.call_ws("ds_random_permutation", lambda b: torch.empty(b, dtype=torch.uint8, device="cuda"),
                             seed, sub, n, perm.data_ptr(), stream=torch.cuda.current_stream().cuda_stream)
            torch.cuda.synchronize()
            outs.append(perm.cpu().numpy().view(np.uint32))
        assert np.array_equal(np.sort(outs[0]), np.arange(n, dtype=np.uint32))
        assert np.array_equal(outs[0], outs[1])
        if n > 1000:
            assert not np.array_equal(outs[0], outs[2]) and not np.array_equal(outs[0], outs[3])


GROUPBY_AGGS = (("sum", "vf"), ("mean", "vf"), ("count", None), ("sum", "vi"),
                ("min", "vi"), ("max", "vf"), ("mean", "vi"))


def _mk_aggs(ds, spec):
    cls = {"sum": ds.Sum, "mean": ds.Mean, "min": ds.Min, "max": ds.Max}
    return tuple(ds.Count() if k == "count" else cls[k](on) for k, on in spec)


def test_groupby_golden(ds, golden):
    g = golden("groupby")
    tabs = [pa.table({"k": g["k"][a:b], "vf": g["vf"][a:b], "vi": g["vi"][a:b]}) for a, b in g["block_cuts"]]
    out = ds.Dataset.from_arrow(tabs).groupby("k").aggregate(*_mk_aggs(ds, GROUPBY_AGGS))
    cat = out.to_arrow_concat()
    assert cat.column_names == list(g["out_names"])
    assert [str(f.type) for f in cat.schema] == list(g["out_types"])
    for name in cat.column_names:
        want, got = g[f"out::{name}"], cat[name].to_numpy()
        if want.dtype.kind == "f":
            assert np.allclose(got, want, rtol=1e-6, atol=0), name  # north_star tolerance
        else:
            assert np.array_equal(got, want), name
    # blocks are key-sorted and key-range disjoint
    last = None
    for t in out.to_arrow():
        if t.num_rows:
            ks = t["k"].to_numpy()
            assert np.all(np.diff(ks) > 0)
            assert last is None or ks[0] > last
            last = ks[-1]


def test_groupby_name_conflict(ds, golden):
    g = golden("groupby")
    tabs = [pa.table({"k": g["k"][a:b], "vi": g["vi"][a:b]}) for a, b in g["block_cuts"]]
    out = ds.Dataset.from_arrow(tabs).groupby("k").aggregate(ds.Sum("vi"), ds.Sum("vi"), ds.Count(), ds.Count())
    assert out.to_arrow_concat().column_names == list(g["conflict_names"])


@pytest.mark.parametrize("groups,n", [(1, 50_000), (7, 50_000), (50_000, 60_000), (300, 5)])
def test_groupby_shapes(ds, groups, n):
    """one giant run (crosses every reduction tile), mid-size runs, mostly-unique keys."""
    rng = np.random.default_rng(groups)
    k = rng.integers(0, groups, n, dtype=np.int64)
    vf = rng.standard_normal(n)
    vi = rng.integers(-10**6, 10**6, n, dtype=np.int64)
    tabs = [pa.table({"k": k[i::3], "vf": vf[i::3], "vi": vi[i::3]}) for i in range(3)]
    spec = (("sum", "vf"), ("mean", "vf"), ("count", None), ("sum", "vi"), ("min", "vf"), ("max", "vi"))
    got = ds.Dataset.from_arrow(tabs).groupby("k").aggregate(*_mk_aggs(ds, spec)).to_arrow_concat()
    want = O.groupby_global(tabs, "k", spec)
    assert got.column_names == want.column_names and got.num_rows == want.num_rows
    for name in want.column_names:
        a, b = got[name].to_numpy(), want[name].to_numpy()
        if b.dtype.kind == "f":
            assert np.allclose(a, b, rtol=1e-6, atol=1e-9), name
        else:
            assert np.array_equal(a, b), name


def test_global_aggregate_and_errors(ds):
    t = pa.table({"v": np.arange(100, dtype=np.int64)})
    d = ds.Dataset.from_arrow([t.slice(0, 50), t.slice(50)])
    row = d.aggregate(ds.Sum("v"), ds.Mean("v"), ds.Count(), ds.Min("v"), ds.Max("v"))
    assert row == {"sum(v)": 4950, "mean(v)": 49.5, "count()": 100, "min(v)": 0, "max(v)": 99}
    with pytest.raises(ValueError):
        d.groupby("v").aggregate()
    with pytest.raises(ValueError):
        d.groupby(None).aggregate(ds.Sum())


def test_reference_tabular_groupby_known_answers(ds):
    # test_dataset.py:3050-3106 of the reference: A = x % 3, B = x over range(100)
    xs = np.arange(100, dtype=np.int64)
    rng = np.random.default_rng(0)
    rng.shuffle(xs)
    tabs = [pa.table({"A": xs[i::4] % 3, "B": xs[i::4]}) for i in range(4)]
    d = ds.Dataset.from_arrow(tabs)
    assert [r for r in d.groupby("A").count().iter_rows()] == [
        {"A": 0, "count()": 34}, {"A": 1, "count()": 33}, {"A": 2, "count()": 33}]
    assert [r for r in d.groupby("A").sum("B").iter_rows()] == [
        {"A": 0, "sum(B)": 1683}, {"A": 1, "sum(B)": 1617}, {"A": 2, "sum(B)": 1650}]
    assert [r for r in d.groupby("A").min("B").iter_rows()] == [
        {"A": 0, "min(B)": 0}, {"A": 1, "min(B)": 1}, {"A": 2, "min(B)": 2}]
    assert [r for r in d.groupby("A").max("B").iter_rows()] == [
        {"A": 0, "max(B)": 99}, {"A": 1, "max(B)": 97}, {"A": 2, "max(B)": 98}]
    assert [r for r in d.groupby("A").mean("B").iter_rows()] == [
        {"A": 0, "mean(B)": 49.5}, {"A": 1, "mean(B)": 49.0}, {"A": 2, "mean(B)": 50.0}]


# ------------------------------------------------------------------ combined (partial) blocks
def _ipc(a):
    return pa.ipc.open_stream(a.tobytes()).read_all()


def _assert_partial_equal(got: pa.Table, want: pa.Table):
    """same schema, same nulls, list contents equal (float parts to 1e-9 relative; the path allows 1e-6)"""
    assert got.schema.equals(want.schema), (got.schema, want.schema)
    assert got.num_rows == want.num_rows
    for name in want.column_names:
        g, w = got[name].combine_chunks(), want[name].combine_chunks()
        if not pa.types.is_list(w.type):
            assert g.equals(w), name
            continue
        assert g.is_valid().equals(w.is_valid()), name
        gl, wl = g.to_pylist(), w.to_pylist()
        for a, b in zip(gl, wl):
            if b is None:
                continue
            assert len(a) == len(b), name
            assert np.allclose(np.array(a, dtype=np.float64), np.array(b, dtype=np.float64), rtol=1e-9, atol=0,
                               equal_nan=True), (name, a, b)


def test_combined_blocks_export_the_reference_schema(golden):
    """ArrowBlockAccessor.combine and aggregate_combined_blocks(finalize=False) of the reference
    (recorded by oracle/gen_golden.py: gen_partial_agg) against the device path's combined
    blocks at the export boundary: list columns [accumulator parts..., has_data], the initial
    accumulator with has_data 0 for a group without valid values, None where a null was met and
    nulls are not ignored; and _GroupbyOp.reduce(partial_reduce=True) returns such a block."""
    import ray_memory_management_b200 as ds
    from ray_memory_management_b200.block import DeviceArrowBlockAccessor, PartialAggBlock
    from ray_memory_management_b200.grouped_dataset import _GroupbyOp

    g = golden("partial_agg")
    t = _ipc(g["input"])
    cut = int(g["cut"][0])
    aggs = (ds.Sum("v"), ds.Mean("v"), ds.Count(), ds.Max("w"), ds.Min("v"), ds.Std("v"),
            ds.Sum("v", ignore_nulls=False), ds.Mean("w"), ds.Sum("w"), ds.AbsMax("w"))
    halves = [t.slice(0, cut), t.slice(cut)]
    parts = [DeviceArrowBlockAccessor(ds.DeviceArrowBlock.from_arrow(h)).combine("k", aggs) for h in halves]
    for i, p in enumerate(parts):
        assert isinstance(p, PartialAggBlock)
        _assert_partial_equal(p.to_arrow(), _ipc(g[f"combined_{i}"]))
    merged, _ = _GroupbyOp.reduce("k", aggs, *parts, partial_reduce=True)
    assert isinstance(merged, PartialAggBlock)
    _assert_partial_equal(merged.to_arrow(), _ipc(g["merged_partial"]))
    # the flat (device-side) layout stays available
    assert "sum(v)#sum" in parts[0].to_flat_arrow().column_names
