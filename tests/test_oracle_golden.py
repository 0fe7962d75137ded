"""CPU: pins oracle/datasets_oracle.py against the vectors oracle/gen_golden.py
recorded from the reference's own source (tests/golden/*.npz)."""
import random

import numpy as np
import pyarrow as pa
import pytest

from oracle import datasets_oracle as O


def _tabs(cuts, **cols):
    return [pa.table({k: v[a:b] for k, v in cols.items()}) for a, b in cuts]


@pytest.mark.parametrize("name", ["unique", "ties"])
@pytest.mark.parametrize("desc", [False, True])
def test_sort_int64(golden, name, desc):
    g = golden("sort_int64")
    keys = g[f"{name}_keys"]
    tabs = _tabs(g["block_cuts"], value=np.arange(len(keys), dtype=np.int64), key=keys)
    want = g[f"{name}_{'desc' if desc else 'asc'}_value_order"]
    random.seed(0)  # block boundaries differ from the recorded run; the concatenation must not
    blocks = O.sort_dataset(tabs, "key", desc)
    assert len(blocks) == len(tabs)
    got = pa.concat_tables([b for b in blocks if b.num_rows])["value"].to_numpy()
    assert np.array_equal(got, want)
    assert np.array_equal(O.global_stable_sort(tabs, "key", desc)["value"].to_numpy(), want)


@pytest.mark.parametrize("name", ["uniform", "ties"])
@pytest.mark.parametrize("desc", [False, True])
def test_sort_graysort(golden, name, desc):
    g = golden("sort_graysort")
    keys = g[f"{name}_keys"]
    n = len(keys)
    base = O.graysort_table(n, seed=11)
    t = pa.table({"key": O.fsb_from_numpy(keys), "payload": base["payload"]})
    tabs = [t.slice(a, b - a) for a, b in g["block_cuts"]]
    blocks = O.sort_dataset(tabs, "key", desc)
    cat = pa.concat_tables([b for b in blocks if b.num_rows])
    rowid = O.fsb_to_numpy(cat["payload"])[:, :8].copy().view("<u8").ravel()
    assert np.array_equal(rowid, g[f"{name}_{'desc' if desc else 'asc'}_rowid_order"])


@pytest.mark.parametrize("desc", [False, True])
def test_sort_and_partition(golden, desc):
    g = golden("partition")
    keys = g["keys"]
    t = pa.table({"key": keys, "value": np.arange(len(keys), dtype=np.int64)})
    tag = "desc" if desc else "asc"
    parts = O.sort_and_partition(t, list(g[f"{tag}_boundaries"]), "key", desc)
    assert [p.num_rows for p in parts] == list(g[f"{tag}_sizes"])
    assert np.array_equal(pa.concat_tables(parts)["value"].to_numpy(), g[f"{tag}_value_order"])


def test_merge_sorted_blocks(golden):
    g = golden("partition")
    keys = g["keys"]
    t = pa.table({"key": keys, "value": np.arange(len(keys), dtype=np.int64)})
    blocks = [O.sort_table(t.slice(a, b - a), "key", False) for a, b in g["merge_cuts"]]
    m = O.merge_sorted_blocks(blocks, "key", False)
    assert np.array_equal(m["value"].to_numpy(), g["merge_value_order"])
    assert O.merge_sorted_blocks([O.empty_table()], "key", False).num_rows == 0


def test_empty_block_contracts():
    # test_sort.py:110-127 of the reference
    parts = O.sort_and_partition(O.empty_table(), [1, 5, 10], "A", False)
    assert len(parts) == 4 and all(p.num_rows == 0 for p in parts)


def test_boundaries(golden):
    g = golden("boundaries")
    for nsamp, r in g["q_cases"]:
        got = O.boundaries_from_samples(g[f"q_{nsamp}_{r}_samples"], int(r))
        assert np.array_equal(np.array(got, dtype=np.int64), g[f"q_{nsamp}_{r}_bounds"])
    keys = g["sampled_keys"]
    tabs = _tabs(g["sampled_cuts"], key=keys)
    random.seed(1234)
    got = O.sample_boundaries(tabs, "key", len(tabs))
    assert np.array_equal(np.array(got, dtype=np.int64), g["sampled_bounds"])
    empty = [pa.table({"key": pa.array([], pa.int64())})] * 3
    assert O.sample_boundaries(empty, "key", 3) == [None, None]


@pytest.mark.parametrize("seed", [42, 43])
def test_random_shuffle_bit_exact(golden, seed):
    g = golden("shuffle")
    tabs = [pa.table({"value": np.arange(a, b, dtype=np.int64)}) for a, b in g["block_cuts"]]
    blocks = O.random_shuffle_dataset(tabs, seed)
    assert [b.num_rows for b in blocks] == list(g[f"seed{seed}_sizes"])
    assert np.array_equal(pa.concat_tables(blocks)["value"].to_numpy(), g[f"seed{seed}_values"])


def test_random_shuffle_ragged(golden):
    g = golden("shuffle")
    tabs = [pa.table({"value": np.arange(a, b, dtype=np.int64)}) for a, b in g["ragged_cuts"]]
    blocks = O.random_shuffle_dataset(tabs, 5, num_blocks=3)
    assert [O.num_rows(b) for b in blocks] == list(g["ragged_sizes"])
    vals = np.concatenate([b["value"].to_numpy() for b in blocks if O.num_rows(b)])
    assert np.array_equal(vals, g["ragged_values"])


GROUPBY_AGGS = (
    ("sum", "vf"), ("mean", "vf"), ("count", None), ("sum", "vi"),
    ("min", "vi"), ("max", "vf"), ("mean", "vi"),
)


def test_groupby(golden):
    g = golden("groupby")
    tabs = _tabs(g["block_cuts"], k=g["k"], vf=g["vf"], vi=g["vi"])
    random.seed(1)
    blocks = O.groupby_dataset(tabs, "k", GROUPBY_AGGS)
    cat = pa.concat_tables([b for b in blocks if b.num_rows])
    assert cat.column_names == list(g["out_names"])
    assert [str(f.type) for f in cat.schema] == list(g["out_types"])
    for name in cat.column_names:
        want, got = g[f"out::{name}"], cat[name].to_numpy()
        if want.dtype.kind == "f":
            assert np.allclose(got, want, rtol=1e-12, atol=0), name
        else:
            assert np.array_equal(got, want), name
    glob = O.groupby_global(tabs, "k", GROUPBY_AGGS)
    for name in cat.column_names:
        assert np.allclose(glob[name].to_numpy(), g[f"out::{name}"], rtol=1e-12), name


def test_groupby_name_conflict(golden):
    g = golden("groupby")
    aggs = (("sum", "vi"), ("sum", "vi"), ("count", None), ("count", None))
    assert ["k"] + O.resolved_names(aggs) == list(g["conflict_names"])


def _unpack_rows(offsets, data):
    raw = data.tobytes()
    return [raw[offsets[i]:offsets[i + 1]] for i in range(len(offsets) - 1)]


@pytest.mark.parametrize("desc", [False, True])
def test_sort_varlen_binary(golden, desc):
    """binary keys of length 0..19 with shared prefixes and NUL bytes (tests/golden/varlen.npz,
    recorded from the reference's sort_impl)."""
    g = golden("varlen")
    rows = _unpack_rows(g["bin_offsets"], g["bin_data"])
    tabs = [pa.table({"key": pa.array(rows[a:b], pa.binary()), "row": np.arange(a, b, dtype=np.int64)})
            for a, b in g["block_cuts"]]
    want = g[f"plain_{'desc' if desc else 'asc'}_row_order"]
    random.seed(1)
    blocks = O.sort_dataset(tabs, "key", desc)
    got = pa.concat_tables([b for b in blocks if b.num_rows])["row"].to_numpy()
    assert np.array_equal(got, want)
    assert np.array_equal(O.global_stable_sort(tabs, "key", desc)["row"].to_numpy(), want)


def test_sort_varlen_string(golden):
    g = golden("varlen")
    words = [r.decode() for r in _unpack_rows(g["str_offsets"], g["str_data"])]
    tabs = [pa.table({"key": pa.array(words[a:b], pa.string()), "row": np.arange(a, b, dtype=np.int64)})
            for a, b in g["block_cuts"]]
    assert np.array_equal(O.global_stable_sort(tabs, "key", False)["row"].to_numpy(), g["str_asc_row_order"])


def test_partial_agg_golden_is_the_reference_schema(golden):
    """the combined blocks recorded from the reference: list columns with a trailing has_data
    flag, None for groups that met a null with ignore_nulls=False (null_aggregate.py:29-65)"""
    import pyarrow as pa

    g = golden("partial_agg")
    t = pa.ipc.open_stream(g["combined_0"].tobytes()).read_all()
    assert t.column_names[:4] == ["k", "sum(v)", "mean(v)", "count()"]
    assert pa.types.is_list(t["sum(v)"].type) and pa.types.is_int64(t["count()"].type)
    assert t["sum(v)_2"].null_count > 0 and t["sum(v)"].null_count == 0
    assert all(len(x) == 3 for x in t["mean(v)"].to_pylist()) and all(len(x) == 4 for x in t["std(v)"].to_pylist())
    m = pa.ipc.open_stream(g["merged_partial"].tobytes()).read_all()
    assert m.schema.equals(t.schema)
