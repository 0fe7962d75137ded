"""CPU tests of the host logic around the hot path's callers: window / epoch bookkeeping of
DatasetPipeline and the DatasetStats summary text.  No device work: blocks are opaque tokens.
Where /root/reference is present the same inputs are run through the reference's own
BlockList.split / split_by_bytes and DatasetStats.summary_string (container-only pinning)."""
import math

import numpy as np
import pytest

torch = pytest.importorskip("torch")

from ray_memory_management_b200.block import BlockExecStats, BlockList, BlockMetadata  # noqa: E402
from ray_memory_management_b200.dataset import Dataset  # noqa: E402
from ray_memory_management_b200.dataset_pipeline import (  # noqa: E402
    DatasetPipeline,
    split_block_list,
    split_block_list_by_bytes,
)
from ray_memory_management_b200.stats import DatasetStats, DatasetStatsBuilder  # noqa: E402


def _meta(rows, size, wall=None):
    es = None
    if wall is not None:
        es = BlockExecStats()
        es.wall_time_s, es.cpu_time_s, es.max_rss_bytes, es.node_id = wall, wall / 2, 300 << 20, "cuda:0"
    return BlockMetadata(num_rows=rows, size_bytes=size, schema=None, input_files=None, exec_stats=es)


def _blocklist(n, size=8):
    return BlockList([f"b{i}" for i in range(n)], [_meta(1, size) for _ in range(n)])


def _tokens(ds):
    return ds._blocks.get_blocks()


# ------------------------------------------------------------------ BlockList.split*
@pytest.mark.parametrize("n,size", [(10, 1), (10, 3), (7, 3), (5, 2), (12, 5), (3, 10), (1, 1)])
def test_split_matches_array_split(n, size):
    got = [s.get_blocks() for s in split_block_list(_blocklist(n), size)]
    want = [a.tolist() for a in np.array_split([f"b{i}" for i in range(n)], math.ceil(n / size))]
    assert got == want


def test_split_by_bytes():
    bl = BlockList([f"b{i}" for i in range(6)], [_meta(1, s) for s in (10, 10, 25, 5, 5, 5)])
    assert [s.get_blocks() for s in split_block_list_by_bytes(bl, 20)] == [
        ["b0", "b1"], ["b2"], ["b3", "b4", "b5"]]           # the 25-byte block gets its own window
    assert [len(s.get_blocks()) for s in split_block_list_by_bytes(bl, 1)] == [1] * 6
    bl = BlockList(["x"], [_meta(1, None)])
    with pytest.raises(RuntimeError):
        split_block_list_by_bytes(bl, 10)


@pytest.mark.reference
def test_split_matches_reference():
    from oracle import ref_shim

    ref = ref_shim.load()
    for n, size in [(10, 3), (7, 3), (12, 5), (9, 4)]:
        names = [f"b{i}" for i in range(n)]
        metas = [_meta(1, 3 + i % 4) for i in range(n)]
        rbl = ref.BlockList(list(names), list(metas), owned_by_consumer=False)
        mine = BlockList(list(names), list(metas))
        assert [s.get_blocks() for s in split_block_list(mine, size)] == [
            s._blocks for s in rbl.split(size)]
        assert [s.get_blocks() for s in split_block_list_by_bytes(mine, 9)] == [
            s._blocks for s in rbl.split_by_bytes(9)]


# ------------------------------------------------------------------ windows and epochs
def test_window_repeat_rewindow_bookkeeping():
    ds = Dataset(_blocklist(10))
    pipe = ds.window(blocks_per_window=3)
    assert repr(pipe) == "DatasetPipeline(num_windows=4, num_stages=1)"
    assert [_tokens(d) for d in pipe.iter_datasets()] == [
        ["b0", "b1", "b2"], ["b3", "b4", "b5"], ["b6", "b7"], ["b8", "b9"]]
    with pytest.raises(RuntimeError):
        pipe.iter_datasets()
    with pytest.raises(RuntimeError):
        pipe.foreach_window(lambda d: d)        # derived pipelines share the flag

    assert [len(_tokens(d)) for d in ds.window().iter_datasets()] == [10]   # default 10 blocks
    with pytest.raises(ValueError):
        ds.window(blocks_per_window=2, bytes_per_window=2)
    assert [len(_tokens(d)) for d in ds.window(bytes_per_window=24).iter_datasets()] == [3, 3, 3, 1]

    pipe = ds.window(blocks_per_window=5).repeat(3)
    assert repr(pipe) == "DatasetPipeline(num_windows=6, num_stages=1)"
    got = [(d._epoch, _tokens(d)[0]) for d in pipe.iter_datasets()]
    assert got == [(0, "b0"), (0, "b5"), (1, "b0"), (1, "b5"), (2, "b0"), (2, "b5")]

    with pytest.raises(ValueError):
        ds.repeat(0)
    with pytest.raises(ValueError):
        ds.repeat().repeat()
    pipe = ds.repeat(2).rewindow(blocks_per_window=4)
    got = [(d._epoch, len(_tokens(d))) for d in pipe.iter_datasets()]
    assert got == [(0, 4), (0, 4), (0, 2), (1, 4), (1, 4), (1, 2)]          # never mixes epochs
    pipe = ds.repeat(2).rewindow(blocks_per_window=4, preserve_epoch=False)
    assert [len(_tokens(d)) for d in pipe.iter_datasets()] == [4, 4, 4, 4, 4]


def test_iter_epochs_lazy_order():
    calls = []
    ds = Dataset(_blocklist(4))
    pipe = (ds.window(blocks_per_window=2)
            .foreach_window(lambda d: (calls.append("pre"), d)[1])
            .repeat(3)
            .foreach_window(lambda d: (calls.append(("post", d._epoch)), d)[1]))
    out = []
    for ep in pipe.iter_epochs():
        out.append([(d._epoch, _tokens(d)[0]) for d in ep.iter_datasets()])
    assert out == [[(0, "b0"), (0, "b2")], [(1, "b0"), (1, "b2")], [(2, "b0"), (2, "b2")]]
    assert calls.count("pre") == 2                                   # before repeat(): once per window
    assert [c for c in calls if c != "pre"] == [("post", e) for e in (0, 0, 1, 1, 2, 2)]
    # a consumer that abandons an epoch early still gets the next epoch from its start
    pipe = Dataset(_blocklist(4)).window(blocks_per_window=1).repeat(2)
    firsts = []
    for ep in pipe.iter_epochs():
        firsts.append(next(iter(ep.iter_datasets()))._epoch)
    assert firsts == [0, 1]
    assert len(list(Dataset(_blocklist(4)).window(blocks_per_window=1).repeat(5).iter_epochs(2))) == 2


def test_from_iterable_length():
    p = DatasetPipeline.from_iterable([lambda: Dataset(_blocklist(1)), lambda: Dataset(_blocklist(2))])
    assert repr(p) == "DatasetPipeline(num_windows=2, num_stages=1)"
    assert [len(_tokens(d)) for d in p.iter_datasets()] == [1, 2]
    gen = (f for f in [lambda: Dataset(_blocklist(1))])
    assert repr(DatasetPipeline.from_iterable(gen)) == "DatasetPipeline(num_windows=None, num_stages=1)"


# ------------------------------------------------------------------ stats text
def _stats_chain(cls_stats, builder_of):
    src = cls_stats(stages={"from_blocks": [_meta(100, 800) for _ in range(4)]}, parent=None)
    src.dataset_uuid = "u0"
    b = builder_of(src, "sort")
    s1 = b.build_multistage({
        "map": [_meta(100, 800, wall=0.002 * (i + 1)) for i in range(4)],
        "reduce": [_meta(r, 8 * r, wall=1.5) for r in (90, 110, 100, 100)],
    })
    s1.dataset_uuid, s1.time_total_s = "u1", 3.14159
    b = builder_of(s1, "map_batches")
    s2 = b.build(BlockList(["x", "y"], [_meta(5, 40, wall=0.00001), _meta(7, 56, wall=0.0002)]))
    s2.dataset_uuid, s2.time_total_s = "u2", 0.004
    return s2


def test_stats_summary_text():
    s = _stats_chain(DatasetStats, lambda parent, name: parent.child_builder(name)).summary_string()
    assert s == (
        "Stage 0 from_blocks: 4/4 blocks split from parent in 0s\n"
        "* Output num rows: 100 min, 100 max, 100 mean, 400 total\n"
        "* Output size bytes: 800 min, 800 max, 800 mean, 3200 total\n"
        "\n"
        "Stage 1 sort: executed in 3.14s\n"
        "\n"
        "\tSubstage 0 sort_map: 4/4 blocks executed\n"
        "\t* Remote wall time: 2.0ms min, 8.0ms max, 5.0ms mean, 20.0ms total\n"
        "\t* Remote cpu time: 1000.0us min, 4.0ms max, 2.5ms mean, 10.0ms total\n"
        "\t* Peak heap memory usage (MiB): 300.0 min, 300.0 max, 300 mean\n"
        "\t* Output num rows: 100 min, 100 max, 100 mean, 400 total\n"
        "\t* Output size bytes: 800 min, 800 max, 800 mean, 3200 total\n"
        "\t* Tasks per node: 4 min, 4 max, 4 mean; 1 nodes used\n"
        "\n"
        "\tSubstage 1 sort_reduce: 4/4 blocks executed\n"
        "\t* Remote wall time: 1.5s min, 1.5s max, 1.5s mean, 6.0s total\n"
        "\t* Remote cpu time: 750.0ms min, 750.0ms max, 750.0ms mean, 3.0s total\n"
        "\t* Peak heap memory usage (MiB): 300.0 min, 300.0 max, 300 mean\n"
        "\t* Output num rows: 90 min, 110 max, 100 mean, 400 total\n"
        "\t* Output size bytes: 720 min, 880 max, 800 mean, 3200 total\n"
        "\t* Tasks per node: 4 min, 4 max, 4 mean; 1 nodes used\n"
        "\n"
        "Stage 2 map_batches: 2/2 blocks executed in 0s\n"
        "* Remote wall time: 10.0us min, 200.0us max, 105.0us mean, 210.0us total\n"
        "* Remote cpu time: 5.0us min, 100.0us max, 52.5us mean, 105.0us total\n"
        "* Peak heap memory usage (MiB): 300.0 min, 300.0 max, 300 mean\n"
        "* Output num rows: 5 min, 7 max, 6 mean, 12 total\n"
        "* Output size bytes: 40 min, 56 max, 48 mean, 96 total\n"
        "* Tasks per node: 2 min, 2 max, 2 mean; 1 nodes used\n"
    )


@pytest.mark.reference
def test_stats_summary_matches_reference():
    import importlib

    from oracle import ref_shim

    ref_shim.load()
    rs = importlib.import_module("ray.data._internal.stats")
    ref = _stats_chain(rs.DatasetStats, lambda parent, name: parent.child_builder(name)).summary_string()
    mine = _stats_chain(DatasetStats, lambda parent, name: DatasetStatsBuilder(name, parent)).summary_string()
    assert mine == ref


# ------------------------------------------------------------------ splitter selection
@pytest.mark.parametrize("w", [1, 2, 8, 9, 10, 16, 20])
@pytest.mark.parametrize("spread", [3, 256])
def test_fsb_order_statistics_equal_sorted_picks(w, spread):
    """the multi-GPU plan picks its W-1 splitters by selection, not by sorting the samples:
    the picks must be the same rows a full memcmp sort would give (heavy ties included)."""
    from ray_memory_management_b200.sort import fsb_order, fsb_order_statistics

    rng = np.random.default_rng(w * 1000 + spread)
    rows = rng.integers(0, spread, (3000, w), dtype=np.uint8)
    srt = sorted(r.tobytes() for r in rows)
    picks = [0, 1, 374, 1499, 1500, 2999, 374]
    assert fsb_order_statistics(rows, picks) == [srt[i] for i in picks]
    order = fsb_order(rows)
    assert [rows[i].tobytes() for i in order] == srt
    assert all(order[i] < order[i + 1] for i in range(2999) if srt[i] == srt[i + 1])   # stable
