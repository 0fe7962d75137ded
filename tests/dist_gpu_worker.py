"""Worker for tests/test_gpu_distributed.py: run under torchrun, one rank per GPU.
Checks the multi-GPU sort / random_shuffle / groupby plans against the oracle."""
import os
import sys

import numpy as np
import pyarrow as pa
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import datasets_oracle as O  # noqa: E402


def gather_tables(t: pa.Table, rank, world):
    """all ranks' tables on every rank (pickled through all_gather_object)."""
    sink = pa.BufferOutputStream()
    with pa.ipc.new_stream(sink, t.schema) as w:
        w.write_table(t)
    out = [None] * world
    dist.all_gather_object(out, sink.getvalue().to_pybytes())
    return [pa.ipc.open_stream(b).read_all() for b in out]


def tables_equal(a: pa.Table, b: pa.Table) -> bool:
    """Table.equals with NaN == NaN (Arrow's equality treats NaNs as different)."""
    if a.schema != b.schema or a.num_rows != b.num_rows:
        return False
    for name in a.column_names:
        x, y = a[name].combine_chunks(), b[name].combine_chunks()
        if pa.types.is_floating(x.type):
            if not x.is_valid().equals(y.is_valid()):
                return False
            if not np.array_equal(x.fill_null(0).to_numpy(), y.fill_null(0).to_numpy(), equal_nan=True):
                return False
        elif not x.equals(y):
            return False
    return True


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    import ray_memory_management_b200 as ds
    from ray_memory_management_b200 import distributed as D

    # ---- sort, GraySort records with heavy ties, ragged per-rank sizes
    n = 200_000 + 1000 * rank
    t = O.graysort_table(n, seed=100 + rank)
    keys = O.fsb_to_numpy(t["key"]).copy()
    keys[:, :7] = 3
    keys[:, 7] &= 0x0F
    payload = O.fsb_to_numpy(t["payload"]).copy()
    payload[:, 8] = rank  # make rows of different ranks distinguishable
    t = pa.table({"key": O.fsb_from_numpy(keys), "payload": O.fsb_from_numpy(payload)})
    inputs = gather_tables(t, rank, world)
    for desc in (False, True):
        out = D.sort_block(ds.DeviceArrowBlock.from_arrow(t), "key", descending=desc)
        outs = gather_tables(out.to_arrow(), rank, world)
        got = pa.concat_tables([o for o in outs if o.num_rows])
        want = O.global_stable_sort(inputs, "key", desc)
        assert got.equals(want), f"distributed sort mismatch desc={desc}"
        sizes = [o.num_rows for o in outs]
        assert max(sizes) < 1.3 * sum(sizes) / world, f"unbalanced partitions {sizes}"

    # ---- the push-based plan: key sub-ranges travel in rounds; several output blocks per rank
    if os.environ.get("DS_EXCHANGE", "split") == "split":
        inputs = gather_tables(t, rank, world)
        # (rounds, staged): staged rounds wait in a local buffer and are shipped by the copy engines
        for rounds, staged in ((2, None), (2, 0), (3, None), (3, 1)):
            for desc in (False, True):
                outs, stats = D.sort_block_push(ds.DeviceArrowBlock.from_arrow(t), "key", descending=desc,
                                                rounds=rounds, staged=staged)
                assert sorted(stats) == ["map", "merge", "reduce"]
                assert len(outs) == rounds == len(stats["merge"]) == len(stats["reduce"])
                mine = pa.concat_tables([o.to_arrow() for o in outs])
                got = pa.concat_tables([o for o in gather_tables(mine, rank, world) if o.num_rows])
                assert got.equals(O.global_stable_sort(inputs, "key", desc)), f"push plan mismatch r={rounds} staged={staged} desc={desc}"
                assert sum(m.num_rows for m in stats["reduce"]) == mine.num_rows
                assert all(m.exec_stats.wall_time_s >= 0 for ms in stats.values() for m in ms)
        # the reference's flag selects the plan inside sort_block (one concatenated block out)
        from ray_memory_management_b200.context import DatasetContext

        DatasetContext._set_current(DatasetContext(use_push_based_shuffle=True))
        out = D.sort_block(ds.DeviceArrowBlock.from_arrow(t), "key")
        DatasetContext._set_current(DatasetContext())
        got = pa.concat_tables([o for o in gather_tables(out.to_arrow(), rank, world) if o.num_rows])
        assert got.equals(O.global_stable_sort(inputs, "key", False)), "use_push_based_shuffle mismatch"

    # ---- int64 sort
    rng = np.random.default_rng(rank)
    ti = pa.table({"k": rng.integers(-10**6, 10**6, 150_000, dtype=np.int64),
                   "v": np.arange(150_000, dtype=np.int64) + rank * 10**7})
    inputs = gather_tables(ti, rank, world)
    out = D.sort_block(ds.DeviceArrowBlock.from_arrow(ti), "k")
    got = pa.concat_tables(gather_tables(out.to_arrow(), rank, world))
    assert got.equals(O.global_stable_sort(inputs, "k", False)), "int64 distributed sort mismatch"

    # ---- float64 keys with NaNs, a narrow (int32) key, and one rank without rows
    nf = 0 if rank == world - 1 else 40_000 + 77 * rank
    kf = rng.random(nf) * 100
    kf[rng.random(nf) < 0.05] = np.nan
    tf = pa.table({"k": kf, "i": rng.integers(-50, 50, nf).astype(np.int32),
                   "r": np.arange(nf, dtype=np.int64) + rank * 10**6})
    inputs_f = gather_tables(tf, rank, world)
    for col in ("k", "i"):
        for desc in (False, True):
            out = D.sort_block(ds.DeviceArrowBlock.from_arrow(tf), col, descending=desc)
            got = pa.concat_tables(gather_tables(out.to_arrow(), rank, world))
            want = O.global_stable_sort(inputs_f, col, desc)
            assert tables_equal(got, want), f"distributed sort mismatch on {col!r} desc={desc}"

    # ---- sort with null keys and null payload values on some ranks only
    kn = rng.integers(-1000, 1000, 50_000, dtype=np.int64)
    tn = pa.table({"k": pa.array(kn, mask=(rng.random(50_000) < 0.1) if rank % 2 == 0 else None),
                   "v": pa.array(rng.random(50_000), mask=(rng.random(50_000) < 0.2) if rank == 0 else None),
                   "r": np.arange(50_000, dtype=np.int64) + rank * 10**6})
    inputs_n = gather_tables(tn, rank, world)
    for desc in (False, True):
        out = D.sort_block(ds.DeviceArrowBlock.from_arrow(tn), "k", descending=desc)
        got = pa.concat_tables(gather_tables(out.to_arrow(), rank, world))
        want = O.global_stable_sort(inputs_n, "k", desc)
        assert got.equals(want), f"distributed sort with nulls mismatch desc={desc}"

    # ---- variable-length payload columns (strings, with nulls on one rank) through sort and shuffle
    words = ["r%d-%d-" % (rank, i) + "x" * int(i % 7) for i in range(30_000 + 500 * rank)]
    smask = (rng.random(len(words)) < 0.1) if rank == 1 else None
    ts = pa.table({"k": rng.integers(-500, 500, len(words), dtype=np.int64),
                   "s": pa.array(words, pa.string(), mask=smask),
                   "b": pa.array([w.encode()[::-1] for w in words], pa.large_binary())})
    inputs_s = gather_tables(ts, rank, world)
    for desc in (False, True):
        out = D.sort_block(ds.DeviceArrowBlock.from_arrow(ts), "k", descending=desc)
        got = pa.concat_tables(gather_tables(out.to_arrow(), rank, world))
        want = O.global_stable_sort(inputs_s, "k", desc)
        assert got.equals(want), f"distributed sort with string payload mismatch desc={desc}"
    sh = D.random_shuffle_block(ds.DeviceArrowBlock.from_arrow(ts), 7)
    gs = pa.concat_tables(gather_tables(sh.to_arrow(), rank, world))
    alls = pa.concat_tables(inputs_s)
    assert gs.num_rows == alls.num_rows
    key = lambda tb: sorted(zip(tb["b"].to_pylist(), tb["k"].to_pylist(), tb["s"].to_pylist()),  # noqa: E731
                            key=lambda r: r[0])
    assert key(gs) == key(alls), "shuffle lost or mixed up variable-length rows"
    # variable-length KEYS (string with nulls on one rank, large_binary): splitters agreed on
    # through the host, the key column rebuilt after the exchange (reference: test_sort.py:55-97)
    for col in ("s", "b"):
        for desc in (False, True):
            out = D.sort_block(ds.DeviceArrowBlock.from_arrow(ts), col, descending=desc)
            got = pa.concat_tables(gather_tables(out.to_arrow(), rank, world))
            want = O.global_stable_sort(inputs_s, col, desc)
            assert got.equals(want), f"distributed sort on a variable-length key {col!r} mismatch desc={desc}"

    # ---- random_shuffle: permutation + determinism
    a = D.random_shuffle_block(ds.DeviceArrowBlock.from_arrow(ti), 42)
    b = D.random_shuffle_block(ds.DeviceArrowBlock.from_arrow(ti), 42)
    c = D.random_shuffle_block(ds.DeviceArrowBlock.from_arrow(ti), 43)
    ga = pa.concat_tables(gather_tables(a.to_arrow(), rank, world))
    gb = pa.concat_tables(gather_tables(b.to_arrow(), rank, world))
    gc = pa.concat_tables(gather_tables(c.to_arrow(), rank, world))
    assert ga.equals(gb) and not ga.equals(gc)
    allv = np.sort(pa.concat_tables(inputs)["v"].to_numpy())
    assert np.array_equal(np.sort(ga["v"].to_numpy()), allv)
    # rows really moved between ranks
    mine = a.to_arrow()["v"].to_numpy() // 10**7
    assert len(np.unique(mine)) == world

    # ---- groupby
    tg = pa.table({"k": rng.integers(0, 5000, 120_000, dtype=np.int64), "v": rng.random(120_000),
                   "w": rng.integers(-100, 100, 120_000, dtype=np.int64)})
    inputs = gather_tables(tg, rank, world)
    aggs = (ds.Sum("v"), ds.Mean("v"), ds.Count(), ds.Sum("w"), ds.Max("w"))
    out = D.groupby_block(ds.DeviceArrowBlock.from_arrow(tg), "k", aggs)
    got = pa.concat_tables([o for o in gather_tables(out.to_arrow(), rank, world) if o.num_rows])
    want = O.groupby_global(inputs, "k", (("sum", "v"), ("mean", "v"), ("count", None), ("sum", "w"), ("max", "w")))
    assert got.column_names == want.column_names
    for name in want.column_names:
        x, y = got[name].to_numpy(), want[name].to_numpy()
        if y.dtype.kind == "f":
            assert np.allclose(x, y, rtol=1e-6, atol=0), name
        else:
            assert np.array_equal(x, y), name
    # ---- groupby with a narrow key, narrow value columns and Std on an integer column
    # (partial columns of different widths and types on the wire)
    tg2 = pa.table({"k": rng.integers(0, 300, 50_000).astype(np.int32),
                    "a": rng.integers(-1000, 1000, 50_000, dtype=np.int64),
                    "f": rng.random(50_000).astype(np.float32),
                    "s": rng.integers(-5, 5, 50_000).astype(np.int32)})
    inputs2 = gather_tables(tg2, rank, world)
    out = D.groupby_block(ds.DeviceArrowBlock.from_arrow(tg2), "k",
                          (ds.Std("a"), ds.Sum("f"), ds.Mean("s"), ds.Min("f"), ds.Max("s")))
    got = pa.concat_tables([o for o in gather_tables(out.to_arrow(), rank, world) if o.num_rows]).to_pandas()
    g = pa.concat_tables(inputs2).to_pandas().groupby("k")
    assert np.array_equal(got["k"].to_numpy(), np.array(sorted(g.groups), dtype=got["k"].dtype))
    assert np.allclose(got["std(a)"].to_numpy(), g["a"].std(ddof=1).to_numpy(), rtol=1e-6)
    assert np.allclose(got["sum(f)"].to_numpy(), g["f"].apply(lambda x: x.astype(np.float64).sum()).to_numpy(), rtol=1e-6)
    assert np.allclose(got["mean(s)"].to_numpy(), g["s"].mean().to_numpy(), rtol=1e-6)
    assert np.array_equal(got["min(f)"].to_numpy().astype(np.float32), g["f"].min().to_numpy())
    assert np.array_equal(got["max(s)"].to_numpy(), g["s"].max().to_numpy())
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        print("DIST_GPU_OK", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
