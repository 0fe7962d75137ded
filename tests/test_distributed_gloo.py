"""CPU, world_size 2 over gloo: the host logic of the multi-GPU plan -- identical
splitters from all-gathered samples, the W x W count matrix, split sizes and the rank-order
concatenation of the all-to-all-v -- with numpy standing in (in this TEST only) for the
CUDA partition and sort steps.  The product path has no such stand-in."""
import os
import socket

import numpy as np
import pyarrow as pa
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import datasets_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _np_bucket(keys_be, boundaries, descending):
    """numpy stand-in for ds_partition_hist_scatter_fsb (rule of arrow_block.py:373-394)."""
    kb = [bytes(r) for r in keys_be]
    out = np.zeros(len(kb), dtype=np.int64)
    for i, k in enumerate(kb):
        if descending:
            out[i] = sum(1 for b in boundaries if b > k)
        else:
            out[i] = sum(1 for b in boundaries if b <= k)
    return out


def _worker(rank, world, port, descending, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from ray_memory_management_b200 import distributed as D

    n = 3000 + 500 * rank  # ragged
    t = O.graysort_table(8000, seed=4).slice(3000 * rank, n)
    keys = O.fsb_to_numpy(t["key"]).copy()
    keys[:, :9] = 1
    keys[:, 9] &= 0x1F  # heavy ties across both ranks
    payload = O.fsb_to_numpy(t["payload"]).copy()

    rng = np.random.default_rng(rank)
    samp = keys[rng.choice(n, 64, replace=False)]
    samples = D.gather_samples(samp, 10, None, None)
    assert samples.shape == (64 * world, 10)
    bounds = D.splitters_from_samples(samples, pa.binary(10), world)
    if descending:
        bounds.reverse()
    bucket = _np_bucket(keys, bounds, descending)
    perm = np.argsort(bucket, kind="stable")
    counts = [int((bucket == d).sum()) for d in range(world)]
    matrix = D.exchange_counts(counts, None, None)
    assert matrix.shape == (world, world) and list(matrix[rank]) == counts
    rk, _ = D.alltoallv_rows(torch.from_numpy(keys[perm].reshape(-1).copy()), 10, matrix)
    rp, work = D.alltoallv_rows(torch.from_numpy(payload[perm].reshape(-1).copy()), 90, matrix, async_op=True)
    work.wait()
    rk = rk.numpy().reshape(-1, 10)
    rp = rp.numpy().reshape(-1, 90)
    assert rk.shape[0] == int(matrix[:, rank].sum())
    tab = pa.table({"key": O.fsb_from_numpy(rk), "payload": O.fsb_from_numpy(rp)})
    out = O.sort_table(tab, "key", descending) if tab.num_rows else tab
    q.put((rank, keys, payload, O.fsb_to_numpy(out["key"]).copy(), O.fsb_to_numpy(out["payload"]).copy(),
           [bytes(b) for b in bounds]))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("descending", [False, True])
def test_two_rank_exchange_reproduces_global_stable_sort(descending):
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, descending, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][5] == res[1][5]  # identical splitters on every rank
    tabs = [pa.table({"key": O.fsb_from_numpy(r[1]), "payload": O.fsb_from_numpy(r[2])}) for r in res]
    want = O.global_stable_sort(tabs, "key", descending)
    got_k = np.concatenate([r[3] for r in res])
    got_p = np.concatenate([r[4] for r in res])
    assert np.array_equal(got_k, O.fsb_to_numpy(want["key"]))
    assert np.array_equal(got_p, O.fsb_to_numpy(want["payload"]))  # ties keep (rank, row) order


def _varlen_worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from ray_memory_management_b200 import distributed as D
    from ray_memory_management_b200.block import DeviceVarColumn

    # rows of this rank in send order: partition 0 first, then partition 1 (ragged, one empty string)
    rows = [b"r%d-%d" % (rank, i) + b"x" * (i % 5) for i in range(7 + 3 * rank)] + [b""]
    counts = [3 + rank, len(rows) - 3 - rank]
    offs = np.concatenate([[0], np.cumsum([len(r) for r in rows])]).astype(np.int64) + 11   # a slice: offsets start at 11
    data = np.frombuffer(b"\xee" * 11 + b"".join(rows), dtype=np.uint8).copy()
    col = DeviceVarColumn(pa.binary(), torch.from_numpy(offs), torch.from_numpy(data))
    len_wire, payload, byte_counts = D._varlen_send(col, np.concatenate([[0], np.cumsum(counts)]), "cpu")
    assert byte_counts == [sum(len(r) for r in rows[: counts[0]]), sum(len(r) for r in rows[counts[0]:])]
    matrix = D.exchange_counts(counts, None, None)
    lens_recv, _ = D.alltoallv_rows(len_wire.clone(), 8, matrix)
    n_recv = int(matrix[:, rank].sum())
    got = D._varlen_recv(pa.binary(), lens_recv, n_recv, payload, byte_counts, None, None, "cpu")
    o, d = got.offsets.numpy(), got.data.numpy().tobytes()
    q.put((rank, rows, counts, [d[o[i]:o[i + 1]] for i in range(n_recv)]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_variable_length_wire():
    """row lengths on a fixed-width wire + byte-granular all-to-all-v rebuild the strings in
    (source rank, row) order on the receiver."""
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_varlen_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for dst in range(world):
        want = []
        for src in range(world):
            rows, counts = res[src][1], res[src][2]
            start = sum(counts[:dst])
            want += rows[start:start + counts[dst]]
        assert res[dst][3] == want


def test_block_without_columns_is_refused_with_the_way_out():
    """the ranks do not exchange schemas (distributed.sort_block docstring): a block without
    columns would desynchronise the collectives, so every multi-rank entry point refuses it on
    the rank that holds it."""
    from ray_memory_management_b200 import distributed as D
    from ray_memory_management_b200.block import DeviceArrowBlock

    with pytest.raises(ValueError, match="no columns.*empty_table"):
        D._require_columns(DeviceArrowBlock.empty(), "sort_block")
