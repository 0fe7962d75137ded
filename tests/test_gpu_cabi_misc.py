"""GPU: direct C-ABI checks of the smaller entry points against numpy."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from ray_memory_management_b200 import _lib

    return _lib


def _ws(b):
    return torch.empty(b, dtype=torch.uint8, device="cuda")


def _stream():
    return torch.cuda.current_stream().cuda_stream


@pytest.mark.parametrize("desc", [False, True])
def test_searchsorted_i64_f64_fsb(lib, desc):
    rng = np.random.default_rng(1)
    col = np.sort(rng.integers(-50, 50, 10_000, dtype=np.int64))
    bnd = np.array([-60, -50, -7, -7, 0, 13, 49, 50, 70], dtype=np.int64)
    if desc:
        col, bnd = col[::-1].copy(), bnd[::-1].copy()
    n = len(col)
    want = (n - np.searchsorted(col, bnd, sorter=np.arange(n - 1, -1, -1))) if desc else np.searchsorted(col, bnd)
    for kind, c, b in (("i64", col, bnd), ("f64", col.astype(np.float64), bnd.astype(np.float64))):
        dc, db = torch.from_numpy(c).cuda(), torch.from_numpy(b).cuda()
        out = torch.empty(len(b), dtype=torch.int64, device="cuda")
        lib.call(f"ds_searchsorted_{kind}", dc.data_ptr(), n, db.data_ptr(), len(b), int(desc), out.data_ptr(), _stream())
        assert np.array_equal(out.cpu().numpy(), want), kind
    # fixed_size_binary(3): same values as big-endian bytes
    def enc(a):
        u = (a + 100).astype(">u4").view(np.uint8).reshape(-1, 4)[:, 1:]
        return np.ascontiguousarray(u)
    dc, db = torch.from_numpy(enc(col).reshape(-1)).cuda(), torch.from_numpy(enc(bnd).reshape(-1)).cuda()
    out = torch.empty(len(bnd), dtype=torch.int64, device="cuda")
    lib.call("ds_searchsorted_fsb", dc.data_ptr(), 3, n, db.data_ptr(), len(bnd), int(desc), out.data_ptr(), _stream())
    assert np.array_equal(out.cpu().numpy(), want)


@pytest.mark.parametrize("desc", [False, True])
@pytest.mark.parametrize("n", [0, 1, 4095, 4097, 200_003])
def test_partition_hist_scatter_is_stable(lib, desc, n):
    rng = np.random.default_rng(n + desc)
    col = rng.integers(-100, 100, n, dtype=np.int64)
    bnd = np.array([-100, -20, -20, 5, 99, 100], dtype=np.int64)
    if desc:
        bnd = bnd[::-1].copy()
    nb = len(bnd)
    bucket = np.array([(bnd > x).sum() if desc else (bnd <= x).sum() for x in col], dtype=np.int64) if n else np.zeros(0, np.int64)
    dc, db = torch.from_numpy(col).cuda(), torch.from_numpy(bnd).cuda()
    counts = torch.empty(nb + 1, dtype=torch.int64, device="cuda")
    offsets = torch.empty(nb + 1, dtype=torch.int64, device="cuda")
    perm = torch.empty(max(n, 1), dtype=torch.int32, device="cuda")
    ws = lib.call_ws("ds_partition_hist_scatter_i64", _ws, dc.data_ptr(), n, db.data_ptr(), nb, int(desc),
                     counts.data_ptr(), offsets.data_ptr(), perm.data_ptr(), stream=_stream())
    torch.cuda.synchronize()
    del ws
    want_counts = np.bincount(bucket, minlength=nb + 1)
    assert np.array_equal(counts.cpu().numpy(), want_counts)
    assert np.array_equal(offsets.cpu().numpy(), np.concatenate([[0], np.cumsum(want_counts)[:-1]]))
    assert np.array_equal(perm.cpu().numpy()[:n], np.argsort(bucket, kind="stable"))


def test_partition_too_many_parts(lib):
    need = ctypes.c_size_t(0)
    st = lib.lib.ds_partition_hist_scatter_i64(None, 10, None, 256, 0, None, None, None, None, ctypes.byref(need), None)
    assert st == lib.DS_ERR_UNSUPPORTED


def test_sample_rows_without_replacement(lib):
    n, w = 10_000, 10
    col = np.arange(n * w, dtype=np.int64).astype(np.uint8).reshape(n, w)
    col[:, :4] = np.arange(n, dtype="<u4").view(np.uint8).reshape(n, 4)  # row id in the first 4 bytes
    dc = torch.from_numpy(col.reshape(-1)).cuda()
    for k in (1, 10, 333, n):
        out = torch.empty(k * w, dtype=torch.uint8, device="cuda")
        outs = []
        for seed in (5, 5, 6):
            lib.call("ds_sample_rows", dc.data_ptr(), w, n, k, seed, out.data_ptr(), _stream())
            rows = out.cpu().numpy().reshape(k, w)
            ids = rows[:, :4].copy().view("<u4").ravel()
            assert len(np.unique(ids)) == k and np.array_equal(rows, col[ids])  # distinct, intact rows
            assert np.all(np.diff(ids.astype(np.int64)) > 0)  # one per stratum, in order
            outs.append(ids.copy())
        assert np.array_equal(outs[0], outs[1])
        if 1 < k < n:
            assert not np.array_equal(outs[0], outs[2])
    assert lib.lib.ds_sample_rows(dc.data_ptr(), w, 5, 6, 0, dc.data_ptr(), None) == lib.DS_ERR_INVALID_ARG


@pytest.mark.parametrize("n,groups", [(1, 1), (2048, 1), (2049, 3), (50_000, 50_000), (300_000, 17)])
def test_segment_heads_and_reductions(lib, n, groups):
    rng = np.random.default_rng(n)
    keys = np.sort(rng.integers(0, groups, n, dtype=np.int64))
    vals = rng.standard_normal(n)
    valid = (rng.random(n) < 0.8).astype(np.uint8)
    perm = rng.permutation(n).astype(np.int32)         # values stay in "input order"
    vin = np.empty(n); vin[perm] = vals
    valin = np.empty(n, np.uint8); valin[perm] = valid
    dk = torch.from_numpy(keys).cuda()
    starts = torch.empty(n + 1, dtype=torch.int32, device="cuda")
    nseg = torch.zeros(1, dtype=torch.int32, device="cuda")
    ws = lib.call_ws("ds_segment_heads_i64", _ws, dk.data_ptr(), n, starts.data_ptr(), nseg.data_ptr(), stream=_stream())
    k = int(nseg.item())
    ref_starts = np.concatenate([[0], np.flatnonzero(keys[1:] != keys[:-1]) + 1, [n]])
    assert k == len(ref_starts) - 1 and np.array_equal(starts[: k + 1].cpu().numpy(), ref_starts)
    dv, dva, dp = torch.from_numpy(vin).cuda(), torch.from_numpy(valin).cuda(), torch.from_numpy(perm).cuda()
    out = torch.empty(k, dtype=torch.float64, device="cuda")
    masked = np.where(valid != 0, vals, 0.0)
    for op, fn, ident in ((lib.DS_OP_SUM, np.add, 0.0), (lib.DS_OP_MIN, np.minimum, np.inf),
                          (lib.DS_OP_MAX, np.maximum, -np.inf), (lib.DS_OP_ABSMAX, np.maximum, 0.0)):
        src = np.where(valid != 0, np.abs(vals) if op == lib.DS_OP_ABSMAX else vals, ident)
        want = fn.reduceat(src, ref_starts[:-1])
        ws = lib.call_ws("ds_segmented_reduce_f64", _ws, dv.data_ptr(), dva.data_ptr(), dp.data_ptr(),
                         starts.data_ptr(), k, n, op, out.data_ptr(), stream=_stream())
        assert np.allclose(out.cpu().numpy(), want, rtol=1e-9, atol=1e-9), op
    cnt = torch.empty(k, dtype=torch.int64, device="cuda")
    ws = lib.call_ws("ds_segmented_count_valid", _ws, dva.data_ptr(), dp.data_ptr(), starts.data_ptr(), k, n,
                     cnt.data_ptr(), stream=_stream())
    assert np.array_equal(cnt.cpu().numpy(), np.add.reduceat(valid.astype(np.int64), ref_starts[:-1]))
    lib.call("ds_segment_counts", starts.data_ptr(), k, cnt.data_ptr(), _stream())
    assert np.array_equal(cnt.cpu().numpy(), np.diff(ref_starts))
    # reproducible: the same call gives the same bits
    a = torch.empty(k, dtype=torch.float64, device="cuda")
    ws = lib.call_ws("ds_segmented_reduce_f64", _ws, dv.data_ptr(), None, dp.data_ptr(), starts.data_ptr(), k, n,
                     lib.DS_OP_SUM, a.data_ptr(), stream=_stream())
    b = torch.empty(k, dtype=torch.float64, device="cuda")
    ws = lib.call_ws("ds_segmented_reduce_f64", _ws, dv.data_ptr(), None, dp.data_ptr(), starts.data_ptr(), k, n,
                     lib.DS_OP_SUM, b.data_ptr(), stream=_stream())
    assert bool((a == b).all())


@pytest.mark.parametrize("width", [90, 18, 250, 1022])
def test_gather_rows_multi_matches_numpy(lib, width):
    """ds_gather_rows_multi: the source column lies in several buffers of ragged sizes (one of
    them empty, one a single row, one starting at an odd multiple of 2 bytes); idx counts rows
    over their concatenation, incl. the first and last row of every buffer (guarded edges)."""
    rng = np.random.default_rng(width)
    sizes = [1000, 1, 0, 4097, 333, 2]
    parts = [rng.integers(0, 256, (m, width), dtype=np.uint8) for m in sizes]
    allrows = np.concatenate(parts)
    devs, ptrs = [], []
    for k, p in enumerate(parts):
        pad = 2 if k == 3 else 0  # a source that is only 2-byte aligned
        t = torch.from_numpy(np.concatenate([np.zeros(pad, np.uint8), p.reshape(-1)])).cuda()
        devs.append(t)
        ptrs.append(t.data_ptr() + pad)
    n_out = 20_011
    idx = rng.integers(0, len(allrows), n_out, dtype=np.int64)
    edges = np.cumsum([0] + sizes)
    firsts_lasts = [int(x) for b in range(len(sizes)) if sizes[b] for x in (edges[b], edges[b + 1] - 1)]
    idx[: len(firsts_lasts)] = firsts_lasts
    idx_dev = torch.from_numpy(idx.astype(np.int32)).cuda()
    out = torch.zeros(n_out * width + 32, dtype=torch.uint8, device="cuda")
    live = [(p, m) for p, m in zip(ptrs, sizes) if m]
    srcs = (ctypes.c_void_p * len(live))(*[p for p, _ in live])
    rows = (ctypes.c_uint64 * len(live))(*[m for _, m in live])
    lib.call("ds_gather_rows_multi", srcs, rows, len(live), width, idx_dev.data_ptr(), n_out, out.data_ptr(), _stream())
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert np.array_equal(got[: n_out * width].reshape(n_out, width), allrows[idx])
    assert not got[n_out * width:].any()


def test_gather_rows_multi_rejects_other_shapes(lib):
    z = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    i = torch.zeros(4, dtype=torch.int32, device="cuda")
    srcs = (ctypes.c_void_p * 2)(z.data_ptr(), z.data_ptr())
    rows = (ctypes.c_uint64 * 2)(4, 4)
    for width in (16, 17, 8, 96):
        assert lib.lib.ds_gather_rows_multi(srcs, rows, 2, width, i.data_ptr(), 4, z.data_ptr(), None) == lib.DS_ERR_UNSUPPORTED
