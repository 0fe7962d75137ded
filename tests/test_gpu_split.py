"""GPU: the streaming split (ds_split_plan + ds_split_rows) through the C ABI against numpy.

All destinations are buffers of the one GPU under test: the kernel does not care whether a
destination pointer is local or peer-mapped.  The contract checked is the one the multi-GPU
plans rely on: destination d receives exactly the rows whose key falls into its
range (the rule of ArrowBlockAccessor.sort_and_partition, python/ray/data/_internal/
arrow_block.py:373-394), in INPUT order, densely packed from dst_bases[d] on.
"""
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


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _plan(lib, kind, col_dev, width, valid_dev, n, bnd_dev, nb, desc):
    nparts = nb + 1
    ntiles = (n + lib.DS_SPLIT_TILE_ROWS - 1) // lib.DS_SPLIT_TILE_ROWS
    dest = torch.full((max(n, 1),), 255, dtype=torch.uint8, device="cuda")
    offs = torch.zeros(int(lib.lib.ds_split_table_words(n, nparts)), dtype=torch.int32, device="cuda")
    counts = torch.zeros(nparts, dtype=torch.int64, device="cuda")
    lib.call("ds_split_plan", kind, col_dev.data_ptr() if n else None, width,
             valid_dev.data_ptr() if valid_dev is not None else None, n,
             bnd_dev.data_ptr() if nb else None, nb, int(desc), dest.data_ptr(), offs.data_ptr(),
             counts.data_ptr(), _stream())
    torch.cuda.synchronize()
    table = offs[: nparts * (ntiles + 1)].cpu().numpy().reshape(nparts, ntiles + 1)
    return dest[:n].cpu().numpy(), table, counts.cpu().numpy(), dest, offs


def _bucket_i64(col, bnd, desc):
    if desc:  # count of boundaries strictly greater than the key
        return np.array([(bnd > x).sum() for x in col], dtype=np.int64) if len(col) < 5000 else \
            (len(bnd) - np.searchsorted(np.sort(bnd), col, side="right"))
    return np.searchsorted(bnd, col, side="right")


@pytest.mark.parametrize("desc", [False, True])
@pytest.mark.parametrize("n", [0, 1, 255, 256, 257, 70_001, 2_100_000])
def test_split_plan_i64(lib, desc, n):
    rng = np.random.default_rng(n * 2 + desc)
    col = rng.integers(-100, 100, n, dtype=np.int64)
    bnd = np.array([-100, -20, -20, 5, 99, 100], dtype=np.int64)
    if desc:
        bnd = bnd[::-1].copy()
    valid = (rng.random(n) > 0.1).astype(np.uint8)
    want = _bucket_i64(col, bnd, desc)
    want = np.where(valid != 0, want, len(bnd))
    dcol = torch.from_numpy(col).cuda()
    dbnd = torch.from_numpy(bnd).cuda()
    dval = torch.from_numpy(valid).cuda()
    dest, offs, counts, _, _ = _plan(lib, lib.DS_KEY_I64, dcol, 8, dval, n, dbnd, len(bnd), desc)
    assert np.array_equal(dest.astype(np.int64), want)
    nparts = len(bnd) + 1
    assert np.array_equal(counts, np.bincount(want, minlength=nparts))
    T = lib.DS_SPLIT_TILE_ROWS
    ntiles = (n + T - 1) // T
    for d in range(nparts):
        per_tile = np.array([(want[t * T:(t + 1) * T] == d).sum() for t in range(ntiles)], dtype=np.int64)
        assert np.array_equal(offs[d], np.concatenate([[0], np.cumsum(per_tile)])), d


@pytest.mark.parametrize("width,nparts,n", [
    (90, 8, 100_003), (10, 8, 100_003), (90, 2, 5_000), (8, 5, 33_333), (16, 3, 10_000),
    (7, 4, 9_999), (100, 8, 4_097), (200, 3, 2_049), (198, 5, 3_000), (2, 32, 50_000), (90, 1, 1_000), (24, 8, 255),
])
@pytest.mark.parametrize("offset", [0, 2, 1])
def test_split_rows_matches_numpy(lib, width, nparts, n, offset):
    """rows land densely, in input order, at every destination; destination buffers start at
    odd / even / 16-byte-aligned addresses; nothing is written around them."""
    rng = np.random.default_rng(width * 131 + nparts + n)
    rows = rng.integers(0, 256, (n, width), dtype=np.uint8)
    keys = rng.integers(0, 1 << 40, n, dtype=np.int64)
    bnd = np.sort(rng.integers(0, 1 << 40, nparts - 1, dtype=np.int64))
    want_dest = np.searchsorted(bnd, keys, side="right")
    dcol, dbnd = torch.from_numpy(keys).cuda(), torch.from_numpy(bnd).cuda()
    _, _, counts, dest_dev, offs_dev = _plan(lib, lib.DS_KEY_I64, dcol, 8, None, n, dbnd, nparts - 1, False)
    assert np.array_equal(counts, np.bincount(want_dest, minlength=nparts))
    src = torch.from_numpy(rows.reshape(-1).copy()).cuda()
    guard = 64
    bufs, bases = [], []
    for d in range(nparts):
        b = torch.full((int(counts[d]) * width + 2 * guard + offset,), 0xAB, dtype=torch.uint8, device="cuda")
        bufs.append(b)
        bases.append(b.data_ptr() + guard + offset)
    arr = (ctypes.c_void_p * nparts)(*bases)
    for max_ctas in (0, 3):
        for b in bufs:
            b.fill_(0xAB)
        lib.call("ds_split_rows", src.data_ptr(), width, n, dest_dev.data_ptr(), offs_dev.data_ptr(), nparts,
                 arr, 0, max_ctas, _stream())
        torch.cuda.synchronize()
        for d in range(nparts):
            got = bufs[d].cpu().numpy()
            body = got[guard + offset: guard + offset + int(counts[d]) * width].reshape(-1, width)
            assert np.array_equal(body, rows[want_dest == d]), (d, max_ctas)
            assert (got[: guard + offset] == 0xAB).all() and (got[guard + offset + int(counts[d]) * width:] == 0xAB).all()


def test_split_rows_unaligned_source_and_fsb_keys(lib):
    """a source column that starts at an odd byte (plain-load path) and 10-byte keys, both orders"""
    rng = np.random.default_rng(5)
    n, width = 20_000, 90
    keys = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    keys[:, :8] = 0  # heavy ties on the leading bytes: the low bytes decide
    bnd_rows = np.sort(rng.choice(65536, 7, replace=False))
    bnd = np.zeros((7, 10), dtype=np.uint8)
    bnd[:, 8] = bnd_rows >> 8
    bnd[:, 9] = bnd_rows & 255
    k16 = keys[:, 8].astype(np.int64) * 256 + keys[:, 9]
    rows = rng.integers(0, 256, (n, width), dtype=np.uint8)
    whole = torch.from_numpy(np.concatenate([[0], rows.reshape(-1)]).astype(np.uint8)).cuda()
    src = whole[1:]
    for desc in (False, True):
        b = bnd[::-1].copy() if desc else bnd
        want_dest = (7 - np.searchsorted(bnd_rows, k16, side="right")) if desc else np.searchsorted(bnd_rows, k16, side="right")
        dk, db = torch.from_numpy(keys.reshape(-1)).cuda(), torch.from_numpy(b.reshape(-1).copy()).cuda()
        dest, _, counts, dest_dev, offs_dev = _plan(lib, lib.DS_KEY_FSB, dk, 10, None, n, db, 7, desc)
        assert np.array_equal(dest.astype(np.int64), want_dest)
        outs = [torch.zeros(int(c) * width, dtype=torch.uint8, device="cuda") for c in counts]
        arr = (ctypes.c_void_p * 8)(*[o.data_ptr() for o in outs])
        lib.call("ds_split_rows", src.data_ptr(), width, n, dest_dev.data_ptr(), offs_dev.data_ptr(), 8, arr, 0, 0, _stream())
        torch.cuda.synchronize()
        for d in range(8):
            assert np.array_equal(outs[d].cpu().numpy().reshape(-1, width), rows[want_dest == d])


@pytest.mark.parametrize("desc", [False, True])
@pytest.mark.parametrize("nb", [1, 7, 8, 15, 31])
def test_split_plan_fsb10_full_tiles(lib, desc, nb):
    """10-byte keys on the register path of the plan kernel (full tiles of an aligned column, 8
    consecutive rows per lane) and on the general path (the ragged last tile, an unaligned
    column): keys equal to splitters, ties on the leading 8 bytes decided by the last two,
    nulls, every number of destinations' groups of four."""
    import bisect

    rng = np.random.default_rng(nb * 2 + desc)
    n = 256 * 37 + 101
    keys = rng.integers(0, 256, (n, 10), dtype=np.uint8)
    keys[rng.choice(n, n // 3, replace=False), :8] = 7       # ties on the leading 8 bytes
    keys[rng.choice(n, n // 5, replace=False), :9] = 7       # ... and on 9
    spl = keys[rng.choice(n, nb, replace=False)].copy()      # splitters are keys of the column
    if nb >= 7:
        spl[3] = spl[2]                                      # an empty destination
    order = sorted(range(nb), key=lambda i: bytes(spl[i]), reverse=desc)
    spl = spl[order]
    valid = (rng.random(n) > 0.05).astype(np.uint8)
    sb = [bytes(r) for r in spl]
    if desc:   # ArrowBlockAccessor.sort_and_partition on a descending sort: boundaries strictly greater
        asc = sb[::-1]
        want = np.array([nb - bisect.bisect_right(asc, bytes(k)) for k in keys], dtype=np.int64)
    else:
        want = np.array([bisect.bisect_right(sb, bytes(k)) for k in keys], dtype=np.int64)
    want = np.where(valid != 0, want, nb)
    dbnd = torch.from_numpy(spl.reshape(-1).copy()).cuda()
    dval = torch.from_numpy(valid).cuda()
    whole = torch.from_numpy(np.concatenate([np.zeros(16, np.uint8), keys.reshape(-1)])).cuda()
    for shift in (16, 14):                                   # aligned -> register path; 14 -> general path
        col = whole[shift:] if shift == 16 else torch.cat([whole[:14], whole[16:]])[14:]
        assert (col.data_ptr() % 16 == 0) == (shift == 16)
        dest, table, counts, _, _ = _plan(lib, lib.DS_KEY_FSB, col, 10, dval, n, dbnd, nb, desc)
        assert np.array_equal(dest.astype(np.int64), want), shift
        assert np.array_equal(counts, np.bincount(want, minlength=nb + 1))
        T = lib.DS_SPLIT_TILE_ROWS
        for d in range(nb + 1):
            per_tile = np.add.reduceat((want == d).astype(np.int64), np.arange(0, n, T))
            assert np.array_equal(table[d], np.concatenate([[0], np.cumsum(per_tile)])), (shift, d)


def test_split_rejects_bad_arguments(lib):
    z = torch.zeros(16, dtype=torch.uint8, device="cuda")
    arr = (ctypes.c_void_p * 1)(z.data_ptr())
    assert lib.lib.ds_split_rows(z.data_ptr(), lib.DS_SPLIT_MAX_ROW_BYTES + 1, 1, z.data_ptr(), z.data_ptr(), 1, arr, 0, 0, None) == lib.DS_ERR_UNSUPPORTED
    assert lib.lib.ds_split_rows(z.data_ptr(), 8, 1, z.data_ptr(), z.data_ptr(), lib.DS_MAX_SPLIT_PARTS + 1, arr, 0, 0, None) == lib.DS_ERR_UNSUPPORTED
    assert lib.lib.ds_split_plan(lib.DS_KEY_I64, z.data_ptr(), 8, None, 1, z.data_ptr(), lib.DS_MAX_SPLIT_PARTS, 0,
                                 z.data_ptr(), z.data_ptr(), z.data_ptr(), None) == lib.DS_ERR_UNSUPPORTED


@pytest.mark.parametrize("width", [90, 10, 7])
def test_split_rows_in_rounds(lib, width):
    """dest_mask: three calls that each move a third of the destinations together do what one
    call does; destinations outside a call's mask are not touched by it."""
    rng = np.random.default_rng(width)
    n, nparts = 60_001, 6
    rows = rng.integers(0, 256, (n, width), dtype=np.uint8)
    keys = rng.integers(0, 6000, n, dtype=np.int64)
    bnd = np.arange(1000, 6000, 1000, dtype=np.int64)
    want_dest = np.searchsorted(bnd, keys, side="right")
    dcol, dbnd = torch.from_numpy(keys).cuda(), torch.from_numpy(bnd).cuda()
    _, _, counts, dest_dev, offs_dev = _plan(lib, lib.DS_KEY_I64, dcol, 8, None, n, dbnd, nparts - 1, False)
    src = torch.from_numpy(rows.reshape(-1).copy()).cuda()
    outs = [torch.full((int(c) * width,), 0xCD, dtype=torch.uint8, device="cuda") for c in counts]
    arr = (ctypes.c_void_p * nparts)(*[o.data_ptr() for o in outs])
    for r in range(3):
        mask = (1 << r) | (1 << (r + 3))
        lib.call("ds_split_rows", src.data_ptr(), width, n, dest_dev.data_ptr(), offs_dev.data_ptr(), nparts,
                 arr, mask, 0, _stream())
        torch.cuda.synchronize()
        for d in range(nparts):
            got = outs[d].cpu().numpy().reshape(-1, width)
            if d % 3 <= r:
                assert np.array_equal(got, rows[want_dest == d]), (r, d)
            else:
                assert (got == 0xCD).all(), (r, d)
