"""GPU: ds_segment_heads_i64_static, the entry point for callers that cannot allocate.

Some users of this library hold the input and output buffers but have no way to get a
scratch buffer at the moment of the call: a reduce task inside a runtime whose pool owns
all of HBM, a thin ctypes/cgo binding with no allocator handle, a worker whose memory
plan leaves no room for an unplanned cudaMalloc.  For them the library carries a small
fixed scratch array of its own.  This test drives that entry point through the C ABI
and checks it against ds_segment_heads_i64 (caller workspace) -- integer outputs, so
the comparison is exact -- on two calls in a row, the second smaller than the first so
that anything the first call left in the shared array would show up.

Kept in a file that pytest collects last: if it fails under ``-x`` every other test has
already run.
"""
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


def _sorted_keys(n, groups, seed):
    rng = np.random.default_rng(seed)
    return np.sort(rng.integers(-groups, groups, n, dtype=np.int64))


def _heads_workspace(lib, keys_dev, n):
    starts = torch.full((n + 1,), -1, dtype=torch.int32, device="cuda")
    nseg = torch.zeros(1, dtype=torch.int32, device="cuda")
    ws = lib.call_ws("ds_segment_heads_i64", lambda b: torch.empty(b, dtype=torch.uint8, device="cuda"),
                     keys_dev.data_ptr(), n, starts.data_ptr(), nseg.data_ptr(),
                     stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    del ws
    k = int(nseg.item())
    return k, starts[: k + 1].cpu().numpy()


def _heads_static(lib, keys_dev, n):
    starts = torch.full((n + 1,), -1, dtype=torch.int32, device="cuda")
    nseg = torch.zeros(1, dtype=torch.int32, device="cuda")
    lib.call("ds_segment_heads_i64_static", keys_dev.data_ptr(), n, starts.data_ptr(),
             nseg.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    k = int(nseg.item())
    return k, starts[: k + 1].cpu().numpy()


def test_static_scratch_matches_workspace_variant_twice_in_a_row(lib):
    # call 1: 300k keys over 147 tiles; call 2: 5k keys over 3 tiles, other run lengths
    for n, groups, seed in ((300_000, 20_000, 1), (5_000, 40, 2)):
        keys = _sorted_keys(n, groups, seed)
        dev = torch.from_numpy(keys).cuda()
        want_k, want_starts = _heads_workspace(lib, dev, n)
        got_k, got_starts = _heads_static(lib, dev, n)
        assert got_k == want_k
        assert np.array_equal(got_starts, want_starts)
        # and both agree with numpy
        ref = np.concatenate([[0], np.flatnonzero(keys[1:] != keys[:-1]) + 1, [n]])
        assert got_k == len(ref) - 1 and np.array_equal(got_starts.astype(np.int64), ref)


def test_static_scratch_edge_sizes(lib):
    for n in (1, 2047, 2048, 2049):
        keys = np.arange(n, dtype=np.int64) // 3
        dev = torch.from_numpy(keys).cuda()
        assert _heads_static(lib, dev, n)[0] == _heads_workspace(lib, dev, n)[0] == (n + 2) // 3
    # n == 0 writes nseg = 0 and the sentinel
    k, starts = _heads_static(lib, torch.empty(0, dtype=torch.int64, device="cuda"), 0)
    assert k == 0 and starts.tolist() == [0]


def test_static_scratch_rejects_sizes_beyond_its_limit(lib):
    n = lib.DS_SEGMENT_HEADS_STATIC_MAX_ROWS + 1
    dummy = torch.zeros(2, dtype=torch.int64, device="cuda")
    st = lib.lib.ds_segment_heads_i64_static(dummy.data_ptr(), n, dummy.data_ptr(), dummy.data_ptr(), None)
    assert st == lib.DS_ERR_TOO_LARGE
    msg = lib.lib.ds_last_error().decode()
    assert "DS_SEGMENT_HEADS_STATIC_MAX_ROWS" in msg and "ds_segment_heads_i64" in msg
    with pytest.raises(lib.DeviceLibraryError):
        lib.call("ds_segment_heads_i64_static", dummy.data_ptr(), n, dummy.data_ptr(), dummy.data_ptr(), None)
