"""GPU: GraySort record files <-> device blocks, valsort-style validation, repartition."""
import zlib

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


def _records(n, seed):
    t = O.graysort_table(n, seed)
    rec = np.concatenate([O.fsb_to_numpy(t["key"]), O.fsb_to_numpy(t["payload"])], axis=1)
    return t, np.ascontiguousarray(rec)


@pytest.mark.parametrize("n", [1, 255, 256, 257, 10_000])
def test_record_file_roundtrip_and_validate(ds, tmp_path, n):
    from ray_memory_management_b200 import graysort

    t, rec = _records(n, seed=n)
    path = tmp_path / "in.dat"
    rec.tofile(path)
    blk = graysort.read_records(str(path))
    assert blk.to_arrow().equals(t)
    out = tmp_path / "out.dat"
    graysort.write_records(blk, str(out))
    assert np.array_equal(np.fromfile(out, dtype=np.uint8), rec.reshape(-1))
    want_sum = sum(zlib.crc32(r.tobytes()) for r in rec)
    v = graysort.validate(blk)
    assert v["records"] == n and v["checksum"] == want_sum
    srt = ds.Dataset.from_blocks([blk]).sort("key").get_blocks()[0]
    vs = graysort.validate(srt)
    assert vs["checksum"] == want_sum and vs["unordered"] == 0  # same records, in order
    if n >= 10_000:
        assert v["unordered"] > n // 3  # the unsorted input is far from ordered


def test_record_file_streamed_in_chunks(ds, tmp_path):
    """many chunks through the pinned ring (chunk size far below the file size, ragged tail)"""
    from ray_memory_management_b200 import graysort

    n = 12_345
    t, rec = _records(n, seed=3)
    path = tmp_path / "in.dat"
    rec.tofile(path)
    blk = graysort.read_records(str(path), chunk_bytes=100 * 777)  # 16 chunks, the last one short
    assert blk.to_arrow().equals(t)
    out = tmp_path / "out.dat"
    graysort.write_records(blk, str(out), chunk_bytes=100 * 500)
    assert np.array_equal(np.fromfile(out, dtype=np.uint8), rec.reshape(-1))


def test_validate_counts_duplicates_and_disorder(ds):
    from ray_memory_management_b200 import graysort

    keys = np.zeros((6, 10), dtype=np.uint8)
    keys[:, 9] = [1, 1, 3, 2, 2, 2]
    t = pa.table({"key": O.fsb_from_numpy(keys), "payload": O.fsb_from_numpy(np.zeros((6, 90), np.uint8))})
    v = graysort.validate(ds.DeviceArrowBlock.from_arrow(t))
    assert v["duplicate_keys"] == 3 and v["unordered"] == 1


def test_record_arg_errors(ds):
    lib = ds._lib
    assert lib.lib.ds_split_records(None, 100, 100, 5, None, None, None) == lib.DS_ERR_INVALID_ARG
    assert lib.lib.ds_split_records(None, 500, 10, 5, None, None, None) == lib.DS_ERR_INVALID_ARG


@pytest.mark.parametrize("nb", [1, 3, 7])
def test_repartition_matches_oracle(ds, nb):
    tabs = [pa.table({"v": np.arange(a, b, dtype=np.int64)}) for a, b in ((0, 10), (10, 13), (13, 13), (13, 50))]
    out = ds.Dataset.from_arrow(tabs).repartition(nb)
    want = O.repartition_dataset(tabs, nb)
    assert out.num_blocks() == nb
    for g, w in zip(out.to_arrow(), want):
        assert (g["v"].to_pylist() if g.num_columns else []) == (w["v"].to_pylist() if w.num_columns else [])
