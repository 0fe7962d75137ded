"""GraySort-format record files <-> device blocks.

The reference states the format (python/ray/experimental/raysort/constants.py:8: 100-byte
records, 10-byte key first) and delegates generation / validation to the external gensort and
valsort binaries (raysort/main.py:148-176,399-432), which are not in its tree.  Here a file is
read into pinned host memory, copied to HBM and split into the two Arrow columns the sort path
works on (streamed in chunks through a ring of pinned buffers, see ``read_records``);
``validate`` is valsort's summary computed on the device.
"""
import collections
import os
from typing import Dict

import pyarrow as pa
import torch

from . import _lib
from .block import DeviceArrowBlock, DeviceColumn, _empty_u8, _stream

RECORD_BYTES = 100
KEY_BYTES = 10


def block_from_records(records: torch.Tensor, record_bytes: int = RECORD_BYTES,
                       key_bytes: int = KEY_BYTES) -> DeviceArrowBlock:
    """Interleaved records (uint8 tensor on the GPU) -> {key, payload} device block."""
    assert records.dtype == torch.uint8 and records.is_cuda and records.numel() % record_bytes == 0
    n = records.numel() // record_bytes
    key = _empty_u8(n * key_bytes)
    payload = _empty_u8(n * (record_bytes - key_bytes))
    _lib.call("ds_split_records", records.data_ptr(), record_bytes, key_bytes, n, key.data_ptr(),
              payload.data_ptr(), _stream())
    cols = collections.OrderedDict(
        key=DeviceColumn(pa.binary(key_bytes), key),
        payload=DeviceColumn(pa.binary(record_bytes - key_bytes), payload),
    )
    return DeviceArrowBlock(cols, n)


def records_from_block(block: DeviceArrowBlock) -> torch.Tensor:
    """{key, payload} device block -> interleaved records (uint8 tensor on the GPU)."""
    k, p = block.columns["key"], block.columns["payload"]
    rb = k.row_bytes + p.row_bytes
    out = _empty_u8(block.num_rows * rb)
    _lib.call("ds_join_records", k.ptr, p.ptr, rb, k.row_bytes, block.num_rows, out.data_ptr(), _stream())
    return out


CHUNK_BYTES = 64 << 20  # file chunk of the streamed reader / writer (rounded to whole records)
_RING = 3               # pinned staging buffers in flight


def read_records(path: str, record_bytes: int = RECORD_BYTES, key_bytes: int = KEY_BYTES,
                 chunk_bytes: int = CHUNK_BYTES) -> DeviceArrowBlock:
    """Streamed ingest: the file is read in chunks into a small ring of pinned buffers; each
    chunk's host-to-device copy and its split into the key / payload columns (ds_split_records)
    run on a side stream while the next chunk is being read from the file, so the disk read,
    the PCIe copy and the device work overlap and no file-sized host buffer exists
    (python/ray/experimental/raysort/main.py:148-176 streams its input parts likewise)."""
    size = os.path.getsize(path)
    if size % record_bytes:
        raise ValueError(f"{path}: size {size} is not a multiple of the record size {record_bytes}")
    n = size // record_bytes
    key = _empty_u8(n * key_bytes)
    payload = _empty_u8(n * (record_bytes - key_bytes))
    per = max(1, chunk_bytes // record_bytes)  # records per chunk
    if n:
        main = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(main)  # the columns are allocated on the caller's stream
        ring = [torch.empty(min(per, n) * record_bytes, dtype=torch.uint8, pin_memory=True)
                for _ in range(min(_RING, (n + per - 1) // per))]
        dev = [torch.empty_like(r, device="cuda") for r in ring]
        free = [None] * len(ring)  # event after which ring[i] / dev[i] may be reused
        with open(path, "rb", buffering=0) as f:
            for c, row0 in enumerate(range(0, n, per)):
                rows = min(per, n - row0)
                i = c % len(ring)
                if free[i] is not None:
                    free[i].synchronize()
                got = f.readinto(memoryview(ring[i].numpy())[: rows * record_bytes])
                if got != rows * record_bytes:
                    raise IOError(f"{path}: short read ({got} of {rows * record_bytes} bytes)")
                with torch.cuda.stream(side):
                    dev[i][: rows * record_bytes].copy_(ring[i][: rows * record_bytes], non_blocking=True)
                    _lib.call("ds_split_records", dev[i].data_ptr(), record_bytes, key_bytes, rows,
                              key.data_ptr() + row0 * key_bytes,
                              payload.data_ptr() + row0 * (record_bytes - key_bytes), side.cuda_stream)
                    free[i] = torch.cuda.Event()
                    free[i].record(side)
        main.wait_stream(side)
        for t in dev:
            t.record_stream(side)
    cols = collections.OrderedDict(
        key=DeviceColumn(pa.binary(key_bytes), key),
        payload=DeviceColumn(pa.binary(record_bytes - key_bytes), payload),
    )
    return DeviceArrowBlock(cols, n)


def write_records(block: DeviceArrowBlock, path: str, chunk_bytes: int = CHUNK_BYTES) -> None:
    """Streamed export: chunks of rows are joined into records on the device
    (ds_join_records), copied into a ring of pinned buffers on a side stream and written to the
    file while the next chunk is on its way."""
    k, p = block.columns["key"], block.columns["payload"]
    rb = k.row_bytes + p.row_bytes
    n = block.num_rows
    per = max(1, chunk_bytes // rb)
    with open(path, "wb", buffering=0) as f:
        if n == 0:
            return
        main = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(main)
        nring = min(_RING, (n + per - 1) // per)
        ring = [torch.empty(min(per, n) * rb, dtype=torch.uint8, pin_memory=True) for _ in range(nring)]
        dev = [torch.empty_like(r, device="cuda") for r in ring]
        ready = [None] * nring  # (event, bytes) of the chunk sitting in ring[i]

        def flush(i):
            ev, nbytes = ready[i]
            ev.synchronize()
            f.write(memoryview(ring[i].numpy())[:nbytes])
            ready[i] = None

        for c, row0 in enumerate(range(0, n, per)):
            rows = min(per, n - row0)
            i = c % nring
            if ready[i] is not None:
                flush(i)
            with torch.cuda.stream(side):
                _lib.call("ds_join_records", k.ptr + row0 * k.row_bytes, p.ptr + row0 * p.row_bytes, rb,
                          k.row_bytes, rows, dev[i].data_ptr(), side.cuda_stream)
                ring[i][: rows * rb].copy_(dev[i][: rows * rb], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(side)
            ready[i] = (ev, rows * rb)
        first = (n + per - 1) // per  # flush what is left, oldest first
        for c in range(max(0, first - nring), first):
            if ready[c % nring] is not None:
                flush(c % nring)
        main.wait_stream(side)
        for t in dev:
            t.record_stream(side)


def validate(block: DeviceArrowBlock) -> Dict[str, int]:
    """valsort's summary: records, unordered records, duplicate keys, checksum (sum of the
    CRC-32 of every record; equal before and after a correct sort)."""
    k, p = block.columns["key"], block.columns["payload"]
    out = torch.zeros(3, dtype=torch.int64, device="cuda")
    _lib.call("ds_validate_records", k.ptr, p.ptr, k.row_bytes + p.row_bytes, k.row_bytes,
              block.num_rows, out.data_ptr(), _stream())
    c, u, d = (int(x) & (2**64 - 1) for x in out.cpu().tolist())
    return {"records": block.num_rows, "unordered": u, "duplicate_keys": d, "checksum": c}
