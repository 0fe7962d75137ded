"""GraySort-format record files <-> device blocks.

The reference states the format (python/ray/experimental/raysort/constants.py:8: 100-byte
records, 10-byte key first) and delegates generation / validation to the external gensort and
valsort binaries (raysort/main.py:148-176,399-432), which are not in its tree.  Here a file is
read into pinned host memory, copied to HBM and split into the two Arrow columns the sort path
works on; ``validate`` is valsort's summary computed on the device.
"""
import collections
import os
from typing import Dict

import numpy as np
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


def read_records(path: str, record_bytes: int = RECORD_BYTES, key_bytes: int = KEY_BYTES) -> DeviceArrowBlock:
    size = os.path.getsize(path)
    if size % record_bytes:
        raise ValueError(f"{path}: size {size} is not a multiple of the record size {record_bytes}")
    host = torch.empty(size, dtype=torch.uint8, pin_memory=True)
    with open(path, "rb") as f:
        f.readinto(host.numpy())
    return block_from_records(host.to("cuda", non_blocking=True), record_bytes, key_bytes)


def write_records(block: DeviceArrowBlock, path: str) -> None:
    rec = records_from_block(block)
    host = torch.empty(rec.numel(), dtype=torch.uint8, pin_memory=True)
    host.copy_(rec)
    torch.cuda.current_stream().synchronize()
    with open(path, "wb") as f:
        f.write(host.numpy().tobytes())


def validate(block: DeviceArrowBlock) -> Dict[str, int]:
    """valsort's summary: records, unordered records, duplicate keys, checksum (sum of the
    CRC-32 of every record; equal before and after a correct sort)."""
    k, p = block.columns["key"], block.columns["payload"]
    out = torch.zeros(3, dtype=torch.int64, device="cuda")
    _lib.call("ds_validate_records", k.ptr, p.ptr, k.row_bytes + p.row_bytes, k.row_bytes,
              block.num_rows, out.data_ptr(), _stream())
    c, u, d = (int(x) & (2**64 - 1) for x in out.cpu().tolist())
    return {"records": block.num_rows, "unordered": u, "duplicate_keys": d, "checksum": c}
