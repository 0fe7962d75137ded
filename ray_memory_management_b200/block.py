"""Device-resident Arrow-layout block and its accessor.

Mirrors, for one new block kind, the block surface the reference's shuffle ops call
(python/ray/data/block.py:100,148-230,233-422 and the ArrowBlockAccessor methods in
python/ray/data/_internal/arrow_block.py:178-186,226-232,258-272,353-394,396-587).

A ``DeviceArrowBlock`` keeps every column as ONE contiguous uint8 torch tensor in HBM
holding the Arrow data buffer of a fixed-width column (int64, float64,
fixed_size_binary(w), any other fixed-width primitive as payload).  Slices are views
(pointer arithmetic), never copies.  All arithmetic goes through the C ABI in
``_lib`` (CUDA, sm_100a); torch is only the owner of device memory and streams.
"""
import collections
import ctypes
import os
import random
import time
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence, Tuple, Union

import numpy as np
import pyarrow as pa
import torch

from . import _lib
from .aggregate import AggregateFn

# A sort key as the reference normalises it: [(column, "ascending"|"descending")]
SortKeyT = Union[None, List[Tuple[str, str]]]


# --------------------------------------------------------------------------- stats
class BlockExecStats:
    """python/ray/data/block.py:148-205.

    In the reference a task runs synchronously, so its wall time IS the time the work took.
    Here the map / reduce function of a block only enqueues kernels, so ``wall_time_s`` is the
    DEVICE time of those launches: a CUDA-event pair recorded on the launching stream by the
    builder, resolved (one event synchronise) the first time the value is read.
    ``host_time_s`` keeps the host time spent issuing the launches."""

    def __init__(self):
        self._wall: Optional[float] = None
        self._events = None
        self.host_time_s: Optional[float] = None
        self.cpu_time_s: Optional[float] = None
        self.node_id = f"cuda:{torch.cuda.current_device()}" if torch.cuda.is_available() else "host"
        self.max_rss_bytes: int = 0

    @property
    def wall_time_s(self) -> Optional[float]:
        if self._events is not None:
            a, b = self._events
            b.synchronize()
            self._wall = a.elapsed_time(b) / 1e3
            self._events = None
        return self._wall

    @wall_time_s.setter
    def wall_time_s(self, v: Optional[float]) -> None:
        self._events = None
        self._wall = v

    @staticmethod
    def builder() -> "_BlockExecStatsBuilder":
        return _BlockExecStatsBuilder()

    def __repr__(self):
        return repr(
            {"wall_time_s": self.wall_time_s, "cpu_time_s": self.cpu_time_s, "node_id": self.node_id}
        )


class _BlockExecStatsBuilder:
    def __init__(self):
        self.start_time = time.perf_counter()
        self.start_cpu = time.process_time()
        self.start_event = None
        if torch.cuda.is_available():
            self.start_event = torch.cuda.Event(enable_timing=True)
            self.start_event.record()

    def build(self) -> BlockExecStats:
        stats = BlockExecStats()
        stats.host_time_s = time.perf_counter() - self.start_time
        stats.wall_time_s = stats.host_time_s
        if self.start_event is not None:
            end = torch.cuda.Event(enable_timing=True)
            end.record()
            stats._events = (self.start_event, end)
        stats.cpu_time_s = time.process_time() - self.start_cpu
        try:
            import resource

            stats.max_rss_bytes = int(resource.getrusage(resource.RUSAGE_SELF).ru_maxrss * 1e3)
        except Exception:  # pragma: no cover
            stats.max_rss_bytes = 0
        return stats


@dataclass
class BlockMetadata:
    """python/ray/data/block.py:208-230."""

    num_rows: Optional[int]
    size_bytes: Optional[int]
    schema: Optional[Any]
    input_files: Optional[List[str]]
    exec_stats: Optional[BlockExecStats]

    def __post_init__(self):
        if self.input_files is None:
            self.input_files = []


# --------------------------------------------------------------------------- device helpers
def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _empty_u8(nbytes: int) -> torch.Tensor:
    return torch.empty(max(int(nbytes), 0), dtype=torch.uint8, device="cuda")


def _workspace(nbytes: int) -> torch.Tensor:
    # 256-byte alignment is what the library asks for; torch's allocator gives 512.
    return _empty_u8(nbytes)


def _row_bytes(t: pa.DataType) -> int:
    if pa.types.is_fixed_size_binary(t):
        return t.byte_width
    if pa.types.is_boolean(t):
        raise NotImplementedError("boolean (bit-packed) columns are not supported on device")
    try:
        bw = t.bit_width
    except ValueError:
        raise NotImplementedError(f"variable-width Arrow type {t} is not supported on device yet")
    if bw % 8:
        raise NotImplementedError(f"Arrow type {t} is not byte-aligned")
    return bw // 8


def _is_varlen(t: pa.DataType) -> bool:
    return (pa.types.is_string(t) or pa.types.is_binary(t) or pa.types.is_large_string(t)
            or pa.types.is_large_binary(t))


def _key_kind(t: pa.DataType) -> str:
    if pa.types.is_int64(t):
        return "i64"
    if pa.types.is_float64(t):
        return "f64"
    if pa.types.is_fixed_size_binary(t) and 1 <= t.byte_width <= _lib.DS_MAX_KEY_BYTES:
        return "fsb"
    if _is_varlen(t):
        return "var"
    raise NotImplementedError(
        f"sort/groupby key of Arrow type {t} is not supported on device "
        f"(int64, float64, fixed_size_binary(1..{_lib.DS_MAX_KEY_BYTES}), string, binary)"
    )


# Arrow types that are sorted through a widened 64-bit stand-in: (ds_widen_keys code, stand-in
# type).  64-bit temporal types are int64 underneath and are only reinterpreted.
def _stand_in_type(t: pa.DataType):
    T = pa.types
    if T.is_int8(t):
        return 0, pa.int64()
    if T.is_int16(t):
        return 1, pa.int64()
    if T.is_int32(t) or T.is_date32(t) or T.is_time32(t):
        return 2, pa.int64()
    if T.is_uint8(t):
        return 3, pa.int64()
    if T.is_uint16(t):
        return 4, pa.int64()
    if T.is_uint32(t):
        return 5, pa.int64()
    if T.is_float32(t):
        return 6, pa.float64()
    if T.is_float16(t):
        return 7, pa.float64()
    if T.is_uint64(t):
        return 8, pa.int64()
    if T.is_timestamp(t) or T.is_date64(t) or T.is_time64(t) or T.is_duration(t):
        return None, pa.int64()
    return None, None


def _stand_in_value(t: pa.DataType, v):
    """A splitter of the original key type as a value of the stand-in type."""
    if pa.types.is_uint64(t):
        return int(v) - (1 << 63)
    if isinstance(v, (np.datetime64, np.timedelta64)):
        return int(v.astype(np.int64))
    return v


class DeviceColumn:
    """One fixed-width Arrow column in HBM: ``data`` is a uint8 view of n*row_bytes bytes.

    ``valid`` (optional) is the column's validity as one byte per row (1 = value, 0 = null);
    it is unpacked from / packed into the Arrow validity bitmap at the H2D / D2H boundary so
    that it can be sliced and gathered like any 1-byte column.  ``None`` means "no nulls"."""

    __slots__ = ("type", "row_bytes", "data", "valid")
    is_varlen = False

    def __init__(self, arrow_type: pa.DataType, data: torch.Tensor, valid: Optional[torch.Tensor] = None):
        self.type = arrow_type
        self.row_bytes = _row_bytes(arrow_type)
        assert data.dtype == torch.uint8 and data.dim() == 1, (data.dtype, data.shape)
        assert data.numel() % self.row_bytes == 0
        self.data = data
        if valid is not None:
            assert valid.dtype == torch.uint8 and valid.numel() == data.numel() // self.row_bytes
        self.valid = valid

    @property
    def valid_ptr(self) -> Optional[int]:
        return self.valid.data_ptr() if self.valid is not None else None

    @property
    def num_rows(self) -> int:
        return self.data.numel() // self.row_bytes

    @property
    def ptr(self) -> int:
        return self.data.data_ptr()

    def slice(self, start: int, end: int) -> "DeviceColumn":
        return DeviceColumn(
            self.type,
            self.data[start * self.row_bytes : end * self.row_bytes],
            self.valid[start:end] if self.valid is not None else None,
        )

    def typed(self) -> torch.Tensor:
        """int64/float64 view of the buffer (plumbing for tests and finalisation)."""
        if pa.types.is_int64(self.type):
            return self.data.view(torch.int64)
        if pa.types.is_float64(self.type):
            return self.data.view(torch.float64)
        return self.data.view(-1, self.row_bytes)

    @property
    def nbytes(self) -> int:
        return self.data.numel()

    def with_valid(self, valid: Optional[torch.Tensor]) -> "DeviceColumn":
        return DeviceColumn(self.type, self.data, valid)

    def clone(self) -> "DeviceColumn":
        return DeviceColumn(self.type, self.data.clone(), self.valid.clone() if self.valid is not None else None)


class DeviceVarColumn(DeviceColumn):
    """A variable-length Arrow column (string / binary / large_*) in HBM: ``offsets`` are
    n+1 int64 byte positions into ``data`` (whatever the Arrow type's offset width; the width
    is restored at export).  ``offsets[0]`` need not be 0: a slice is a view of the parent's
    offsets and shares its bytes, as in Arrow."""

    __slots__ = ("offsets", "data_bytes")
    is_varlen = True

    def __init__(self, arrow_type: pa.DataType, offsets: torch.Tensor, data: torch.Tensor,
                 valid: Optional[torch.Tensor] = None, data_bytes: Optional[int] = None):
        assert _is_varlen(arrow_type), arrow_type
        assert offsets.dtype == torch.int64 and offsets.dim() == 1 and offsets.numel() >= 1
        assert data.dtype == torch.uint8 and data.dim() == 1
        self.type = arrow_type
        self.row_bytes = None  # arithmetic on it fails loudly where a path is fixed-width only
        self.offsets = offsets
        self.data = data
        if valid is not None:
            assert valid.dtype == torch.uint8 and valid.numel() == offsets.numel() - 1
        self.valid = valid
        # bytes the rows occupy: exact when known (import, take), else an estimate (slices)
        self.data_bytes = data.numel() if data_bytes is None else int(data_bytes)

    @property
    def num_rows(self) -> int:
        return self.offsets.numel() - 1

    def slice(self, start: int, end: int) -> "DeviceVarColumn":
        n = self.num_rows
        est = self.data_bytes * (end - start) // n if n else 0
        return DeviceVarColumn(self.type, self.offsets[start:end + 1], self.data,
                               self.valid[start:end] if self.valid is not None else None, est)

    def typed(self) -> torch.Tensor:
        raise TypeError(f"a {self.type} column has no fixed-width view")

    @property
    def nbytes(self) -> int:
        off_w = 8 if (pa.types.is_large_string(self.type) or pa.types.is_large_binary(self.type)) else 4
        return self.data_bytes + off_w * (self.num_rows + 1)

    def with_valid(self, valid: Optional[torch.Tensor]) -> "DeviceVarColumn":
        return DeviceVarColumn(self.type, self.offsets, self.data, valid, self.data_bytes)

    def clone(self) -> "DeviceVarColumn":
        return _gather_varlen(self, None)


class DeviceArrowBlock:
    """An Arrow-layout table whose column buffers live in HBM."""

    def __init__(self, columns: "collections.OrderedDict[str, DeviceColumn]", num_rows: int):
        self.columns = columns
        self._num_rows = int(num_rows)
        for name, c in columns.items():
            assert c.num_rows == self._num_rows, (name, c.num_rows, num_rows)

    # -- construction --------------------------------------------------------------
    @staticmethod
    def empty() -> "DeviceArrowBlock":
        return DeviceArrowBlock(collections.OrderedDict(), 0)

    @staticmethod
    def from_arrow(table: pa.Table) -> "DeviceArrowBlock":
        """H2D import of a pyarrow.Table (python/ray/data/_internal/arrow_block.py:135-138
        is where the reference materialises Arrow blocks)."""
        if not torch.cuda.is_available():
            raise RuntimeError("DeviceArrowBlock needs a CUDA device; there is no CPU fallback")
        cols = collections.OrderedDict()
        n = table.num_rows if table.num_columns > 0 else 0
        for name in table.column_names:
            ca = table[name]
            if pa.types.is_null(ca.type):
                # an all-null column as Arrow infers it from [None, ...]: kept as float64 nulls
                ca = pa.chunked_array([pa.array([None] * n, pa.float64())])
            arr = ca.combine_chunks() if ca.num_chunks != 1 else ca.chunk(0)
            if isinstance(arr, pa.ChunkedArray):
                arr = arr.chunk(0) if arr.num_chunks else pa.array([], type=ca.type)
            valid = None
            if arr.null_count and n:
                bits = np.frombuffer(arr.buffers()[0], dtype=np.uint8)
                mask = np.unpackbits(bits, bitorder="little")[arr.offset : arr.offset + n]
                valid = torch.from_numpy(np.ascontiguousarray(mask)).to("cuda")
            if _is_varlen(arr.type):
                cols[name] = _import_varlen(arr, n, valid)
                continue
            w = _row_bytes(arr.type)
            buf = arr.buffers()[1]
            if buf is None or n == 0:
                host = np.zeros(0, dtype=np.uint8)
            else:
                host = np.frombuffer(buf, dtype=np.uint8)[arr.offset * w : (arr.offset + n) * w]
            dev = torch.from_numpy(host.copy() if not host.flags.writeable else host).to("cuda", non_blocking=False)
            cols[name] = DeviceColumn(arr.type, dev, valid)
        return DeviceArrowBlock(cols, n)

    @staticmethod
    def from_host_buffers(
        schema: Sequence[Tuple[str, pa.DataType]], buffers: Sequence[torch.Tensor], num_rows: int
    ) -> "DeviceArrowBlock":
        """H2D import of raw (ideally pinned) uint8 host buffers, one per column; the
        copies are asynchronous on the current stream."""
        cols = collections.OrderedDict()
        for (name, t), host in zip(schema, buffers):
            assert host.dtype == torch.uint8
            dev = torch.empty(host.numel(), dtype=torch.uint8, device="cuda")
            dev.copy_(host, non_blocking=True)
            cols[name] = DeviceColumn(t, dev)
        return DeviceArrowBlock(cols, num_rows)

    def to_host_buffers(self, out: Sequence[torch.Tensor]) -> None:
        """D2H export of the data buffers into caller-provided (ideally pinned) uint8 buffers
        (columns with nulls need ``to_arrow``, which also exports validity)."""
        for name, c in self.columns.items():
            if c.valid is not None and not bool(c.valid.all()):
                raise NotImplementedError(f"to_host_buffers: column {name!r} has nulls; use to_arrow()")
        for c, host in zip(self.columns.values(), out):
            if c.is_varlen:
                raise NotImplementedError("to_host_buffers: variable-length columns need to_arrow()")
            if host.numel() < c.data.numel():
                raise ValueError(f"host buffer too small: {host.numel()} < {c.data.numel()} bytes")
            host[: c.data.numel()].copy_(c.data, non_blocking=True)

    def to_arrow(self) -> pa.Table:
        """D2H export (synchronises)."""
        if not self.columns:
            return pa.Table.from_pydict({})
        arrays, fields = [], []
        for name, c in self.columns.items():
            vbuf, nulls = None, 0
            if c.valid is not None:
                mask = c.valid.cpu().numpy()
                nulls = int(self._num_rows - np.count_nonzero(mask))
                if nulls:
                    vbuf = pa.py_buffer(np.packbits(mask != 0, bitorder="little").tobytes())
            if c.is_varlen:
                bufs = [vbuf] + _export_varlen(c)
            else:
                bufs = [vbuf, pa.py_buffer(c.data.cpu().numpy().tobytes())]
            arrays.append(pa.Array.from_buffers(c.type, self._num_rows, bufs, null_count=nulls))
            fields.append(pa.field(name, c.type))
        return pa.Table.from_arrays(arrays, schema=pa.schema(fields))

    # -- shape ---------------------------------------------------------------------
    @property
    def num_rows(self) -> int:
        return self._num_rows if self.columns else 0

    @property
    def num_columns(self) -> int:
        return len(self.columns)

    @property
    def column_names(self) -> List[str]:
        return list(self.columns)

    @property
    def nbytes(self) -> int:
        # like pyarrow.Table.nbytes: data buffers plus validity bitmaps where present
        return sum(
            c.nbytes + ((self._num_rows + 7) // 8 if c.valid is not None else 0)
            for c in self.columns.values()
        )

    @property
    def schema(self) -> pa.Schema:
        return pa.schema([pa.field(n, c.type) for n, c in self.columns.items()])

    def slice(self, start: int, end: int) -> "DeviceArrowBlock":
        start = max(0, min(start, self.num_rows))
        end = max(start, min(end, self.num_rows))
        cols = collections.OrderedDict((n, c.slice(start, end)) for n, c in self.columns.items())
        return DeviceArrowBlock(cols, end - start)

    def select(self, names: Sequence[str]) -> "DeviceArrowBlock":
        return DeviceArrowBlock(
            collections.OrderedDict((n, self.columns[n]) for n in names), self._num_rows
        )

    def column(self, name: str) -> DeviceColumn:
        if name not in self.columns:
            raise ValueError(
                f"Cannot find column {name}, available columns: {self.column_names}"
            )
        return self.columns[name]

    @staticmethod
    def concat(blocks: Sequence["DeviceArrowBlock"]) -> "DeviceArrowBlock":
        """Row-wise concatenation in the given order (pyarrow.concat_tables at
        _internal/arrow_ops/transform_pyarrow.py:73)."""
        blocks = [b for b in blocks if b.num_columns > 0]
        if not blocks:
            return DeviceArrowBlock.empty()
        if len(blocks) == 1:
            return blocks[0]
        names = blocks[0].column_names
        for b in blocks[1:]:
            if b.column_names != names or any(
                b.columns[n].type != blocks[0].columns[n].type for n in names
            ):
                raise ValueError("cannot concatenate device blocks with different schemas")
        cols = collections.OrderedDict()
        for n in names:
            valid = None
            if any(b.columns[n].valid is not None for b in blocks):
                valid = torch.cat([
                    b.columns[n].valid if b.columns[n].valid is not None
                    else torch.ones(b.num_rows, dtype=torch.uint8, device="cuda")
                    for b in blocks
                ])
            if blocks[0].columns[n].is_varlen:
                cols[n] = _concat_varlen([b.columns[n] for b in blocks], valid)
                continue
            cols[n] = DeviceColumn(
                blocks[0].columns[n].type, torch.cat([b.columns[n].data for b in blocks]), valid
            )
        return DeviceArrowBlock(cols, sum(b.num_rows for b in blocks))


# --------------------------------------------------------------------------- variable-length columns
def _import_varlen(arr: pa.Array, n: int, valid: Optional[torch.Tensor]) -> DeviceVarColumn:
    large = pa.types.is_large_string(arr.type) or pa.types.is_large_binary(arr.type)
    bufs = arr.buffers()
    if n == 0 or bufs[1] is None:
        return DeviceVarColumn(arr.type, torch.zeros(1, dtype=torch.int64, device="cuda"), _empty_u8(0), valid, 0)
    offs = np.frombuffer(bufs[1], dtype=np.int64 if large else np.int32)[arr.offset : arr.offset + n + 1]
    first, last = int(offs[0]), int(offs[-1])
    offs = offs.astype(np.int64) - first
    if bufs[2] is None or last == first:
        data = np.zeros(0, dtype=np.uint8)
    else:
        data = np.frombuffer(bufs[2], dtype=np.uint8)[first:last]
    return DeviceVarColumn(
        arr.type, torch.from_numpy(offs).to("cuda"),
        torch.from_numpy(data.copy() if not data.flags.writeable else data).to("cuda"), valid, last - first)


def _export_varlen(c: DeviceVarColumn) -> list:
    """[offsets buffer, data buffer] for pyarrow.Array.from_buffers; offsets are rebased to 0
    and narrowed to the Arrow type's width."""
    large = pa.types.is_large_string(c.type) or pa.types.is_large_binary(c.type)
    offs = c.offsets.cpu().numpy()
    first, last = int(offs[0]), int(offs[-1])
    offs = offs - first
    if not large:
        if last - first > np.iinfo(np.int32).max:
            raise ValueError(f"{c.type} column holds {last - first} bytes, more than 32-bit offsets can address")
        offs = offs.astype(np.int32)
    data = c.data[first:last].cpu().numpy()
    return [pa.py_buffer(offs.tobytes()), pa.py_buffer(data.tobytes())]


def _concat_varlen(cols: Sequence[DeviceVarColumn], valid: Optional[torch.Tensor]) -> DeviceVarColumn:
    # one device-to-host copy for every block's first and last offset
    ends = torch.stack([torch.stack([c.offsets[0], c.offsets[-1]]) for c in cols]).cpu().tolist()
    offs, data, base = [], [], 0
    for c, (first, last) in zip(cols, ends):
        offs.append(c.offsets[:-1] + (base - first))
        data.append(c.data[first:last])
        base += last - first
    offs.append(torch.tensor([base], dtype=torch.int64, device="cuda"))
    return DeviceVarColumn(cols[0].type, torch.cat(offs), torch.cat(data), valid, base)


def _gather_varlen(col: DeviceVarColumn, idx: Optional[torch.Tensor]) -> DeviceVarColumn:
    """ds_varlen_take_offsets + ds_varlen_take_bytes; ``idx`` None compacts rows 0..n-1."""
    n = col.num_rows if idx is None else idx.numel()
    iptr = idx.data_ptr() if idx is not None else None
    dst_off = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    ws = _lib.call_ws("ds_varlen_take_offsets", _workspace, col.offsets.data_ptr(), col.num_rows, iptr, n,
                      dst_off.data_ptr(), stream=_stream())
    ws.record_stream(torch.cuda.current_stream())
    total = int(dst_off[-1].item())  # host sync: the output byte count sizes the allocation
    out = _empty_u8(total)
    _lib.call("ds_varlen_take_bytes", col.offsets.data_ptr(), col.data.data_ptr(), iptr, n,
              dst_off.data_ptr(), out.data_ptr(), _stream())
    valid = None
    if col.valid is not None:
        valid = col.valid.clone() if idx is None else _gather_raw(col.valid, 1, col.num_rows, idx)
    return DeviceVarColumn(col.type, dst_off, out, valid, total)


def _sort_u64(keys: torch.Tensor, descending: bool) -> torch.Tensor:
    n = keys.numel()
    idx = torch.empty(n, dtype=torch.int32, device="cuda")
    ws = _lib.call_ws("ds_sort_indices_u64", _workspace, keys.data_ptr(), n, int(descending), idx.data_ptr(),
                      None, 0, stream=_stream())
    ws.record_stream(torch.cuda.current_stream())
    return idx


def _argsort_varlen(col: DeviceVarColumn, descending: bool, want_sorted: bool):
    """Stable sort of a variable-length key: LSD over 64-bit parts (length, then 8-byte
    chunks from the last to the first), each part one stable radix sort; see csrc/varlen.cu."""
    n = col.num_rows
    stream = _stream()
    keys = torch.empty(n, dtype=torch.int64, device="cuda")
    max_len = torch.empty(1, dtype=torch.int64, device="cuda")
    _lib.call("ds_varlen_key_lengths", col.offsets.data_ptr(), None, n, keys.data_ptr(), max_len.data_ptr(), stream)
    longest = int(max_len.item())  # host sync: number of chunk rounds
    perm = _sort_u64(keys, descending)
    for chunk in range((longest + 7) // 8 - 1, -1, -1):
        _lib.call("ds_varlen_key_chunk", col.offsets.data_ptr(), col.data.data_ptr(), perm.data_ptr(), n,
                  8 * chunk, keys.data_ptr(), stream)
        order = _sort_u64(keys, descending)
        perm = _gather_raw(perm.view(torch.uint8), 4, n, order).view(torch.int32)
    return perm, (_gather_varlen(DeviceVarColumn(col.type, col.offsets, col.data), perm) if want_sorted else None)


def _varlen_bounds_to_device(boundaries: Sequence):
    raw = [b.encode("utf-8") if isinstance(b, str) else bytes(b) for b in boundaries]
    offs = np.zeros(len(raw) + 1, dtype=np.int64)
    np.cumsum([len(r) for r in raw], out=offs[1:])
    data = np.frombuffer(b"".join(raw), dtype=np.uint8)
    return torch.from_numpy(offs).to("cuda"), torch.from_numpy(data.copy()).to("cuda")


Block = Union[DeviceArrowBlock, pa.Table]


# --------------------------------------------------------------------------- kernels (thin wrappers)
def _sortable(col: DeviceColumn) -> DeviceColumn:
    """``col`` itself if the kernels take its type as a key, else a 64-bit column that orders
    and ties like it (ds_widen_keys; a reinterpretation for 64-bit temporal types)."""
    if col.is_varlen:
        return col
    code, stand_in = _stand_in_type(col.type)
    if stand_in is None:
        return col
    if code is None:
        return DeviceColumn(stand_in, col.data, col.valid)
    n = col.num_rows
    out = _empty_u8(n * 8)
    _lib.call("ds_widen_keys", col.ptr, code, n, out.data_ptr(), _stream())
    return DeviceColumn(stand_in, out, col.valid)


def _argsort(col: DeviceColumn, descending: bool, want_sorted: bool, options: int = 0):
    """ds_sort_indices_* -> (idx int32 tensor holding uint32 values, sorted key column or None)."""
    wide = _sortable(col)
    if wide is not col:
        idx, _ = _argsort(wide, descending, False, options)
        return idx, (_gather(col.with_valid(None), idx) if want_sorted else None)
    n = col.num_rows
    kind = _key_kind(col.type)
    if kind == "var":
        return _argsort_varlen(col, descending, want_sorted)
    idx = torch.empty(n, dtype=torch.int32, device="cuda")
    out = _empty_u8(n * col.row_bytes) if want_sorted else None
    optr = out.data_ptr() if out is not None else None
    stream = _stream()
    if kind == "fsb":
        args = (col.ptr, col.row_bytes, n, int(descending), idx.data_ptr(), optr, options)
    else:
        args = (col.ptr, n, int(descending), idx.data_ptr(), optr, options)
    ws = _lib.call_ws(f"ds_sort_indices_{kind}", _workspace, *args, stream=stream)
    ws.record_stream(torch.cuda.current_stream())
    return idx, (DeviceColumn(col.type, out) if want_sorted else None)


def _gather_raw(src: torch.Tensor, row_bytes: int, src_rows: int, idx: torch.Tensor) -> torch.Tensor:
    n = idx.numel()
    out = _empty_u8(n * row_bytes)
    _lib.call("ds_gather_rows", src.data_ptr(), row_bytes, src_rows, idx.data_ptr(), n, out.data_ptr(), _stream())
    return out


def _gather(col: DeviceColumn, idx: torch.Tensor) -> DeviceColumn:
    if col.is_varlen:
        return _gather_varlen(col, idx)
    out = _gather_raw(col.data, col.row_bytes, col.num_rows, idx)
    valid = _gather_raw(col.valid, 1, col.num_rows, idx) if col.valid is not None else None
    return DeviceColumn(col.type, out, valid)


def _multi_gather_ok(cols: Sequence[DeviceColumn]) -> bool:
    """ds_gather_rows_multi takes even rows of 18..1024 bytes that are not a multiple of 16 from
    at most 64 2-byte-aligned sources (the asynchronous-copy kernel)."""
    w = cols[0].row_bytes if not cols[0].is_varlen else 0
    return (1 < len(cols) <= 64 and w >= 18 and w <= 1024 and w % 2 == 0 and w % 16 != 0
            and all((not c.is_varlen) and c.type == cols[0].type and c.ptr % 2 == 0 for c in cols))


def _gather_multi(cols: Sequence[DeviceColumn], idx: torch.Tensor) -> DeviceColumn:
    """Row gather from a column that lies in several blocks (idx counts rows over their
    concatenation) without concatenating them first."""
    import ctypes

    n = idx.numel()
    rb = cols[0].row_bytes
    live = [c for c in cols if c.num_rows]
    out = _empty_u8(n * rb)
    srcs = (ctypes.c_void_p * len(live))(*[c.ptr for c in live])
    rows = (ctypes.c_uint64 * len(live))(*[c.num_rows for c in live])
    _lib.call("ds_gather_rows_multi", srcs, rows, len(live), rb, idx.data_ptr(), n, out.data_ptr(), _stream())
    valid = None
    if any(c.valid is not None for c in cols):
        v = torch.cat([c.valid if c.valid is not None else torch.ones(c.num_rows, dtype=torch.uint8, device="cuda")
                       for c in cols])
        valid = _gather_raw(v, 1, v.numel(), idx)
    return DeviceColumn(cols[0].type, out, valid)


def _split_nulls(col: DeviceColumn) -> Tuple[torch.Tensor, int]:
    """Row ids with the non-null rows first and the null rows last, input order kept in
    both groups (a stable two-way partition on the validity byte), and the non-null count."""
    vcol = DeviceColumn(pa.binary(1), col.valid)
    perm, counts = _range_partition(vcol, [b"\x01"], True)  # descending: valid(1) -> bucket 0
    return perm, counts[0]


def _argsort_nulls_last(col: DeviceColumn, descending: bool):
    """Arrow's sort_indices order: by key, nulls at the end in input order (both directions).
    Returns (idx, sorted key column, number of non-null rows)."""
    n = col.num_rows
    if col.valid is None:
        idx, sk = _argsort(col, descending, want_sorted=True)
        return idx, sk, n
    perm0, nv = _split_nulls(col)
    if nv == n:
        idx, sk = _argsort(col.with_valid(None), descending, want_sorted=True)
        return idx, sk, n
    bare = col.with_valid(None)
    idx = torch.empty(n, dtype=torch.int32, device="cuda")
    pieces = []
    if nv:
        idx2, sk = _argsort(_gather(bare, perm0[:nv]), descending, want_sorted=True)
        idx[:nv] = _gather_raw(perm0.view(torch.uint8), 4, n, idx2).view(torch.int32)
        pieces.append(sk)
    idx[nv:] = perm0[nv:]
    pieces.append(_gather(bare, perm0[nv:]))
    valid = torch.zeros(n, dtype=torch.uint8, device="cuda")
    valid[:nv] = 1
    if col.is_varlen:
        return idx, _concat_varlen(pieces, valid), nv
    return idx, DeviceColumn(col.type, torch.cat([p.data for p in pieces]), valid), nv


def _take_block(block: DeviceArrowBlock, idx: torch.Tensor, ready: Optional[Dict[str, DeviceColumn]] = None) -> DeviceArrowBlock:
    cols = collections.OrderedDict()
    for name, c in block.columns.items():
        cols[name] = ready[name] if ready and name in ready else _gather(c, idx)
    return DeviceArrowBlock(cols, idx.numel())


def _boundaries_to_device(boundaries: Sequence, col: DeviceColumn) -> torch.Tensor:
    kind = _key_kind(col.type)
    if kind == "i64":
        host = np.asarray([int(b) for b in boundaries], dtype=np.int64).view(np.uint8)
    elif kind == "f64":
        host = np.asarray([float(b) for b in boundaries], dtype=np.float64).view(np.uint8)
    else:
        raw = b"".join(bytes(b) for b in boundaries)
        if len(raw) != len(boundaries) * col.row_bytes:
            raise ValueError("fixed_size_binary boundaries must have the key's width")
        host = np.frombuffer(raw, dtype=np.uint8)
    return torch.from_numpy(host.copy()).to("cuda")


def _searchsorted(sorted_col: DeviceColumn, boundaries: Sequence, descending: bool) -> List[int]:
    wide = _sortable(sorted_col)
    if wide is not sorted_col:
        boundaries = [_stand_in_value(sorted_col.type, b) for b in boundaries]
        sorted_col = wide
    nb = len(boundaries)
    kind = _key_kind(sorted_col.type)
    out = torch.empty(nb, dtype=torch.int64, device="cuda")
    if kind == "var":
        boff, bdata = _varlen_bounds_to_device(boundaries)
        _lib.call("ds_searchsorted_varlen", sorted_col.offsets.data_ptr(), sorted_col.data.data_ptr(),
                  sorted_col.num_rows, boff.data_ptr(), bdata.data_ptr(), nb, int(descending),
                  out.data_ptr(), _stream())
        return [int(x) for x in out.cpu().tolist()]
    b = _boundaries_to_device(boundaries, sorted_col)
    if kind == "fsb":
        _lib.call("ds_searchsorted_fsb", sorted_col.ptr, sorted_col.row_bytes, sorted_col.num_rows,
                  b.data_ptr(), nb, int(descending), out.data_ptr(), _stream())
    else:
        _lib.call(f"ds_searchsorted_{kind}", sorted_col.ptr, sorted_col.num_rows, b.data_ptr(), nb,
                  int(descending), out.data_ptr(), _stream())
    return [int(x) for x in out.cpu().tolist()]


def _range_partition(col: DeviceColumn, boundaries: Sequence, descending: bool):
    """ds_partition_hist_scatter_* -> (perm tensor, counts list)."""
    wide = _sortable(col)
    if wide is not col:
        boundaries = [_stand_in_value(col.type, b) for b in boundaries]
        col = wide
    kind = _key_kind(col.type)
    if kind in ("f64", "var"):
        # no one-pass bucket kernel for these key types: a stable sort and a searchsorted give
        # a partition that is also ordered inside each part (the reducer's stable sort makes
        # the final result the same either way)
        idx, sorted_key = _argsort(col, descending, want_sorted=True)
        edges = [0] + _searchsorted(sorted_key, boundaries, descending) + [col.num_rows]
        for j in range(1, len(edges)):
            edges[j] = max(edges[j], edges[j - 1])
        return idx, [edges[i + 1] - edges[i] for i in range(len(boundaries) + 1)]
    n, nb = col.num_rows, len(boundaries)
    b = _boundaries_to_device(boundaries, col)
    counts = torch.empty(nb + 1, dtype=torch.int64, device="cuda")
    offsets = torch.empty(nb + 1, dtype=torch.int64, device="cuda")
    perm = torch.empty(n, dtype=torch.int32, device="cuda")
    if kind == "fsb":
        args = (col.ptr, col.row_bytes, n, b.data_ptr(), nb, int(descending))
    else:
        args = (col.ptr, n, b.data_ptr(), nb, int(descending))
    ws = _lib.call_ws(f"ds_partition_hist_scatter_{kind}", _workspace, *args, counts.data_ptr(),
                      offsets.data_ptr(), perm.data_ptr(), stream=_stream())
    ws.record_stream(torch.cuda.current_stream())
    return perm, [int(x) for x in counts.cpu().tolist()]


def _random_permutation(n: int, seed: int, subsequence: int = 0) -> torch.Tensor:
    perm = torch.empty(n, dtype=torch.int32, device="cuda")
    ws = _lib.call_ws("ds_random_permutation", _workspace, seed & (2**64 - 1), subsequence, n,
                      perm.data_ptr(), stream=_stream())
    ws.record_stream(torch.cuda.current_stream())
    return perm


def _group_ids(sorted_keys: DeviceColumn) -> DeviceColumn:
    """int64 column with the same runs of equal values as the sorted key column."""
    if pa.types.is_float32(sorted_keys.type) or pa.types.is_float16(sorted_keys.type):
        sorted_keys = _sortable(sorted_keys)  # numeric comparison (-0.0 == 0.0) needs doubles
    n = sorted_keys.num_rows
    gids = torch.empty(n, dtype=torch.int64, device="cuda")
    if sorted_keys.is_varlen:
        ws = _lib.call_ws("ds_group_ids_varlen", _workspace, sorted_keys.offsets.data_ptr(),
                          sorted_keys.data.data_ptr(), n, gids.data_ptr(), stream=_stream())
    else:
        ws = _lib.call_ws("ds_group_ids_fixed", _workspace, sorted_keys.ptr, sorted_keys.row_bytes,
                          int(pa.types.is_float64(sorted_keys.type)), n, gids.data_ptr(), stream=_stream())
    ws.record_stream(torch.cuda.current_stream())
    return DeviceColumn(pa.int64(), gids.view(torch.uint8))


def _segment_heads(sorted_keys: DeviceColumn):
    if not pa.types.is_int64(sorted_keys.type):
        sorted_keys = _group_ids(sorted_keys)
    n = sorted_keys.num_rows
    starts = torch.empty(n + 1, dtype=torch.int32, device="cuda")
    nseg = torch.empty(1, dtype=torch.int32, device="cuda")
    ws = _lib.call_ws("ds_segment_heads_i64", _workspace, sorted_keys.ptr, n, starts.data_ptr(),
                      nseg.data_ptr(), stream=_stream())
    ws.record_stream(torch.cuda.current_stream())
    return starts, int(nseg.item())


_OPS = {"sum": _lib.DS_OP_SUM, "min": _lib.DS_OP_MIN, "max": _lib.DS_OP_MAX, "absmax": _lib.DS_OP_ABSMAX}


def _numeric64(vals: DeviceColumn) -> DeviceColumn:
    """Value column as int64 / float64 for the reductions: narrower integers and floats are
    widened (pyarrow's sum/min/max of an int32 column land in int64 rows too)."""
    t = vals.type
    if vals.is_varlen or pa.types.is_int64(t) or pa.types.is_float64(t) or pa.types.is_uint64(t) or pa.types.is_temporal(t):
        return vals
    code, stand_in = _stand_in_type(t)
    return _sortable(vals) if code is not None else vals


def _segmented_reduce(vals: DeviceColumn, idx: Optional[torch.Tensor], starts: torch.Tensor,
                      nseg: int, n: int, op: str) -> DeviceColumn:
    vals = _numeric64(vals)
    if pa.types.is_float64(vals.type):
        fn = "ds_segmented_reduce_f64"
    elif pa.types.is_int64(vals.type):
        fn = "ds_segmented_reduce_i64"
    else:
        raise NotImplementedError(f"aggregation over Arrow type {vals.type} is not supported on device (int64, float64)")
    out = _empty_u8(nseg * 8)
    ws = _lib.call_ws(fn, _workspace, vals.ptr, vals.valid_ptr, idx.data_ptr() if idx is not None else None,
                      starts.data_ptr(), nseg, n, _OPS[op], out.data_ptr(), stream=_stream())
    ws.record_stream(torch.cuda.current_stream())
    return DeviceColumn(vals.type, out)


def _segmented_count_valid(vals: DeviceColumn, idx: Optional[torch.Tensor], starts: torch.Tensor,
                           nseg: int, n: int) -> DeviceColumn:
    out = _empty_u8(nseg * 8)
    ws = _lib.call_ws("ds_segmented_count_valid", _workspace, vals.valid_ptr,
                      idx.data_ptr() if idx is not None else None, starts.data_ptr(), nseg, n,
                      out.data_ptr(), stream=_stream())
    ws.record_stream(torch.cuda.current_stream())
    return DeviceColumn(pa.int64(), out)


def _segmented_sumsqdiff(vals: DeviceColumn, idx: Optional[torch.Tensor], starts: torch.Tensor, nseg: int,
                         n: int, means: DeviceColumn, weights: Optional[DeviceColumn] = None,
                         addend: Optional[DeviceColumn] = None) -> DeviceColumn:
    vals = _numeric64(vals)
    out = _empty_u8(nseg * 8)
    ip = idx.data_ptr() if idx is not None else None
    if pa.types.is_float64(vals.type):
        args = (vals.ptr, vals.valid_ptr, weights.ptr if weights is not None else None,
                addend.ptr if addend is not None else None, ip, starts.data_ptr(), nseg, n, means.ptr,
                out.data_ptr())
        fn = "ds_segmented_sumsqdiff_f64"
    elif pa.types.is_int64(vals.type) and weights is None and addend is None:
        args = (vals.ptr, vals.valid_ptr, ip, starts.data_ptr(), nseg, n, means.ptr, out.data_ptr())
        fn = "ds_segmented_sumsqdiff_i64"
    else:
        raise NotImplementedError(f"std over Arrow type {vals.type} is not supported on device (int64, float64)")
    ws = _lib.call_ws(fn, _workspace, *args, stream=_stream())
    ws.record_stream(torch.cuda.current_stream())
    return DeviceColumn(pa.float64(), out)


def _scale(a: DeviceColumn, w: DeviceColumn) -> DeviceColumn:
    n = a.num_rows
    out = _empty_u8(n * 8)
    _lib.call("ds_scale_f64_by_i64", a.ptr, w.ptr, n, out.data_ptr(), _stream())
    return DeviceColumn(pa.float64(), out)


def _finalize_std(m2: DeviceColumn, count: DeviceColumn, ddof: int) -> DeviceColumn:
    n = count.num_rows
    out = _empty_u8(n * 8)
    _lib.call("ds_finalize_std", m2.ptr, count.ptr, n, int(ddof), out.data_ptr(), _stream())
    return DeviceColumn(pa.float64(), out)


def _finalize_validity(nvalid: DeviceColumn, nrows: DeviceColumn, ignore_nulls: bool) -> torch.Tensor:
    n = nvalid.num_rows
    out = _empty_u8(n)
    _lib.call("ds_finalize_validity", nvalid.ptr, nrows.ptr, n, int(ignore_nulls), out.data_ptr(), _stream())
    return out


def _segment_counts(starts: torch.Tensor, nseg: int) -> DeviceColumn:
    out = _empty_u8(nseg * 8)
    _lib.call("ds_segment_counts", starts.data_ptr(), nseg, out.data_ptr(), _stream())
    return DeviceColumn(pa.int64(), out)


def _finalize_mean(s: DeviceColumn, c: DeviceColumn) -> DeviceColumn:
    n = c.num_rows
    out = _empty_u8(n * 8)
    _lib.call("ds_finalize_mean", s.ptr, int(pa.types.is_int64(s.type)), c.ptr, n, out.data_ptr(), _stream())
    return DeviceColumn(pa.float64(), out)


# --------------------------------------------------------------------------- accessor
def _normalize_key(key) -> Tuple[str, bool]:
    """[(col, order)] -> (col, descending); arrow_block.py:356-359 for the error."""
    if isinstance(key, str):
        return key, False
    if key is None or callable(key):
        raise NotImplementedError(f"Table sort key must be a column name, was: {key}")
    if len(key) > 1:
        raise NotImplementedError("sorting by multiple columns is not supported yet")
    col, order = key[0]
    return col, order == "descending"


class PartialAggBlock(DeviceArrowBlock):
    """A combined (partially aggregated) block: key column + SoA accumulator columns.

    The reference stores each accumulator as an Arrow list column with a trailing
    has_data flag (python/ray/data/_internal/null_aggregate.py:29-46); on device the
    numeric parts are kept as plain columns named ``<agg>#sum`` / ``<agg>#nvalid`` / ...
    (``to_flat_arrow``), and the list wrapper is built only at the export boundary: ``to_arrow``
    of a block that knows its aggregations returns the reference's table."""

    def __init__(self, columns, num_rows, key: Optional[str] = None, aggs=None):
        super().__init__(columns, num_rows)
        self.key = key
        self.aggs = aggs

    def to_flat_arrow(self) -> pa.Table:
        return DeviceArrowBlock.to_arrow(self)

    def to_arrow(self) -> pa.Table:
        """The block as ArrowBlockAccessor.combine / aggregate_combined_blocks(finalize=False)
        return it (arrow_block.py:396-467, 488-587): the key column, then per aggregation a
        ``list`` column ``[accumulator parts..., has_data]`` -- ``count()`` stays int64 -- where
        a group without any valid value holds the initial accumulator with has_data 0
        (ignore_nulls) and a group that met a null holds None (ignore_nulls=False)
        (null_aggregate.py:29-65, 93-138)."""
        flat = self.to_flat_arrow()
        if self.aggs is None:
            return flat
        n = self._num_rows
        arrays, names = [], []
        if self.key is not None:
            arrays.append(flat[self.key].combine_chunks())
            names.append(self.key)
        for name, agg in zip(resolve_agg_names(self.aggs), self.aggs):
            parts = {p: flat[f"{name}#{p}"].to_numpy(zero_copy_only=False) for p in agg._device_merge_ops()}
            names.append(name)
            if set(parts) == {"count"}:
                arrays.append(pa.array(parts["count"], pa.int64()))
                continue
            kind = agg._kind
            if kind == "mean":
                cols, init = [parts["sum"], parts["nvalid"]], [0, 0]
            elif kind == "std":
                cols, init = [parts["m2"], parts["mean"], parts["nvalid"]], [0, 0, 0]
            else:
                cols = [parts[kind]]
                init = [{"sum": 0, "absmax": 0, "min": np.inf, "max": -np.inf}[kind]]
            has = parts["nvalid"] > 0
            null = np.zeros(n, dtype=bool) if agg._ignore_nulls else parts["nvalid"] < parts["nrows"]
            # integer columns keep integer accumulators (list<int64>); their min / max have no
            # integer "empty" value (the reference holds +-inf there, which turns that one list
            # into floats): an integer group without valid values exports 0 with has_data 0
            is_int = all(c.dtype.kind in "iu" for c in cols)
            if is_int:
                init = [v if np.isfinite(v) else 0 for v in init]
            dt = np.int64 if is_int else np.float64
            width = len(cols) + 1
            mat = np.empty((n, width), dtype=dt)
            for c, (col, v0) in enumerate(zip(cols, init)):
                mat[:, c] = np.where(has, col.astype(dt), dt(v0) if is_int else v0)
            mat[:, width - 1] = has
            offsets = pa.array(np.arange(n + 1, dtype=np.int32) * width, pa.int32())
            arrays.append(pa.ListArray.from_arrays(offsets, pa.array(mat.reshape(-1)),
                                                   mask=pa.array(null) if null.any() else None))
        return pa.Table.from_arrays(arrays, names=names)


class DeviceArrowBlockAccessor:
    """BlockAccessor for DeviceArrowBlock (python/ray/data/block.py:233-422)."""

    def __init__(self, block: DeviceArrowBlock):
        self._block = block

    # -- plain accessors -------------------------------------------------------------
    def num_rows(self) -> int:
        return self._block.num_rows

    def size_bytes(self) -> int:
        return self._block.nbytes  # arrow_block.py:231-232 (table.nbytes)

    def schema(self) -> pa.Schema:
        return self._block.schema

    def to_block(self) -> DeviceArrowBlock:
        return self._block

    def to_arrow(self) -> pa.Table:
        return self._block.to_arrow()

    def to_numpy(self, columns=None):
        t = self._block.to_arrow()
        if columns is None:
            columns = t.column_names
        single = not isinstance(columns, list)
        cols = [columns] if single else columns
        for c in cols:
            if c not in t.column_names:
                raise ValueError(f"Cannot find column {c}, available columns: {t.column_names}")
        arrs = [t[c].combine_chunks().to_numpy(zero_copy_only=False) for c in cols]
        return arrs[0] if len(arrs) == 1 else dict(zip(cols, arrs))

    def get_metadata(self, input_files, exec_stats) -> BlockMetadata:
        return BlockMetadata(
            num_rows=self.num_rows(),
            size_bytes=self.size_bytes(),
            schema=self.schema(),
            input_files=input_files,
            exec_stats=exec_stats,
        )

    def slice(self, start: int, end: int, copy: bool = False) -> DeviceArrowBlock:
        """arrow_block.py:178-182.  Device slices are views; ``copy=True`` clones."""
        view = self._block.slice(start, end)
        if copy:
            cols = collections.OrderedDict((n, c.clone()) for n, c in view.columns.items())
            view = DeviceArrowBlock(cols, view.num_rows)
        return view

    def take(self, indices) -> DeviceArrowBlock:
        """arrow_block.py:258-267."""
        if not isinstance(indices, torch.Tensor):
            indices = torch.as_tensor(np.asarray(indices, dtype=np.int64).astype(np.int32))
        idx = indices.to(device="cuda", dtype=torch.int32)
        return _take_block(self._block, idx)

    @staticmethod
    def _empty_table() -> DeviceArrowBlock:
        return DeviceArrowBlock.empty()

    # -- sampling (table_block.py:186-196, arrow_block.py:269-272) ------------------
    def sample(self, n_samples: int, key: SortKeyT) -> DeviceArrowBlock:
        if key is None or callable(key):
            raise NotImplementedError(f"Table sort key must be a column name, was: {key}")
        if self.num_rows() == 0:
            return self._empty_table()
        k = min(n_samples, self.num_rows())
        names = [key] if isinstance(key, str) else [kk[0] for kk in key]
        seed = random.getrandbits(63)  # the reference draws from the global `random` too
        cols = collections.OrderedDict()
        for name in names:
            c = self._block.column(name)
            if c.is_varlen:  # sample row ids with the same kernel, then take those rows
                rows = torch.arange(c.num_rows, dtype=torch.int32, device="cuda")
                picked = torch.empty(k, dtype=torch.int32, device="cuda")
                if k:
                    _lib.call("ds_sample_rows", rows.data_ptr(), 4, c.num_rows, k, seed, picked.data_ptr(), _stream())
                cols[name] = _gather_varlen(c, picked)
                continue
            out = _empty_u8(k * c.row_bytes)
            valid = None
            if k:
                _lib.call("ds_sample_rows", c.ptr, c.row_bytes, c.num_rows, k, seed, out.data_ptr(), _stream())
                if c.valid is not None:  # same seed -> same rows
                    valid = _empty_u8(k)
                    _lib.call("ds_sample_rows", c.valid_ptr, 1, c.num_rows, k, seed, valid.data_ptr(), _stream())
            cols[name] = DeviceColumn(c.type, out, valid)
        return DeviceArrowBlock(cols, k)

    # -- sort (arrow_block.py:353-394) ------------------------------------------------
    def _sorted(self, col: str, descending: bool, with_valid_count: bool = False):
        keycol = self._block.column(col)
        idx, sorted_key, nv = _argsort_nulls_last(keycol, descending)
        out = _take_block(self._block, idx, ready={col: sorted_key})
        return (out, nv) if with_valid_count else out

    def sort_and_partition(self, boundaries: List, key: SortKeyT, descending: bool) -> List[DeviceArrowBlock]:
        if not isinstance(key, str) and key is not None and not callable(key) and len(key) > 1:
            raise NotImplementedError("sorting by multiple columns is not supported yet")
        if self._block.num_rows == 0:
            return [self._empty_table() for _ in range(len(boundaries) + 1)]
        col, _ = _normalize_key(key)
        table, nv = self._sorted(col, descending, with_valid_count=True)
        if len(boundaries) == 0:
            return [table]
        # null keys sit behind the last cut: they fall into the last partition
        bounds = _searchsorted(table.column(col).slice(0, nv), boundaries, descending)
        parts, last = [], 0
        for b in bounds:
            b = max(b, last)  # cuts never go backwards, whatever the splitters are
            parts.append(table.slice(last, b))
            last = b
        parts.append(table.slice(last, table.num_rows))
        return parts

    @staticmethod
    def sort_blocks_and_partition(blocks: List[DeviceArrowBlock], boundaries: List, key: SortKeyT,
                                  descending: bool) -> List[DeviceArrowBlock]:
        """sort_and_partition of the CONCATENATION of ``blocks`` without building it: only the key
        columns are concatenated (they have to be contiguous for the radix sort); every other
        column is gathered straight from the blocks (ds_gather_rows_multi) where its shape allows,
        which saves one full copy of the dataset."""
        if not isinstance(key, str) and key is not None and not callable(key) and len(key) > 1:
            raise NotImplementedError("sorting by multiple columns is not supported yet")
        col, _ = _normalize_key(key)
        blocks = [b for b in blocks if b.num_rows > 0]
        names = blocks[0].column_names
        if len(blocks) == 1 or any(b.column_names != names for b in blocks):
            return DeviceArrowBlockAccessor(DeviceArrowBlock.concat(blocks)).sort_and_partition(
                boundaries, key, descending)
        keycol = DeviceArrowBlock.concat([b.select([col]) for b in blocks]).column(col)
        idx, sorted_key, nv = _argsort_nulls_last(keycol, descending)
        cols = collections.OrderedDict()
        for name in names:
            if name == col:
                cols[name] = sorted_key
                continue
            parts = [b.columns[name] for b in blocks]
            if _multi_gather_ok(parts):
                cols[name] = _gather_multi(parts, idx)
            else:
                cols[name] = _gather(DeviceArrowBlock.concat([b.select([name]) for b in blocks]).column(name), idx)
        table = DeviceArrowBlock(cols, idx.numel())
        if len(boundaries) == 0:
            return [table]
        bounds = _searchsorted(table.column(col).slice(0, nv), boundaries, descending)
        out, last = [], 0
        for b in bounds:
            b = max(b, last)  # cuts never go backwards, whatever the splitters are
            out.append(table.slice(last, b))
            last = b
        out.append(table.slice(last, table.num_rows))
        return out

    def partition_perm(self, boundaries: List, key: SortKeyT, descending: bool):
        """(perm, counts) of the stable range partition: ``perm`` lists the row ids grouped
        by partition with input order kept inside each; null keys go, in input order, to the
        end of the last partition."""
        col, _ = _normalize_key(key)
        keycol = self._block.column(col)
        if keycol.valid is None:
            return _range_partition(keycol, boundaries, descending)
        n = keycol.num_rows
        perm0, nv = _split_nulls(keycol)
        head = _gather(keycol.with_valid(None), perm0[:nv])
        permv, counts = _range_partition(head, boundaries, descending)
        perm = torch.empty(n, dtype=torch.int32, device="cuda")
        perm[:nv] = _gather_raw(perm0.view(torch.uint8), 4, n, permv).view(torch.int32)
        perm[nv:] = perm0[nv:]
        counts[-1] += n - nv
        return perm, counts

    def partition_unsorted(self, boundaries: List, key: SortKeyT, descending: bool) -> List[DeviceArrowBlock]:
        """Stable range partition WITHOUT sorting: same partition rule and same
        in-partition input order as slicing the stably sorted block would give, up to
        the order inside a partition (which the reducer's stable sort restores).  Used by
        the device plan, which sorts once on the reduce side."""
        nparts = len(boundaries) + 1
        if self._block.num_rows == 0:
            return [self._empty_table() for _ in range(nparts)]
        if nparts == 1:
            return [self._block]
        perm, counts = self.partition_perm(boundaries, key, descending)
        table = _take_block(self._block, perm)
        parts, last = [], 0
        for c in counts:
            parts.append(table.slice(last, last + c))
            last += c
        return parts

    @staticmethod
    def merge_sorted_blocks(blocks: List[Block], key: SortKeyT, _descending: bool):
        """arrow_block.py:473-486: concat in mapper order + stable sort."""
        stats = BlockExecStats.builder()
        blocks = [BlockAccessor.for_block(b).to_block() for b in blocks]
        blocks = [b for b in blocks if b.num_rows > 0]
        if len(blocks) == 0:
            ret = DeviceArrowBlock.empty()
        else:
            col, desc_from_key = _normalize_key(key)
            # (no concatenated copy of the payload: the columns are gathered from the blocks)
            ret = DeviceArrowBlockAccessor.sort_blocks_and_partition(
                blocks, [], col, desc_from_key if not isinstance(key, str) else _descending
            )[0]
        return ret, DeviceArrowBlockAccessor(ret).get_metadata(None, exec_stats=stats.build())

    # -- random_shuffle (arrow_block.py:184-186) ----------------------------------------
    def random_shuffle(self, random_seed: Optional[int]) -> DeviceArrowBlock:
        if random_seed is None:
            random_seed = int.from_bytes(os.urandom(8), "little")
        n = self.num_rows()
        if n == 0:
            return self._block
        perm = _random_permutation(n, int(random_seed))
        return _take_block(self._block, perm)

    # -- groupby (arrow_block.py:396-467, 488-587) ----------------------------------------
    def combine(self, key: Optional[str], aggs: Tuple[AggregateFn]) -> PartialAggBlock:
        """Partial aggregation of a key-sorted block: one row per run of equal keys."""
        if key is not None and not isinstance(key, str):
            raise ValueError(
                "key must be a string or None when aggregating on Arrow blocks, but "
                f"got: {type(key)}."
            )
        blk = self._block
        n = blk.num_rows
        cols = collections.OrderedDict()
        if n == 0:
            return PartialAggBlock(cols, 0, key, aggs)
        if key is None:
            starts = torch.tensor([0, n], dtype=torch.int32, device="cuda")
            nseg = 1
        else:
            keycol = blk.column(key)
            _key_kind(_sortable(keycol).type)  # raises for key types the device path cannot order
            if keycol.valid is not None and not bool(keycol.valid.all()):
                raise NotImplementedError("null groupby keys are not supported on device yet")
            starts, nseg = _segment_heads(keycol)
            cols[key] = _gather(keycol, starts[:nseg])
        for name, agg in zip(resolve_agg_names(aggs), aggs):
            for part, col in agg._device_accumulate(blk, starts, nseg, n).items():
                cols[f"{name}#{part}"] = col
        return PartialAggBlock(cols, nseg, key, aggs)

    @staticmethod
    def aggregate_combined_blocks(blocks: List[PartialAggBlock], key: Optional[str],
                                  aggs: Tuple[AggregateFn], finalize: bool):
        """Merge partials of one key range (arrow_block.py:488-587): concat in mapper
        order, stable sort by key, merge accumulators of equal keys left to right."""
        stats = BlockExecStats.builder()
        blocks = [b for b in blocks if b.num_rows > 0]
        names = resolve_agg_names(aggs)
        if not blocks:
            ret = DeviceArrowBlock.empty()
            return ret, DeviceArrowBlockAccessor(ret).get_metadata(None, exec_stats=stats.build())
        cat = DeviceArrowBlock.concat(blocks)
        n = cat.num_rows
        cols = collections.OrderedDict()
        if key is None:
            idx = None
            starts = torch.tensor([0, n], dtype=torch.int32, device="cuda")
            nseg = 1
        else:
            keycol = cat.columns[key]
            if len(blocks) == 1:
                idx, sorted_key = None, keycol  # a single combined block is already key-sorted
            else:
                idx, sorted_key = _argsort(keycol, False, want_sorted=True)
            starts, nseg = _segment_heads(sorted_key)
            cols[key] = _gather(sorted_key, starts[:nseg])
        for name, agg in zip(names, aggs):
            if hasattr(agg, "_device_merge"):  # merges that are not plain per-part reductions
                parts = agg._device_merge(lambda p: cat.columns[f"{name}#{p}"], idx, starts, nseg, n)
            else:
                parts = {
                    p: _segmented_reduce(cat.columns[f"{name}#{p}"], idx, starts, nseg, n, op)
                    for p, op in agg._device_merge_ops().items()
                }
            if finalize:
                cols[name] = agg._device_finalize(parts)
            else:
                for p, c in parts.items():
                    cols[f"{name}#{p}"] = c
        ret = DeviceArrowBlock(cols, nseg) if finalize else PartialAggBlock(cols, nseg, key, aggs)
        return ret, DeviceArrowBlockAccessor(ret).get_metadata(None, exec_stats=stats.build())


def resolve_agg_names(aggs) -> List[str]:
    """arrow_block.py:456-463,469-471: duplicate names become name_2, name_3, ..."""
    count = collections.defaultdict(int)
    out = []
    for agg in aggs:
        name = agg.name
        if count[name] > 0:
            name = f"{name}_{count[name] + 1}"
        count[agg.name] += 1
        out.append(name)
    return out


def _validate_key_fn(ds, key) -> None:
    """python/ray/data/block.py:60-93 for table datasets: a key is a column name that exists
    in the schema; ``None`` and callables belong to simple (list) datasets."""
    schema = ds.schema()
    if schema is None:  # empty or cleared dataset: validation not possible
        return
    if isinstance(key, str):
        if len(schema.names) > 0 and key not in schema.names:
            raise ValueError("The column '{}' does not exist in the schema '{}'.".format(key, schema))
    elif key is None:
        raise ValueError("The `None` key '{}' requires dataset format to be 'simple', was 'arrow'.".format(key))
    elif callable(key):
        raise ValueError("Callable key '{}' requires dataset format to be 'simple', was 'arrow'.".format(key))
    else:
        raise TypeError("Invalid key type {} ({}).".format(key, type(key)))


class BlockAccessor:
    """python/ray/data/block.py:370-394: dispatch on the block kind."""

    @staticmethod
    def for_block(block) -> DeviceArrowBlockAccessor:
        if isinstance(block, DeviceArrowBlock):
            return DeviceArrowBlockAccessor(block)
        if isinstance(block, pa.Table):
            # host Arrow tables enter the device path through an H2D import
            return DeviceArrowBlockAccessor(DeviceArrowBlock.from_arrow(block))
        raise TypeError(f"Not a device block type: {block!r} ({type(block)})")


class BlockList:
    """python/ray/data/_internal/block_list.py:10-60,172-220 -- the subset the plans use.
    Holds device blocks directly (there is no object store between stages)."""

    def __init__(self, blocks: List[Block], metadata: List[BlockMetadata], *, owned_by_consumer: bool = False):
        assert len(blocks) == len(metadata), (blocks, metadata)
        self._blocks = blocks
        self._num_blocks = len(blocks)
        self._metadata = metadata
        self._owned_by_consumer = owned_by_consumer

    def get_metadata(self, fetch_if_missing: bool = False) -> List[BlockMetadata]:
        return self._metadata.copy()

    def copy(self) -> "BlockList":
        return BlockList(self._blocks, self._metadata, owned_by_consumer=self._owned_by_consumer)

    def clear(self) -> None:
        self._blocks = None

    def is_cleared(self) -> bool:
        return self._blocks is None

    def _check_if_cleared(self) -> None:
        if self.is_cleared():
            raise ValueError(
                "This Dataset's blocks have been moved, which means that you "
                "can no longer use this Dataset."
            )

    def get_blocks(self) -> List[Block]:
        self._check_if_cleared()
        return list(self._blocks)

    def initial_num_blocks(self) -> int:
        return self._num_blocks

    def executed_num_blocks(self) -> int:
        return len(self.get_blocks())

    @staticmethod
    def from_blocks(blocks: Sequence[Block]) -> "BlockList":
        blocks = [BlockAccessor.for_block(b).to_block() for b in blocks]
        metas = [DeviceArrowBlockAccessor(b).get_metadata(None, None) for b in blocks]
        return BlockList(blocks, metas, owned_by_consumer=False)
