"""Multi-GPU plans: one process per GPU, the M x N block exchange as an all-to-all-v.

In the reference the exchange between map and reduce tasks is implicit: every reduce
task lists M object refs and the object manager pulls the chunks
(python/ray/data/_internal/shuffle.py:106-112, src/ray/object_manager/object_manager.cc:336-559).
Here every rank holds the input blocks of its position in block order, the partitions a
rank produces are contiguous slices of one device buffer per column, and the exchange is
``torch.distributed.all_to_all_single`` with per-peer split sizes -- on the NCCL backend a
grouped ncclSend/ncclRecv over NVLink -- after a tiny all-gather of the count matrix.
Receivers get the chunks concatenated in source-rank order, which together with the
stable partition and the stable sort makes the global result the stable sort of the
concatenated input (ties keep (rank, row) order), bit-exact to Dataset.sort.

The sort plan sends the key column first and starts the local radix sort of the received
keys while the 9x larger payload column is still in flight on NCCL's stream.

Only the host logic here (splitters from gathered samples, count matrix, split sizes) is
backend independent and is what the gloo CPU tests exercise; the compute is CUDA only.
"""
import collections
from typing import List, Optional, Sequence, Tuple

import numpy as np
import pyarrow as pa
import torch
import torch.distributed as dist

SAMPLES_PER_RANK = 8192

# Optional phase timing (DS_DIST_TIMING=1): CUDA events on the current stream; the list of
# (phase, ms) of the last call is left in LAST_TIMING.  Measurement aid only.
import os as _os

_TIMING = _os.environ.get("DS_DIST_TIMING", "0") == "1"
LAST_TIMING: list = []
# bench.py: when a list, every split-plan sort appends its exchange events here (exchange_log)
EXCHANGE_LOG = None
# measurement aid: issue the payload transfer AFTER the key sort instead of beside it
PAYLOAD_AFTER_SORT = False


def set_phase_timing(on: bool) -> None:
    global _TIMING
    _TIMING = bool(on)


class _Phases:
    """Phase boundaries of one multi-GPU plan: CUDA events when phase timing is on, and one
    NVTX range per phase when ranges are on (each mark closes the previous phase's range)."""

    def __init__(self):
        self.ev = []
        self.open = False

    def mark(self, name):
        from . import _nvtx

        if _nvtx.enabled():
            if self.open:
                torch.cuda.nvtx.range_pop()
            torch.cuda.nvtx.range_push("after:" + name)
            self.open = True
        if _TIMING:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self.ev.append((name, e))

    def finish(self):
        if self.open:
            torch.cuda.nvtx.range_pop()
            self.open = False
        if _TIMING and self.ev:
            torch.cuda.synchronize()
            LAST_TIMING[:] = [(self.ev[i + 1][0], self.ev[i][1].elapsed_time(self.ev[i + 1][1]))
                              for i in range(len(self.ev) - 1)]


# ----------------------------------------------------------------------------- host logic
def exchange_counts(counts: Sequence[int], group=None, device=None) -> np.ndarray:
    """All-gather of each rank's per-destination row counts -> W x W matrix M with
    M[src, dst] = rows src sends to dst."""
    world = dist.get_world_size(group)
    assert len(counts) == world
    mine = torch.tensor(list(counts), dtype=torch.int64, device=device)
    allc = torch.empty(world * world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(allc, mine, group=group)
    return allc.cpu().numpy().reshape(world, world)


def alltoallv_rows(send: torch.Tensor, row_bytes: int, matrix: np.ndarray, group=None,
                   async_op: bool = False):
    """Exchange rows of ``row_bytes`` bytes: ``send`` holds this rank's partitions back to
    back (partition d goes to rank d).  Returns (recv uint8 tensor, work handle or None);
    recv holds the chunks of ranks 0..W-1 back to back."""
    rank = dist.get_rank(group)
    send_rows = [int(x) for x in matrix[rank]]
    recv_rows = [int(x) for x in matrix[:, rank]]
    assert send.numel() == sum(send_rows) * row_bytes, (send.numel(), send_rows, row_bytes)
    recv = torch.empty(sum(recv_rows) * row_bytes, dtype=torch.uint8, device=send.device)
    work = dist.all_to_all_single(
        recv, send,
        output_split_sizes=[r * row_bytes for r in recv_rows],
        input_split_sizes=[r * row_bytes for r in send_rows],
        group=group, async_op=async_op,
    )
    return recv, work


def gather_samples(local: np.ndarray, width: int, group=None, device=None) -> np.ndarray:
    """All-gather of up to SAMPLES_PER_RANK sampled keys (rows of ``width`` bytes) per
    rank; ranks may contribute fewer.  Returns the concatenation in rank order."""
    world = dist.get_world_size(group)
    k = local.shape[0]
    buf = np.zeros((SAMPLES_PER_RANK + 1, width), dtype=np.uint8)
    buf[:k] = local
    buf[SAMPLES_PER_RANK, : min(width, 4)] = np.frombuffer(np.uint32(k).tobytes(), np.uint8)[: min(width, 4)]
    cnt = torch.tensor([k], dtype=torch.int64, device=device)
    cnts = torch.empty(world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(cnts, cnt, group=group)
    mine = torch.from_numpy(buf[:SAMPLES_PER_RANK].reshape(-1).copy()).to(device)
    allb = torch.empty(world * mine.numel(), dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(allb, mine, group=group)
    allb = allb.cpu().numpy().reshape(world, SAMPLES_PER_RANK, width)
    return np.concatenate([allb[r, : int(c)] for r, c in enumerate(cnts.cpu().tolist())], axis=0)


def splitters_from_samples(samples: np.ndarray, arrow_type: pa.DataType, world: int) -> list:
    """Identical W-1 splitters on every rank: the i/W quantiles (i = 1..W-1) of the sorted
    gathered samples.

    The reference takes the quantiles i/(R-1), i = 1..R-1, of 10 samples per block
    (python/ray/data/_internal/sort.py:110-115), whose last splitter is the largest sample,
    so its last partition is almost empty and the others are uneven ("splitter
    quality").  Output block sizes are not part of its contract -- only the concatenation
    is -- so the multi-GPU plan uses evenly spaced quantiles of 2048 samples per rank."""
    from .sort import fsb_order_statistics, sort_samples

    n, w = samples.shape
    if n == 0:
        return [None] * (world - 1)
    picks = [min(n - 1, (i * n) // world) for i in range(1, world)]
    if pa.types.is_fixed_size_binary(arrow_type):  # only the W-1 picked rows become Python objects
        return fsb_order_statistics(np.ascontiguousarray(samples), picks)
    buf = pa.py_buffer(np.ascontiguousarray(samples).tobytes())
    srt = sort_samples(pa.Array.from_buffers(arrow_type, n, [None, buf]))
    return [srt[i] for i in picks]


# ----------------------------------------------------------------------------- peer-memory exchange
# The all-to-all-v without NCCL: every rank owns one symmetric (peer-mapped) receive buffer;
# a sender copies each of its partitions straight into the destination GPU's buffer at the
# offset its source rank owns there (copy engines over NVLink / NVSwitch, no SMs, no staging
# on the receiver), so the receiver finds the chunks already concatenated in source-rank
# order.  Device-side barriers on the buffer's signal pad order the phases.  Measured on
# 2 x B200 this moves the 9 GB payload exchange from 28 ms (NCCL send/recv sharing SMs with
# the radix sort) to copy-engine speed while the key sort runs.  (A variant that also cut
# the partition into 4 row ranges and pipelined gather / copy per range was measured slower
# on 2 and on 8 GPUs -- the ranges' gathers compete with the key sort for HBM -- and was
# dropped; see profiles/r01_multi_gpu.md.)
# "split" (default): streaming split with bulk stores into peer memory (_sort_block_split);
# "p2p": materialised partition + copy engines; "fused": per-thread peer stores; "nccl"
_EXCHANGE = _os.environ.get("DS_EXCHANGE", "split")
# persistent CTAs of the payload split that runs beside the key sort (0 = one per SM)
_SPLIT_CTAS = int(_os.environ.get("DS_SPLIT_CTAS", "-1"))  # -1: one per SM
_SPLIT_TILE = int(_os.environ.get("DS_SPLIT_TILE", "0"))   # tuning: tile = k * 256 rows
_EARLY_PAYLOAD = int(_os.environ.get("DS_EARLY_PAYLOAD", "0"))  # 1: payload beside the key transfer
_LANES = _os.environ.get("DS_P2P_LANES", "1") == "1"  # one copy stream per destination
_peer_cache: dict = {}


class _PeerBuffer:
    def __init__(self, nbytes: int, group):
        import torch.distributed._symmetric_memory as symm_mem

        self.nbytes = nbytes
        pg = group if group is not None else dist.group.WORLD
        enable = getattr(symm_mem, "enable_symm_mem_for_group", None)
        if enable is not None:  # needed by older torch, a deprecated no-op in newer ones
            import warnings

            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                try:
                    enable(pg.group_name)
                except Exception:
                    pass
        self.buf = symm_mem.empty(nbytes, dtype=torch.uint8, device=_dev())
        self.hdl = symm_mem.rendezvous(self.buf, pg)
        self.side = torch.cuda.Stream()
        self.lanes = [torch.cuda.Stream() for _ in range(self.hdl.world_size)]  # one per destination


def _peer_buffer(nbytes: int, group) -> "_PeerBuffer":
    """Grow-only cache; collective (every rank calls with the same nbytes)."""
    key = id(group) if group is not None else 0
    pb = _peer_cache.get(key)
    if pb is None or pb.nbytes < nbytes:
        _peer_cache.pop(key, None)
        pb = _PeerBuffer(int(nbytes * 1.25) + 4096, group)
        _peer_cache[key] = pb
    return pb


def _align(x: int, a: int = 256) -> int:
    return (x + a - 1) // a * a


class _PeerLayout:
    """Where each column's received rows live inside the symmetric buffer."""

    def __init__(self, names, row_bytes, matrix: np.ndarray):
        self.cap_rows = int(matrix.sum(axis=0).max())
        self.off, total = {}, 0
        for nme in names:
            self.off[nme] = total
            total += _align(self.cap_rows * row_bytes[nme])
        self.total = max(total, 256)


def _p2p_send_column(pb, layout, name, rb, send: torch.Tensor, matrix: np.ndarray, rank: int,
                     after: "torch.cuda.Stream"):
    """Copy this rank's partitions of one column into every destination's buffer, one CUDA
    stream per destination so that the copy engines drive all NVLink directions at once.
    The copies start after the work already queued on ``after``."""
    world = matrix.shape[0]
    offs = np.concatenate([[0], np.cumsum(matrix[rank])])
    for step in range(world):
        d = (rank + step) % world
        rows = int(matrix[rank, d])
        if rows == 0:
            continue
        before = int(matrix[:rank, d].sum())  # rows lower ranks place in d's buffer
        dst = pb.hdl.get_buffer(d, (rows * rb,), torch.uint8, layout.off[name] + before * rb)
        lane = pb.lanes[d] if _LANES else pb.lanes[0]
        lane.wait_stream(after)
        with torch.cuda.stream(lane):
            dst.copy_(send[int(offs[d]) * rb: int(offs[d + 1]) * rb], non_blocking=True)
        send.record_stream(lane)


def _join_lanes(pb, stream):
    for lane in pb.lanes:
        stream.wait_stream(lane)


def _p2p_exchange_all(names, rbs, send, matrix, group):
    """Exchange every column through peer memory; returns {name: view of this rank's
    receive buffer} (valid until the next exchange on this group), or None when peer memory
    cannot be used (the caller then takes the NCCL path)."""
    world = matrix.shape[0]
    if not (_EXCHANGE in ("p2p", "fused", "split") and world > 1 and dist.get_backend(group) == "nccl"):
        return None
    try:
        layout = _PeerLayout(names, rbs, matrix)
        pb = _peer_buffer(layout.total, group)
    except Exception as e:  # pragma: no cover - depends on the system
        import warnings

        warnings.warn(f"peer-memory exchange unavailable ({e!r}); using NCCL all_to_all")
        return None
    rank = dist.get_rank(group)
    main = torch.cuda.current_stream()
    pb.hdl.barrier(channel=0)
    for name in names:
        _p2p_send_column(pb, layout, name, rbs[name], send[name], matrix, rank, main)
    _join_lanes(pb, main)
    pb.hdl.barrier(channel=1)
    n_recv = int(matrix[:, rank].sum())
    return {nm: pb.buf[layout.off[nm]: layout.off[nm] + n_recv * rbs[nm]] for nm in names}


# ----------------------------------------------------------------------------- plans
def _dev():
    return torch.device("cuda", torch.cuda.current_device())


def _sample_local(block, col: str) -> np.ndarray:
    from .block import DeviceArrowBlockAccessor

    c = block.column(col) if block.num_rows else None
    if c is None:
        return np.zeros((0, 1), dtype=np.uint8)
    k = min(SAMPLES_PER_RANK, block.num_rows)
    s = DeviceArrowBlockAccessor(block).sample(k, col)
    rows = s.columns[col].data.cpu().numpy().reshape(k, c.row_bytes)
    if s.columns[col].valid is not None:  # null keys take no part in the splitters
        rows = rows[s.columns[col].valid.cpu().numpy() != 0]
    return rows


def _key_width(block, col, group) -> Tuple[int, pa.DataType]:
    return block.column(col).row_bytes, block.column(col).type


_VALID = "\x00valid"  # wire-name suffix of a column's validity bytes


_LEN = "\x00len"  # wire-name suffix of a variable-length column's row lengths


def _fixed_width_only(block, what: str, key: Optional[str] = None, payload_ok: bool = False,
                      key_ok: bool = False) -> List[str]:
    """Variable-length columns travel as a fixed-width wire of row lengths plus one
    all-to-all-v of their bytes (``payload_ok``); a variable-length KEY (``key_ok``: sort only)
    also needs variable-length splitters, which the ranks agree on through the host.
    Returns the names of the variable-length columns."""
    names = []
    for name, c in block.columns.items():
        if not c.is_varlen:
            continue
        if payload_ok and (name != key or key_ok):
            names.append(name)
            continue
        raise NotImplementedError(
            f"{what}: column {name!r} has the variable-length type {c.type}; the multi-GPU "
            "plans take variable-length columns in sort (key or payload) and as payload of random_shuffle")
    return names


def _varlen_send(whole, row_bounds: Sequence[int], dev):
    """``whole``: a variable-length column in send order, partition d = rows
    row_bounds[d]:row_bounds[d+1].  Returns (row lengths as an 8-byte-per-row wire, the bytes
    in the same order, bytes per destination)."""
    off = whole.offsets
    lens = (off[1:] - off[:-1]).contiguous()
    at = torch.as_tensor(list(row_bounds), dtype=torch.int64, device=dev)
    byte_bounds = [int(x) for x in off[at].cpu().tolist()]  # host sync: W + 1 numbers
    data = whole.data[byte_bounds[0]: byte_bounds[-1]]
    return lens.view(torch.uint8), data, [byte_bounds[d + 1] - byte_bounds[d] for d in range(len(row_bounds) - 1)]


def _varlen_recv(arrow_type, len_wire: torch.Tensor, n_recv: int, data, byte_counts, valid, group, dev):
    """Received lengths (chunks of ranks 0..W-1 back to back) -> offsets; the bytes follow in
    the same order through one all-to-all-v with byte granularity."""
    from .block import DeviceVarColumn

    lens = len_wire[: n_recv * 8].view(torch.int64)
    offsets = torch.zeros(n_recv + 1, dtype=torch.int64, device=dev)
    if n_recv:
        torch.cumsum(lens, 0, out=offsets[1:])
    bm = exchange_counts(byte_counts, group, dev)
    recv, _ = alltoallv_rows(data.contiguous(), 1, bm, group)
    return DeviceVarColumn(arrow_type, offsets, recv, valid, int(bm[:, dist.get_rank(group)].sum()))


def _varlen_splitters(block, key: str, world: int, group) -> list:
    """W-1 splitters of a string / binary key column, identical on every rank: up to
    SAMPLES_PER_RANK sampled keys per rank travel as Python objects (all_gather_object), are
    ordered bytewise -- Arrow's order for binary and for UTF-8 strings -- and the i/W quantiles
    are picked (python/ray/data/_internal/sort.py:105-115 does the same on the driver)."""
    from .block import DeviceArrowBlockAccessor

    mine = []
    if block.num_rows:
        s = DeviceArrowBlockAccessor(block).sample(min(SAMPLES_PER_RANK, block.num_rows), key)
        mine = [v for v in s.to_arrow()[key].to_pylist() if v is not None]
    gathered = [None] * world
    dist.all_gather_object(gathered, mine, group=group)
    allv = [v for part in gathered for v in part]
    if not allv:
        return [None] * (world - 1)
    allv.sort(key=lambda v: v.encode("utf-8") if isinstance(v, str) else bytes(v))
    n = len(allv)
    return [allv[min(n - 1, (i * n) // world)] for i in range(1, world)]


def _agree_on_validity(block, group, dev):
    """Every rank must exchange the same set of buffers: a column carries validity on the
    wire if it has a mask on ANY rank (ranks without one send all-ones)."""
    from .block import DeviceArrowBlock, DeviceColumn

    flags = torch.tensor([int(c.valid is not None) for c in block.columns.values()],
                         dtype=torch.int32, device=dev)
    if flags.numel():
        dist.all_reduce(flags, op=dist.ReduceOp.MAX, group=group)
    need = [bool(x) for x in flags.cpu().tolist()]
    if not any(need):
        return block, set()
    cols = collections.OrderedDict()
    for (name, c), nd in zip(block.columns.items(), need):
        if nd and c.valid is None:
            c = c.with_valid(torch.ones(c.num_rows, dtype=torch.uint8, device=dev))
        cols[name] = c
    return DeviceArrowBlock(cols, block.num_rows), {n for n, nd in zip(block.columns, need) if nd}


def _wire_columns(parts, names, with_valid, key, dev, row_bytes=None):
    """[(wire name, row bytes, send buffer)] with the key column(s) first.  ``row_bytes``:
    {column: bytes per row} from the block's schema, so that a rank without rows puts the same
    widths on the wire as the others."""
    data = _spanning(parts, names, dev)
    wires = []
    order = [key] + [n for n in names if n != key] if key is not None else list(names)
    live = [p for p in parts if p.num_rows]
    for name in order:
        if row_bytes is not None:
            rb = row_bytes[name]
        else:
            rb = live[0].columns[name].row_bytes if live else 1
        wires.append((name, rb, data[name]))
        if name in with_valid:
            vs = [p.columns[name].valid for p in live]
            v = torch.cat(vs) if len(vs) != 1 else vs[0]
            if not vs:
                v = torch.empty(0, dtype=torch.uint8, device=dev)
            wires.append((name + _VALID, 1, v.contiguous()))
    return wires


def _sort_block_fused(block, key, descending, boundaries, with_valid, group, ph, consume=False):
    """Peer-memory sort plan whose map side is ONE kernel per column that slices the block by
    destination and stores every row straight into the destination GPU's receive buffer
    (ds_scatter_rows_to_parts over peer-mapped pointers): no send buffer, no copy pass.
    Keys go first; the payload scatter runs on a second stream beside the key sort.
    Returns None when peer memory is unavailable."""
    import ctypes

    from . import _lib
    from .block import DeviceArrowBlock, DeviceArrowBlockAccessor, DeviceColumn, _argsort_nulls_last, _gather

    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = _dev()
    n = block.num_rows
    names = block.column_names
    types = {nm: block.columns[nm].type for nm in names}
    sort_key = [(key, "descending" if descending else "ascending")]
    if n:
        perm, counts = DeviceArrowBlockAccessor(block).partition_perm(boundaries, sort_key, descending)
    else:
        perm, counts = None, [0] * world
    ph.mark("partition perm")
    matrix = exchange_counts(counts, group, dev)
    # wire columns: (wire name, row bytes, source tensor), key column(s) first
    wires = []
    for name in [key] + [nm for nm in names if nm != key]:
        c = block.columns[name]
        wires.append((name, c.row_bytes, c.data))
        if name in with_valid:
            wires.append((name + _VALID, 1, c.valid))
    nkey = 2 if key in with_valid else 1
    rbs = {w: rb for w, rb, _ in wires}
    try:
        layout = _PeerLayout([w for w, _, _ in wires], rbs, matrix)
        pb = _peer_buffer(layout.total, group)
    except Exception as e:  # pragma: no cover - depends on the system
        import warnings

        warnings.warn(f"peer-memory exchange unavailable ({e!r}); using NCCL all_to_all")
        return None
    n_recv = int(matrix[:, rank].sum())
    offs = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint64)
    offs_c = (ctypes.c_uint64 * (world + 1))(*[int(x) for x in offs])
    ptrs = [int(p) for p in pb.hdl.buffer_ptrs]
    before = [int(matrix[:rank, d].sum()) for d in range(world)]
    ph.mark("counts")

    def scatter(w, rb, src):
        if n == 0:
            return
        bases = (ctypes.c_void_p * world)(*[ptrs[d] + layout.off[w] + before[d] * rb for d in range(world)])
        _lib.call("ds_scatter_rows_to_parts", src.data_ptr(), rb, n, perm.data_ptr(), n, world, offs_c,
                  bases, torch.cuda.current_stream().cuda_stream)

    main = torch.cuda.current_stream()
    pb.hdl.barrier(channel=0)  # every rank is done with the buffer's previous contents
    for w, rb, src in wires[:nkey]:
        scatter(w, rb, src)
    pb.hdl.barrier(channel=1)  # all keys have landed everywhere
    ph.mark("scatter keys")
    pb.side.wait_stream(main)
    with torch.cuda.stream(pb.side):  # payload rows travel beside the key sort
        for w, rb, src in wires[nkey:]:
            scatter(w, rb, src)
            src.record_stream(pb.side)
        if perm is not None:
            perm.record_stream(pb.side)
    recv = {w: pb.buf[layout.off[w]: layout.off[w] + n_recv * rb] for w, rb, _ in wires}

    def column(name):
        return DeviceColumn(types[name], recv[name], recv.get(name + _VALID))

    idx, sorted_key, _ = _argsort_nulls_last(column(key), descending)
    ph.mark("key sort")
    main.wait_stream(pb.side)
    pb.hdl.barrier(channel=2)  # all payload rows have landed everywhere
    ph.mark("wait payload")
    if consume:  # every row of the input now lives in some rank's receive buffer
        wires = None
        block.columns.clear()
        del perm
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
    cols = collections.OrderedDict()
    for name in names:
        cols[name] = sorted_key if name == key else _gather(column(name), idx)
    ph.mark("gather payload")
    ph.finish()
    return DeviceArrowBlock(cols, n_recv)


# Exchange measurement of the last sort_block call on this rank (always on: six CUDA events).
# Filled by _sort_block_split; read with exchange_stats() after the stream has been synchronised.
_last_exchange: dict = {}


def _exchange_entry(e: dict) -> dict:
    return {"bytes_out": e["bytes_out"], "bytes_in": e["bytes_in"], "key_bytes_out": e["key_bytes_out"],
            "ms_keys": e["k0"].elapsed_time(e["k1"]), "ms_payload": e["p0"].elapsed_time(e["p1"]),
            "ms_exposed_wait": e["w0"].elapsed_time(e["w1"]), "ms_first_to_last": e["k0"].elapsed_time(e["w1"])}


def exchange_stats() -> Optional[dict]:
    """Device-side timing of the last split-plan exchange on this rank: bytes that left this
    GPU for peers and arrived from peers, the duration of the key and payload transfers
    (CUDA events on the streams that ran them: ms_keys includes the two device barriers around
    the key transfer), how long the main stream waited for the payload after the key sort, and
    the span from the first key byte to the last payload byte.  Call after synchronising."""
    return _exchange_entry(_last_exchange) if _last_exchange else None


def exchange_log() -> list:
    """exchange_stats() of every call since EXCHANGE_LOG was set to a list (synchronise first)."""
    return [_exchange_entry(e) for e in (EXCHANGE_LOG or [])]


def _device_splitters(block, key: str, descending: bool, group, dev, nparts: Optional[int] = None):
    """nparts-1 (default W-1) splitters chosen ON THE DEVICE, identical on every rank: each rank contributes
    SAMPLES_PER_RANK slots (Philox-stratified rows of its key column; unused slots, null keys
    and NaNs are replaced by a value that sorts last and do not count), the slots are
    all-gathered, sorted with the library's own radix sort and the i/nparts quantiles of the m valid
    samples are picked with a device-side index -- no host round trip.  Replaces the driver-side
    np.sort + quantile of the reference (python/ray/data/_internal/sort.py:105-115).
    Returns (key column as the kernels compare it, kind, key bytes, bounds tensor)."""
    import random

    from . import _lib
    from .block import DeviceColumn, _argsort, _key_kind, _sortable

    world = dist.get_world_size(group)
    wide = _sortable(block.column(key))
    kind = _key_kind(wide.type)
    rb = wide.row_bytes
    S = SAMPLES_PER_RANK
    n = block.num_rows
    k = min(S, n)
    stream = torch.cuda.current_stream().cuda_stream
    pack = torch.zeros(S * rb + S, dtype=torch.uint8, device=dev)
    slots, flags = pack[: S * rb], pack[S * rb:]
    if k:
        seed = random.getrandbits(63)
        _lib.call("ds_sample_rows", wide.ptr, rb, n, k, seed, slots.data_ptr(), stream)
        if wide.valid is not None:  # same seed -> same rows
            _lib.call("ds_sample_rows", wide.valid_ptr, 1, n, k, seed, flags.data_ptr(), stream)
        else:
            flags[:k] = 1
    ok = flags != 0
    if kind == "i64":
        v = slots.view(torch.int64)
        v.copy_(torch.where(ok, v, torch.full_like(v, torch.iinfo(torch.int64).max)))
    elif kind == "f64":
        v = slots.view(torch.float64)
        ok = ok & ~torch.isnan(v)
        flags.copy_(ok.to(torch.uint8))
        v.copy_(torch.where(ok, v, torch.full_like(v, float("inf"))))
    else:
        v = slots.view(S, rb)
        v.copy_(torch.where(ok[:, None], v, torch.full_like(v, 255)))
    allp = torch.empty(world * pack.numel(), dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(allp, pack, group=group)
    allp = allp.view(world, S * rb + S)
    allslots = allp[:, : S * rb].reshape(world * S * rb)
    m = allp[:, S * rb:].to(torch.int64).sum()
    _, srt = _argsort(DeviceColumn(wide.type, allslots), False, True, options=_lib.DS_SORT_NO_PASS_SKIP)
    nparts = world if nparts is None else nparts
    picks = (torch.arange(1, nparts, dtype=torch.int64, device=dev) * m) // nparts
    picks = picks.clamp_(0, world * S - 1)
    bounds = srt.data.view(world * S, rb)[picks]
    if descending:
        bounds = bounds.flip(0)
    return wide, kind, rb, bounds.contiguous()


def _split_supported(block, key, world: int) -> bool:
    from . import _lib

    if world > _lib.DS_MAX_SPLIT_PARTS or block.num_columns == 0:
        return False
    for name, c in block.columns.items():
        if c.is_varlen or c.row_bytes > _lib.DS_SPLIT_MAX_ROW_BYTES:
            return False
    return True


def _sort_block_split(block, key, descending, group, ph, consume=False, rounds: int = 1,
                      staged: int = 0):
    """The default multi-GPU sort plan: the map side is ONE streaming pass per column
    (ds_split_plan + ds_split_rows) that stores every row, in input order, straight into the
    receive buffer of the GPU that owns its key range -- bulk stores from shared memory over
    NVLink, no partitioned copy of the block, no permutation.  Keys go first; the payload
    columns travel on a second stream beside the radix sort of the received keys.  One host
    read-back per call (the count matrix together with the validity flags); splitters are
    chosen on the device.

    ``rounds`` > 1 is the push-based form (python/ray/data/_internal/push_based_shuffle.py:
    363-526): every rank's key range is cut into ``rounds`` sub-ranges by more splitters, the
    payload travels one sub-range ("round") at a time, and the receiver finishes round r --
    gather of the rows whose keys it sorted while they were in flight -- while round r+1 is
    still arriving.  Where the reference merges the map outputs of a round and reduces the
    merged runs at the end, the rounds here ARE key sub-ranges, so the "merge" of a round is
    the sort of its received keys and the final reduce is the gather; the rounds' outputs are
    consecutive output blocks, no re-merge.

    ``staged`` (< rounds): the last ``staged`` rounds do not cross NVLink during the split pass.
    The ONE pass over the input stores the rows of the first rounds straight into the peers'
    buffers and the rows of the staged rounds into a local staging buffer (the split kernel does
    not care where a destination lives); the copy engines then ship the staged rounds, one after
    the other, while the SMs gather the rounds that have landed.  With ``staged`` = 0 every round
    is its own masked pass of the split kernel over the input.

    Returns ([output block per round], {"map", "merge", "reduce"} event pairs) or None when
    peer memory is unavailable."""
    import ctypes

    from . import _lib
    from .block import DeviceArrowBlock, DeviceColumn, _argsort_nulls_last, _gather

    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = _dev()
    n = block.num_rows
    names = block.column_names
    types = {nm: block.columns[nm].type for nm in names}
    R = max(1, min(int(rounds), _lib.DS_MAX_SPLIT_PARTS // world))
    nd = world * R  # destinations: sub-range r of rank d is destination d * R + r
    stream = lambda: torch.cuda.current_stream().cuda_stream  # noqa: E731
    wide, kind, kb, bounds = _device_splitters(block, key, descending, group, dev, nd)
    ph.mark("sample+splitters")
    dest = torch.empty(max(n, 1), dtype=torch.uint8, device=dev)
    offs = torch.empty(int(_lib.lib.ds_split_table_words(n, nd)), dtype=torch.int32, device=dev)
    # counts of this rank's rows per destination, followed by one "has validity" flag per column
    info = torch.zeros(nd + len(names), dtype=torch.int64, device=dev)
    code = {"i64": _lib.DS_KEY_I64, "f64": _lib.DS_KEY_F64, "fsb": _lib.DS_KEY_FSB}[kind]
    _lib.call("ds_split_plan", code, wide.ptr if n else None, kb, wide.valid_ptr if n else None, n,
              bounds.data_ptr() if nd > 1 else None, nd - 1, int(descending), dest.data_ptr(),
              offs.data_ptr(), info.data_ptr(), stream())
    has_valid = [int(block.columns[nm].valid is not None) for nm in names]
    if any(has_valid):
        info[nd:] = torch.tensor(has_valid, dtype=torch.int64, device=dev)
    allinfo = torch.empty(world * info.numel(), dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(allinfo, info, group=group)
    ph.mark("plan")
    allinfo = allinfo.cpu().numpy().reshape(world, -1)  # the one host read-back of the call
    matrix = allinfo[:, :nd]                            # [source rank, destination]
    need_valid = allinfo[:, nd:].max(axis=0) != 0
    with_valid = {nm for nm, f in zip(names, need_valid) if f}
    wires = []  # (wire name, row bytes, source tensor), key column first
    for name in [key] + [nm for nm in names if nm != key]:
        c = block.columns[name]
        wires.append((name, c.row_bytes, c.data))
        if name in with_valid:
            v = c.valid if c.valid is not None else torch.ones(n, dtype=torch.uint8, device=dev)
            wires.append((name + _VALID, 1, v))
    nkey = 2 if key in with_valid else 1
    rbs = {w: rb for w, rb, _ in wires}
    key_rb = sum(rb for _w, rb, _t in wires[:nkey])
    # receive layout, identical on every rank: one region per (wire, round), each with room for
    # the fullest (rank, round) anywhere
    cap_rows = int(matrix.sum(axis=0).max())
    region, total = {}, 0
    wire_rb = [(w, rb) for w, rb, _t in wires]  # names and widths only: no tensor references
    for w, rb in wire_rb:
        for r in range(R):
            region[(w, r)] = total
            total += _align(cap_rows * rb)
    try:
        pb = _peer_buffer(max(total, 256), group)
    except Exception as e:  # pragma: no cover - depends on the system
        import warnings

        warnings.warn(f"peer-memory exchange unavailable ({e!r}); using NCCL all_to_all")
        return None
    ptrs = [int(p) for p in pb.hdl.buffer_ptrs]
    before = matrix[:rank].sum(axis=0)                  # rows lower ranks place, per destination
    n_round = [int(matrix[:, rank * R + r].sum()) for r in range(R)]
    own = sum(int(matrix[rank, rank * R + r]) for r in range(R))
    remote_rows = int(matrix[rank].sum()) - own
    ph.mark("counts")

    S = max(0, min(int(staged), R - 1))
    direct_rounds = R - S
    # staging: per payload wire and staged round, this rank's rows for ranks 0..W-1 back to back
    stage_buf, stage_off = {}, {}
    for r in range(direct_rounds, R):
        cnt = [int(matrix[rank, d * R + r]) for d in range(world)]
        stage_off[r] = np.concatenate([[0], np.cumsum(cnt)])
        for w, rb in wire_rb[nkey:]:
            stage_buf[(w, r)] = torch.empty(max(int(stage_off[r][-1]) * rb, 1), dtype=torch.uint8, device=dev)

    def base_of(w, rb, j):
        d, r = divmod(j, R)
        if (w, r) in stage_buf:
            return stage_buf[(w, r)].data_ptr() + int(stage_off[r][d]) * rb
        return ptrs[d] + region[(w, r)] + int(before[j]) * rb

    def split(w, rb, src, mask, max_ctas):
        if n == 0:
            return
        bases = (ctypes.c_void_p * nd)(*[base_of(w, rb, j) for j in range(nd)])
        _lib.call("ds_split_rows", src.data_ptr(), rb, n, dest.data_ptr(), offs.data_ptr(), nd, bases,
                  mask, max_ctas, stream())

    payload_ctas = _SPLIT_CTAS
    if payload_ctas < 0:  # beside the key sort: one persistent CTA per SM keeps NVLink busy
        payload_ctas = torch.cuda.get_device_properties(dev).multi_processor_count
    if _SPLIT_TILE:
        _lib.lib.ds_debug_set_split_tile(_SPLIT_TILE)
    ev = {k: torch.cuda.Event(enable_timing=True) for k in ("k0", "k1", "p0", "p1", "w0", "w1")}
    main = torch.cuda.current_stream()
    pb.hdl.barrier(channel=0)  # every rank is done with the buffer's previous contents
    # DS_EARLY_PAYLOAD=1 starts the payload transfer beside the key transfer instead of after it.
    # Measured slower on 2 and on 8 GPUs (the key split is starved by the payload kernel: 5.8
    # instead of 1.7 ms at N=8, and the key sort starts that much later; profiles/r02_b8_early*.json),
    # so the keys get the link to themselves by default.
    early_payload = _EARLY_PAYLOAD == 1
    ev_start = torch.cuda.Event(enable_timing=True)
    ev_start.record()

    def recv(w, r):
        rb = rbs[w]
        return pb.buf[region[(w, r)]: region[(w, r)] + n_round[r] * rb]

    def column(name, r):
        v = recv(name + _VALID, r) if name in with_valid else None
        return DeviceColumn(types[name], recv(name, r), v)

    landed = [torch.cuda.Event() for _ in range(R)]

    def send_payload():
        pb.side.wait_stream(main)
        with torch.cuda.stream(pb.side):
            # (no record_stream: the main stream waits for this stream's events before the
            # function returns, so every buffer used here outlives the work that reads it)
            ev["p0"].record()
            if S == 0:
                for r in range(R):
                    mask = sum(1 << (d * R + r) for d in range(world)) if R > 1 else 0
                    for w, rb, src in wires[nkey:]:
                        split(w, rb, src, mask, payload_ctas)
                    if r == R - 1:
                        ev["p1"].record()
                    pb.hdl.barrier(channel=2 + r)  # round r has landed everywhere
                    landed[r].record()
                return
            for w, rb, src in wires[nkey:]:  # one pass: direct rounds to the peers, the rest staged
                split(w, rb, src, 0, payload_ctas)
            pb.hdl.barrier(channel=2)
            for r in range(direct_rounds):
                landed[r].record()
        for r in range(direct_rounds, R):  # the copy engines ship the staged rounds in order
            for step in range(world):
                d = (rank + step) % world
                rows = int(matrix[rank, d * R + r])
                if rows == 0:
                    continue
                lane = pb.lanes[d]
                lane.wait_stream(pb.side)
                with torch.cuda.stream(lane):
                    for w, rb in wire_rb[nkey:]:
                        dst = pb.hdl.get_buffer(d, (rows * rb,), torch.uint8,
                                                region[(w, r)] + int(before[d * R + r]) * rb)
                        o = int(stage_off[r][d]) * rb
                        dst.copy_(stage_buf[(w, r)][o: o + rows * rb], non_blocking=True)
            _join_lanes(pb, pb.side)
            with torch.cuda.stream(pb.side):
                if r == R - 1:
                    ev["p1"].record()  # the last payload byte of this rank has been copied out
                pb.hdl.barrier(channel=2 + r)  # staged round r has landed everywhere
                landed[r].record()

    if early_payload and not PAYLOAD_AFTER_SORT:
        send_payload()
    ev["k0"].record()
    for w, rb, src in wires[:nkey]:
        split(w, rb, src, 0, 0)  # the keys of every round go first
    pb.hdl.barrier(channel=1)  # all keys have landed everywhere
    ev["k1"].record()
    ph.mark("split keys")
    if not PAYLOAD_AFTER_SORT and not early_payload:
        send_payload()  # payload rows travel beside the key sort
    marks = {"merge": [], "reduce": []}
    order = []
    for r in range(R):  # "merge": the received keys of a round become a sorted run
        a = torch.cuda.Event(enable_timing=True)
        a.record()
        order.append(_argsort_nulls_last(column(key, r), descending)[:2])
        b = torch.cuda.Event(enable_timing=True)
        b.record()
        marks["merge"].append((a, b))
    ph.mark("key sort")
    if PAYLOAD_AFTER_SORT:
        send_payload()
    if consume:
        # clear_input_blocks: every row of the input must have left before the input's memory can
        # make room for the output, so the gathers wait for the whole payload here
        main.wait_stream(pb.side)
        torch.cuda.synchronize()
        wires = c = src = v = None  # (loop variables above still name the last column's tensors)
        block.columns.clear()
        stage_buf.clear()
        del dest, offs, wide
        torch.cuda.empty_cache()
    outs = []
    for r in range(R):  # "reduce": the round's output block, as soon as its payload is here
        if r == R - 1:
            ev["w0"].record()
        main.wait_event(landed[r])
        if r == R - 1:
            ev["w1"].record()
            ph.mark("wait payload")
        a = torch.cuda.Event(enable_timing=True)
        a.record()
        idx, sorted_key = order[r]
        cols = collections.OrderedDict()
        for name in names:
            cols[name] = sorted_key if name == key else _gather(column(name, r), idx)
        outs.append(DeviceArrowBlock(cols, n_round[r]))
        b = torch.cuda.Event(enable_timing=True)
        b.record()
        marks["reduce"].append((a, b))
    main.wait_stream(pb.side)
    marks["map"] = [(ev_start, ev["w1"])]  # first key byte sent .. last payload byte landed (this stream)
    _last_exchange.clear()
    _last_exchange.update(ev)
    _last_exchange.update(
        bytes_out=remote_rows * sum(rbs.values()),
        bytes_in=(sum(n_round) - own) * sum(rbs.values()),
        key_bytes_out=remote_rows * key_rb)
    if EXCHANGE_LOG is not None:
        EXCHANGE_LOG.append(dict(_last_exchange))
    ph.mark("gather payload")
    ph.finish()
    return outs, marks


def reserve_peer_buffer(nbytes: int, group=None) -> None:
    """Allocate (collectively) the symmetric receive buffer up front, e.g. before a run
    whose HBM budget leaves no room for the grow-by-25% policy of the on-demand path."""
    key = id(group) if group is not None else 0
    pb = _peer_cache.get(key)
    if pb is None or pb.nbytes < nbytes:
        _peer_cache.pop(key, None)
        del pb
        torch.cuda.empty_cache()
        _peer_cache[key] = _PeerBuffer(int(nbytes), group)


PUSH_ROUNDS = int(_os.environ.get("DS_PUSH_ROUNDS", "2"))  # rounds of the push-based plan


PUSH_STAGED = int(_os.environ.get("DS_PUSH_STAGED", "-1"))  # staged rounds (-1: all but the first)


def _require_columns(block, fn: str) -> None:
    """The ranks do not exchange schemas: a block without columns (DeviceArrowBlock.empty()) would
    send this rank into different collectives than the others and hang the job.  Fail here, on the
    rank that holds it, with the way out."""
    if block.num_columns == 0:
        raise ValueError(
            f"{fn}: this rank's block has no columns.  Every rank must pass a block with the dataset's "
            "columns; a rank without rows passes zero-row columns, e.g. "
            "DeviceArrowBlock.from_arrow(schema.empty_table()).")


def sort_block_push(block, key: str, descending: bool = False, group=None, consume: bool = False,
                    rounds: Optional[int] = None, staged: Optional[int] = None):
    """Dataset.sort across the ranks of ``group`` with the push-based plan
    (python/ray/data/_internal/push_based_shuffle.py:363-526): this rank's input block in,
    (this rank's output blocks in key order, {"map", "merge", "reduce": [BlockMetadata]}) out --
    the return shape of PushBasedShufflePlan.execute (:513-517).  Global order = rank order,
    then block order.  See _sort_block_split for what a round is."""
    from .block import BlockExecStats, BlockMetadata, DeviceArrowBlock

    world = dist.get_world_size(group)
    if world > 1:
        _require_columns(block, "sort_block_push")
    var_names = _fixed_width_only(block, "sort_block_push", key, payload_ok=True)
    ph = _Phases()
    ph.mark("start")
    in_rows, in_bytes = block.num_rows, block.nbytes
    R = PUSH_ROUNDS if rounds is None else rounds
    ok = world > 1 and dist.get_backend(group) == "nccl" and not var_names and _split_supported(block, key, world)
    if staged is None:
        staged = PUSH_STAGED if PUSH_STAGED >= 0 else R - 1
    res = _sort_block_split(block, key, descending, group, ph, consume, R, staged) if ok else None
    if res is None:  # inputs the split plan does not take: one round through sort_block
        out = sort_block(block, key, descending, group, consume, rounds=1)
        st = BlockExecStats.builder().build()
        meta = BlockMetadata(num_rows=out.num_rows, size_bytes=out.nbytes, schema=None, input_files=None, exec_stats=st)
        return [out], {"map": [meta], "merge": [meta], "reduce": [meta]}
    outs, marks = res

    def meta_of(rows, nbytes, pair):
        st = BlockExecStats()
        st.cpu_time_s = 0.0
        st._events = pair  # device time of the phase, resolved when read
        return BlockMetadata(num_rows=rows, size_bytes=nbytes, schema=None, input_files=None, exec_stats=st)

    stats = {
        "map": [meta_of(in_rows, in_bytes, marks["map"][0])],
        "merge": [meta_of(o.num_rows, o.columns[key].nbytes, p) for o, p in zip(outs, marks["merge"])],
        "reduce": [meta_of(o.num_rows, o.nbytes, p) for o, p in zip(outs, marks["reduce"])],
    }
    return outs, stats


def sort_block(block, key: str, descending: bool = False, group=None, consume: bool = False,
               rounds: Optional[int] = None):
    """Dataset.sort across the ranks of ``group``: this rank's input block in, this rank's
    slice of the globally sorted dataset out (rank order = key order).

    Every rank passes a block with the dataset's columns -- a rank without rows passes a block
    whose columns are empty, not a block without columns (the ranks do not exchange schemas).

    ``consume=True`` is ``clear_input_blocks`` (python/ray/data/_internal/shuffle.py:95-101):
    the input block's columns are dropped as soon as the partitioned copy exists and the send
    buffers as soon as they have been shipped, which is what lets 62.5 GB per GPU fit."""
    from .block import (DeviceArrowBlock, DeviceArrowBlockAccessor, DeviceColumn,
                        _argsort_nulls_last, _gather)

    world = dist.get_world_size(group)
    if world > 1:
        _require_columns(block, "sort_block")
    var_names = _fixed_width_only(block, "sort_block", key, payload_ok=True, key_ok=True)
    key_is_var = key in var_names
    rank = dist.get_rank(group)
    dev = _dev()
    ph = _Phases()
    ph.mark("start")
    if rounds is None:  # the plan selection point of the reference (sort.py:142-146)
        from .context import DatasetContext

        rounds = PUSH_ROUNDS if DatasetContext.get_current().use_push_based_shuffle else 1
    if consume:  # the caller's handle stays valid (and emptied); work on our own view of it
        src_block = block
        from .block import DeviceArrowBlock as _DAB
        block = _DAB(collections.OrderedDict(src_block.columns), src_block.num_rows)
        src_block.columns.clear()
    if (_EXCHANGE == "split" and world > 1 and dist.get_backend(group) == "nccl" and not var_names
            and _split_supported(block, key, world)):
        res = _sort_block_split(block, key, descending, group, ph, consume, rounds)
        if res is not None:
            outs, _ = res
            return outs[0] if len(outs) == 1 else DeviceArrowBlock.concat(outs)
    block, with_valid = _agree_on_validity(block, group, dev)
    names = block.column_names
    types = {nm: block.columns[nm].type for nm in names}
    if key_is_var:
        boundaries = _varlen_splitters(block, key, world, group)
    else:
        width, ktype = _key_width(block, key, group)
        local = _sample_local(block, key)
        if local.shape[1] != width:
            local = np.zeros((0, width), dtype=np.uint8)
        samples = gather_samples(local, width, group, dev)
        boundaries = splitters_from_samples(samples, ktype, world)
    if boundaries and boundaries[0] is None:
        return block  # no non-null key anywhere: nothing to order
    if descending:
        boundaries.reverse()

    acc = DeviceArrowBlockAccessor(block)
    sort_key = [(key, "descending" if descending else "ascending")]
    ph.mark("sample+splitters")
    if _EXCHANGE == "fused" and 1 < world <= 16 and dist.get_backend(group) == "nccl" and not var_names:
        out = _sort_block_fused(block, key, descending, boundaries, with_valid, group, ph, consume)
        if out is not None:
            return out
    col_rb = {nm: block.columns[nm].row_bytes for nm in names if nm not in var_names}
    parts = acc.partition_unsorted(boundaries, sort_key, descending)
    counts = [p.num_rows for p in parts]
    if consume and world > 1:
        block.columns.clear()  # the partitioned copy replaces the input
        del acc
    ph.mark("partition")
    matrix = exchange_counts(counts, group, dev)
    wires = _wire_columns(parts, [nm for nm in names if nm not in var_names], with_valid,
                          None if key_is_var else key, dev, col_rb)
    var_send = {}
    for nm in var_names:  # lengths ride with the fixed-width wires, the bytes follow below
        from .block import DeviceVarColumn, _concat_varlen

        pcs = [p.columns[nm] for p in parts if p.num_rows]
        if not pcs:  # this rank holds no rows
            whole = DeviceVarColumn(types[nm], torch.zeros(1, dtype=torch.int64, device=dev),
                                    torch.empty(0, dtype=torch.uint8, device=dev),
                                    torch.empty(0, dtype=torch.uint8, device=dev) if nm in with_valid else None)
        elif len(pcs) == 1:
            whole = pcs[0]
        else:
            whole = _concat_varlen(pcs, torch.cat([c.valid for c in pcs]) if nm in with_valid else None)
        len_wire, data, byte_counts = _varlen_send(whole, np.concatenate([[0], np.cumsum(counts)]), dev)
        wires.append((nm + _LEN, 8, len_wire))
        if nm in with_valid:
            wires.append((nm + _VALID, 1, whole.valid.contiguous()))
        var_send[nm] = (data, byte_counts)
    del parts
    if consume:
        torch.cuda.empty_cache()
    # wires that make up the key column (a variable-length key is rebuilt after the exchange)
    nkey = 0 if key_is_var else (2 if key in with_valid else 1)
    n_recv = int(matrix[:, rank].sum())
    rbs = {w: rb for w, rb, _ in wires}

    use_p2p = _EXCHANGE in ("p2p", "fused", "split") and world > 1 and dist.get_backend(group) == "nccl"
    if use_p2p:
        try:
            layout = _PeerLayout([w for w, _, _ in wires], rbs, matrix)
            pb = _peer_buffer(layout.total, group)
        except Exception as e:  # symmetric memory unavailable on this system: NCCL path
            import warnings

            warnings.warn(f"peer-memory exchange unavailable ({e!r}); using NCCL all_to_all")
            use_p2p = False
    ph.mark("counts")
    recv = {}

    def column(name):
        v = recv.get(name + _VALID)
        return DeviceColumn(types[name], recv[name], v)

    if use_p2p:
        main = torch.cuda.current_stream()
        pb.hdl.barrier(channel=0)  # every rank is done with the buffer's previous contents
        for w, rb, buf in wires[:nkey]:
            _p2p_send_column(pb, layout, w, rb, buf, matrix, rank, main)
        _join_lanes(pb, main)
        pb.hdl.barrier(channel=1)  # all keys have landed everywhere
        for w, rb, buf in wires[nkey:]:  # payload copies run beside the key sort
            _p2p_send_column(pb, layout, w, rb, buf, matrix, rank, main)
        ph.mark("p2p keys")
        for w, rb, _ in wires:
            recv[w] = pb.buf[layout.off[w]: layout.off[w] + n_recv * rb]
        if not key_is_var:
            idx, sorted_key, _ = _argsort_nulls_last(column(key), descending)
        ph.mark("key sort")
        _join_lanes(pb, main)
        pb.hdl.barrier(channel=2)  # all payload rows have landed everywhere
        ph.mark("wait payload")
        if consume:  # the send buffers have been shipped
            wires = [(w, rb, None) for w, rb, _ in wires]
            torch.cuda.synchronize()
            torch.cuda.empty_cache()
    else:
        # keys first; sort them while the payload columns are still being exchanged
        for w, rb, buf in wires[:nkey]:
            recv[w], _ = alltoallv_rows(buf, rb, matrix, group)
        ph.mark("a2a keys")
        pending = []
        for w, rb, buf in wires[nkey:]:
            r, work = alltoallv_rows(buf, rb, matrix, group, async_op=True)
            recv[w] = r
            pending.append(work)
        if not key_is_var:
            idx, sorted_key, _ = _argsort_nulls_last(column(key), descending)
        ph.mark("key sort")
        for work in pending:
            work.wait()
        ph.mark("wait payload")
    var_cols = {}
    for name in names:  # the bytes of the variable-length columns follow their lengths
        if name in var_names:
            data, byte_counts = var_send[name]
            v = recv.get(name + _VALID)
            var_cols[name] = _varlen_recv(types[name], recv[name + _LEN], n_recv, data, byte_counts,
                                          v[:n_recv] if v is not None else None, group, dev)
    if key_is_var:
        idx, sorted_key, _ = _argsort_nulls_last(var_cols[key], descending)
    cols = collections.OrderedDict()
    for name in names:
        if name == key:
            cols[name] = sorted_key
        elif name in var_names:
            cols[name] = _gather(var_cols[name], idx)
        else:
            cols[name] = _gather(column(name), idx)
    ph.mark("gather payload")
    ph.finish()
    return DeviceArrowBlock(cols, n_recv)


def _spanning(parts, names, dev):
    """One send buffer per column holding the partitions back to back."""
    base = _contiguous_base(parts, names)
    if base is not None:
        return base
    out = {}
    for name in names:
        cols = [p.columns[name].data for p in parts if p.num_rows]
        out[name] = torch.cat(cols) if cols else torch.empty(0, dtype=torch.uint8, device=dev)
    return out


def _contiguous_base(parts, names):
    """If the partitions are back-to-back views of one storage per column (what
    partition_unsorted returns), return {name: the spanning tensor} without copying."""
    live = [p for p in parts if p.num_rows]
    if not live:
        return None
    out = {}
    for name in names:
        ts = [p.columns[name].data for p in live]
        start = ts[0].data_ptr()
        total = 0
        for t in ts:
            if t.data_ptr() != start + total or t.untyped_storage().data_ptr() != ts[0].untyped_storage().data_ptr():
                return None
            total += t.numel()
        off = ts[0].storage_offset()
        out[name] = torch.as_strided(ts[0], (total,), (1,), off)
    return out


def random_shuffle_block(block, seed: Optional[int], group=None):
    """Dataset.random_shuffle across ranks, structured as the reference's op
    (python/ray/data/_internal/shuffle_and_partition.py:29-97): permute locally with
    seed+rank, cut into W slices of ceil(n/W) rows, rotate which slice goes where with a
    RandomState(seed+rank) list shuffle, exchange, permute the received rows with seed."""
    import math

    from .block import DeviceArrowBlock, DeviceArrowBlockAccessor, DeviceColumn

    world = dist.get_world_size(group)
    if world > 1:
        _require_columns(block, "random_shuffle_block")
    var_names = _fixed_width_only(block, "random_shuffle_block", None, payload_ok=True)
    rank = dist.get_rank(group)
    dev = _dev()
    seed_i = seed + rank if seed is not None else None
    block, with_valid = _agree_on_validity(block, group, dev)
    shuffled = DeviceArrowBlockAccessor(block).random_shuffle(seed_i) if block.num_rows else block
    n = shuffled.num_rows
    slice_sz = max(1, math.ceil(n / world))
    slices = [(min(i * slice_sz, n), min((i + 1) * slice_sz, n)) for i in range(world)]
    order = list(range(world))
    np.random.RandomState(seed_i).shuffle(order)  # slice order[d] goes to rank d
    counts = [slices[order[d]][1] - slices[order[d]][0] for d in range(world)]
    matrix = exchange_counts(counts, group, dev)
    cols = collections.OrderedDict()
    wires = []
    var_send = {}
    for name, c in shuffled.columns.items():
        if name in var_names:
            from .block import _concat_varlen

            whole = _concat_varlen([c.with_valid(None).slice(*slices[order[d]]) for d in range(world)], None)
            len_wire, data, byte_counts = _varlen_send(whole, np.concatenate([[0], np.cumsum(counts)]), dev)
            wires.append((name + _LEN, 8, len_wire))
            var_send[name] = (data, byte_counts)
        else:
            rb = c.row_bytes
            wires.append((name, rb, torch.cat([c.data[slices[order[d]][0] * rb: slices[order[d]][1] * rb] for d in range(world)])))
        if name in with_valid:
            wires.append((name + _VALID, 1, torch.cat([c.valid[slices[order[d]][0]: slices[order[d]][1]] for d in range(world)])))
    rbs = {w: rb for w, rb, _ in wires}
    got = _p2p_exchange_all([w for w, _, _ in wires], rbs, {w: b for w, _, b in wires}, matrix, group)
    recv = {}
    for w, rb, buf in wires:
        recv[w] = got[w] if got is not None else alltoallv_rows(buf, rb, matrix, group)[0]
    n_recv = int(matrix[:, rank].sum())
    for name, c in shuffled.columns.items():
        v = recv.get(name + _VALID)
        if name in var_names:
            data, byte_counts = var_send[name]
            cols[name] = _varlen_recv(c.type, recv[name + _LEN], n_recv, data, byte_counts,
                                      v[:n_recv].clone() if v is not None else None, group, dev)
        else:
            cols[name] = DeviceColumn(c.type, recv[name], v)
    out = DeviceArrowBlock(cols, n_recv)
    # every reducer permutes with the same seed in the reference (:86-89); the
    # subsequence keeps the streams of different ranks apart without changing that contract
    return DeviceArrowBlockAccessor(out).random_shuffle(seed) if out.num_rows else out


def groupby_block(block, key: str, aggs, group=None):
    """groupby(key).aggregate(*aggs) across ranks (python/ray/data/grouped_dataset.py:24-68):
    local sort + combine, range-partition the partial rows, exchange, merge + finalize."""
    from .block import DeviceArrowBlock, DeviceArrowBlockAccessor, DeviceColumn, PartialAggBlock, _searchsorted

    world = dist.get_world_size(group)
    if world > 1:
        _require_columns(block, "groupby_block")
    _fixed_width_only(block, "groupby_block", key)
    rank = dist.get_rank(group)
    dev = _dev()
    width, ktype = _key_width(block, key, group)
    samples = gather_samples(_sample_local(block, key), width, group, dev)
    boundaries = splitters_from_samples(samples, ktype, world)
    acc = DeviceArrowBlockAccessor(block)
    if boundaries and boundaries[0] is None:
        return DeviceArrowBlock.empty()
    sorted_blk = acc.sort_and_partition([], [(key, "ascending")], False)[0] if block.num_rows else block
    partial = DeviceArrowBlockAccessor(sorted_blk).combine(key, aggs)
    if partial.num_rows:
        bounds = _searchsorted(partial.columns[key], boundaries, False)
    else:
        bounds = [0] * (world - 1)
    edges = [0] + bounds + [partial.num_rows]
    counts = [edges[i + 1] - edges[i] for i in range(world)]
    matrix = exchange_counts(counts, group, dev)
    names = _partial_names(key, aggs, block)
    cols = collections.OrderedDict()
    from .block import _row_bytes

    for name, t in names:
        c = partial.columns.get(name)
        if c is not None:
            if c.type != t:  # the layout every rank derives must be the one this rank produced
                raise AssertionError(f"partial column {name!r} is {c.type}, expected {t}")
            send = c.data[: partial.num_rows * c.row_bytes]
        else:
            send = torch.empty(0, dtype=torch.uint8, device=dev)
        rb = _row_bytes(t)
        recv, _ = alltoallv_rows(send.contiguous(), rb, matrix, group)
        cols[name] = DeviceColumn(t, recv)
    n_recv = int(matrix[:, rank].sum())
    # received chunks are key-sorted per source and concatenated in rank order: hand them
    # to the merge as W blocks so that it sorts them stably by key
    chunks, off = [], 0
    for src in range(world):
        r = int(matrix[src, rank])
        if r:
            chunks.append(PartialAggBlock(collections.OrderedDict(
                (nm, cc.slice(off, off + r)) for nm, cc in cols.items()), r))
        off += r
    out, _ = DeviceArrowBlockAccessor.aggregate_combined_blocks(chunks, key, aggs, finalize=True)
    return out


def _partial_names(key, aggs, block) -> List[Tuple[str, pa.DataType]]:
    """Column layout of a combined block, derived from the schema alone so that ranks whose
    partial block is empty put the same buffers on the wire: run counts (count, nvalid, nrows)
    are int64, Std's mean and m2 are float64, and the sum / min / max / absmax parts have the
    64-bit type the reductions compute in (int64 or float64; narrower value columns are
    widened by block._numeric64)."""
    from .block import _stand_in_type, resolve_agg_names

    def reduced_type(t: pa.DataType) -> pa.DataType:
        code, stand_in = _stand_in_type(t)
        return stand_in if code is not None else t

    out = [(key, block.column(key).type)]
    for name, agg in zip(resolve_agg_names(aggs), aggs):
        for part in agg._device_merge_ops():
            if part in ("count", "nvalid", "nrows"):
                t = pa.int64()
            elif part in ("mean", "m2"):
                t = pa.float64()
            else:
                t = reduced_type(block.column(agg._key_fn).type)
            out.append((f"{name}#{part}", t))
    return out
