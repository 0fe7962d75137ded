"""Dataset.sort on device blocks: splitter sampling, the sort op, sort_impl.

Mirrors python/ray/data/_internal/sort.py:41-158 of the reference.  Differences that
do not change the observable result (the concatenated output is the
global stable sort of the concatenated input):
  * the map side only range-partitions (stable), the reduce side sorts once; the
    reference sorts on both sides (sort.py:52-54 and :67-69).
  * samples are drawn by a Philox-stratified kernel instead of random.sample.
"""
from typing import Any, Callable, List, Tuple, TypeVar, Union

import numpy as np
import pyarrow as pa

from .block import (
    Block,
    BlockAccessor,
    BlockExecStats,
    BlockList,
    BlockMetadata,
    DeviceArrowBlock,
)
from .context import DatasetContext
from .shuffle import DeviceShufflePlan, ShuffleOp

T = TypeVar("T")
SortKeyT = Union[None, List[Tuple[str, str]], Callable[[T], Any]]


class _SortOp(ShuffleOp):
    @staticmethod
    def map(idx: int, block: Block, output_num_blocks: int, boundaries: List[T], key: SortKeyT,
            descending: bool) -> List[Union[BlockMetadata, Block]]:
        stats = BlockExecStats.builder()
        acc = BlockAccessor.for_block(block)
        out = acc.partition_unsorted(boundaries, key, descending)
        meta = acc.get_metadata(input_files=None, exec_stats=stats.build())
        return out + [meta]

    @staticmethod
    def reduce(key: SortKeyT, descending: bool, *mapper_outputs: List[Block],
               partial_reduce: bool = False) -> (Block, BlockMetadata):
        return BlockAccessor.for_block(mapper_outputs[0]).merge_sorted_blocks(
            mapper_outputs, key, descending
        )


class DeviceSortOp(_SortOp, DeviceShufflePlan):
    @staticmethod
    def fused(blocks: List[Block], output_num_blocks: int, boundaries: List[T], key: SortKeyT,
              descending: bool):
        """One stable sort of the concatenated blocks, cut at the splitters.

        Output block j of the map/reduce plan is the stable sort of the rows whose key falls
        into range j, in (block, row) order -- i.e. exactly the j-th slice, cut at
        searchsorted(boundaries), of the stable sort of the concatenation.  Returns None when
        the concatenation would exceed one sort call."""
        from . import _lib

        stats = BlockExecStats.builder()
        accs = [BlockAccessor.for_block(b) for b in blocks]
        if sum(a.num_rows() for a in accs) > _lib.DS_MAX_ROWS:
            return None
        live = [a.to_block() for a in accs if a.num_rows() > 0]
        map_meta = [a.get_metadata(input_files=None, exec_stats=stats.build()) for a in accs]
        if not live:
            return [DeviceArrowBlock.empty() for _ in range(output_num_blocks)], map_meta
        from .block import DeviceArrowBlockAccessor

        return DeviceArrowBlockAccessor.sort_blocks_and_partition(live, boundaries, key, descending), map_meta


def nearest_rank_positions(num_samples: int, num_reducers: int) -> List[int]:
    """Positions np.quantile(sorted, q, interpolation="nearest") picks for
    q in linspace(0, 1, R), first dropped (sort.py:110-115): round-half-even of
    q * (n - 1)."""
    qs = np.linspace(0, 1, num_reducers)
    return [int(np.around((num_samples - 1) * q)) for q in qs][1:]


def fsb_order(rows: np.ndarray) -> np.ndarray:
    """Stable memcmp order of n fixed-width byte rows ((n, w) uint8): the rows are compared as
    big-endian integers, 8 bytes at a time (least significant chunk first for lexsort)."""
    n, w = rows.shape
    keys = []
    for b in range(0, w, 8):
        chunk = np.zeros((n, 8), dtype=np.uint8)
        chunk[:, : min(8, w - b)] = rows[:, b : b + 8]
        keys.append(chunk.view(">u8").ravel())
    return np.argsort(keys[0], kind="stable") if len(keys) == 1 else np.lexsort(tuple(reversed(keys)))


def fsb_order_statistics(rows: np.ndarray, picks) -> list:
    """rows[order][i] for i in picks, where order is the memcmp order of the (n, w <= 16) byte
    rows, WITHOUT sorting them: a selection (np.partition) on the leading 8 bytes, then the
    rank inside the group of rows that tie on those (the multi-GPU plan needs W-1 splitters
    out of thousands of samples on every step)."""
    n, w = rows.shape
    if w > 16:
        order = fsb_order(rows)
        return [rows[order[i]].tobytes() for i in picks]
    pad = np.zeros((n, 16), dtype=np.uint8)
    pad[:, :w] = rows
    hi = pad[:, :8].copy().view(">u8").ravel().astype(np.uint64)
    lo = pad[:, 8:].copy().view(">u8").ravel().astype(np.uint64)
    picks = [int(i) for i in picks]
    by_hi = np.partition(hi, sorted(set(picks))) if picks else hi
    out = []
    for i in picks:
        v = by_hi[i]
        tie = lo[hi == v]
        below = int(np.count_nonzero(hi < v))
        l = np.partition(tie, i - below)[i - below] if len(tie) > 1 else tie[0]
        out.append((int(v).to_bytes(8, "big") + int(l).to_bytes(8, "big"))[:w])
    return out


def sort_samples(samples: pa.Array) -> list:
    """Host sort of the few sampled keys (np.sort at sort.py:110); fixed_size_binary keys
    are ordered bytewise (memcmp), as Arrow orders them."""
    n = len(samples)
    t = samples.type
    if pa.types.is_fixed_size_binary(t):
        w = t.byte_width
        raw = np.frombuffer(samples.buffers()[1], dtype=np.uint8)[samples.offset * w : (samples.offset + n) * w]
        rows = raw.reshape(n, w)
        return [r.tobytes() for r in rows[fsb_order(rows)]]
    if pa.types.is_temporal(t):  # splitters of temporal keys travel as their integer values
        samples = samples.cast(pa.int32() if t.bit_width == 32 else pa.int64())
    return list(np.sort(samples.to_numpy(zero_copy_only=False)))


def boundaries_from_samples(samples: pa.Array, num_reducers: int) -> list:
    """Driver-side half of sample_boundaries (sort.py:104-115): sort the <= 10*R sampled
    keys on the host and take nearest-rank quantiles."""
    n = len(samples)
    if n == 0:
        return [None] * (num_reducers - 1)
    srt = sort_samples(samples)
    return [srt[i] for i in nearest_rank_positions(n, num_reducers)]


def sample_boundaries(blocks: List[Block], key: SortKeyT, num_reducers: int) -> List[T]:
    """Return (num_reducers - 1) items in ascending order that split the key domain
    into ranges of about equally many rows (sort.py:80-115)."""
    if isinstance(key, list) and len(key) > 1:
        raise ValueError("Multiple boundary sampling keys not supported.")
    if num_reducers <= 1:
        # linspace(0, 1, 1)[1:] is empty: one output partition has no splitters, whatever the
        # samples are -- so none are drawn (the reference samples 10 rows and discards them)
        return []
    n_samples = int(num_reducers * 10 / len(blocks))
    column = key[0][0] if isinstance(key, list) else key
    if len(blocks) > 1 and all(isinstance(b, DeviceArrowBlock) for b in blocks):
        # device blocks of one GPU: one sampling launch over the concatenated key column
        # (the same n_samples per block on average, stratified over the whole column)
        live = [b.select([column]) for b in blocks if b.num_rows > 0]
        samples = []
        if live:
            whole = BlockAccessor.for_block(DeviceArrowBlock.concat(live))
            samples = [whole.sample(min(n_samples * len(live), whole.num_rows()), key)]
    else:
        samples = [_sample_block(block, n_samples, key) for block in blocks]
    samples = [s for s in samples if s.num_rows > 0]
    if len(samples) == 0:
        return [None] * (num_reducers - 1)
    # one device-side concatenation and ONE device-to-host copy for all blocks' samples
    host = DeviceArrowBlock.concat(samples).to_arrow()[column].combine_chunks().drop_null()
    if isinstance(host, pa.ChunkedArray):
        host = host.combine_chunks()
    if pa.types.is_floating(host.type):
        # a NaN splitter cuts nothing sensibly (NaNs sort after every number in both directions
        # and np.searchsorted on them is order-dependent in the reference): never pick one
        import pyarrow.compute as pc

        host = host.filter(pc.invert(pc.is_nan(host)))
    return boundaries_from_samples(host, num_reducers)


def sort_impl(blocks: BlockList, clear_input_blocks: bool, key: SortKeyT,
              descending: bool = False) -> Tuple[BlockList, dict]:
    """sort.py:120-154."""
    stage_info = {}
    blocks_list = blocks.get_blocks()
    if len(blocks_list) == 0:
        return BlockList([], []), stage_info

    if isinstance(key, str):
        key = [(key, "descending" if descending else "ascending")]
    if isinstance(key, list):
        descending = key[0][1] == "descending"

    num_mappers = len(blocks_list)
    num_reducers = num_mappers  # same number of output partitions (sort.py:134-136)
    boundaries = sample_boundaries(blocks_list, key, num_reducers)
    if descending:
        boundaries.reverse()
    del blocks_list

    DatasetContext.get_current()  # plan selection point (sort.py:142-146): device plan only
    sort_op = DeviceSortOp(map_args=[boundaries, key, descending], reduce_args=[key, descending])
    return sort_op.execute(blocks, num_reducers, clear_input_blocks)


def _sample_block(block: Block, n_samples: int, key: SortKeyT) -> DeviceArrowBlock:
    return BlockAccessor.for_block(block).sample(n_samples, key)
