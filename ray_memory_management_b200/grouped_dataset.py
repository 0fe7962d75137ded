"""groupby().aggregate() on device blocks.

Mirrors python/ray/data/grouped_dataset.py:24-68 (the op) and :118-154 (the driver
lines of GroupedDataset.aggregate) of the reference: range partition by sampled
boundaries, per-partition combine on the map side, merge + finalize on the reduce side.
"""
from typing import List, Optional, Tuple, Union

from . import sort
from .aggregate import AbsMax, AggregateFn, Count, Max, Mean, Min, Std, Sum
from .block import Block, BlockAccessor, BlockExecStats, BlockList, BlockMetadata
from .context import DatasetContext
from .shuffle import DeviceShufflePlan, ShuffleOp


class _GroupbyOp(ShuffleOp):
    @staticmethod
    def map(idx: int, block: Block, output_num_blocks: int, boundaries: List,
            key: Optional[str], aggs: Tuple[AggregateFn]) -> List[Union[BlockMetadata, Block]]:
        """Partition the block and combine rows with the same key."""
        stats = BlockExecStats.builder()
        acc = BlockAccessor.for_block(block)
        if key is None:
            partitions = [acc.to_block()]
        else:
            partitions = acc.sort_and_partition(
                boundaries, [(key, "ascending")] if isinstance(key, str) else key, descending=False
            )
        parts = [BlockAccessor.for_block(p).combine(key, aggs) for p in partitions]
        meta = acc.get_metadata(input_files=None, exec_stats=stats.build())
        return parts + [meta]

    @staticmethod
    def reduce(key: Optional[str], aggs: Tuple[AggregateFn], *mapper_outputs: List[Block],
               partial_reduce: bool = False) -> (Block, BlockMetadata):
        """Aggregate sorted and partially combined blocks."""
        return BlockAccessor.for_block(mapper_outputs[0]).aggregate_combined_blocks(
            list(mapper_outputs), key, aggs, finalize=not partial_reduce
        )


class DeviceGroupbyOp(_GroupbyOp, DeviceShufflePlan):
    @staticmethod
    def fused(blocks: List[Block], output_num_blocks: int, boundaries: List, key: Optional[str],
              aggs: Tuple[AggregateFn]):
        """One sort + one segmented reduction over the concatenated blocks, cut at the
        splitters: output block j of the map/reduce plan holds the finalized groups whose key
        falls into range j, which is the j-th slice of the key-sorted global result.  (Float
        sums associate differently than per-block partial sums would; both are within the
        path's 1e-6 tolerance.)"""
        from . import _lib
        from .block import DeviceArrowBlock, DeviceArrowBlockAccessor, _searchsorted

        stats = BlockExecStats.builder()
        accs = [BlockAccessor.for_block(b) for b in blocks]
        if sum(a.num_rows() for a in accs) > _lib.DS_MAX_ROWS:
            return None
        map_meta = [a.get_metadata(input_files=None, exec_stats=stats.build()) for a in accs]
        live = [a.to_block() for a in accs if a.num_rows() > 0]
        if not live:
            return [DeviceArrowBlock.empty() for _ in range(output_num_blocks)], map_meta
        whole = DeviceArrowBlockAccessor(DeviceArrowBlock.concat(live))
        if key is not None:
            whole = DeviceArrowBlockAccessor(whole.sort_and_partition([], [(key, "ascending")], False)[0])
        partial = whole.combine(key, aggs)
        result, _ = DeviceArrowBlockAccessor.aggregate_combined_blocks([partial], key, aggs, finalize=True)
        if key is None or output_num_blocks == 1:
            return [result], map_meta
        bounds = _searchsorted(result.columns[key], boundaries, False)
        edges = [0] + bounds + [result.num_rows]
        return [result.slice(edges[j], edges[j + 1]) for j in range(output_num_blocks)], map_meta


def aggregate_impl(blocks: BlockList, clear_input_blocks: bool, key: Optional[str],
                   aggs: Tuple[AggregateFn]):
    """do_agg of the reference (grouped_dataset.py:118-154)."""
    stage_info = {}
    if len(aggs) == 0:
        raise ValueError("Aggregate requires at least one aggregation")
    if blocks.initial_num_blocks() == 0:
        return blocks, stage_info
    num_mappers = blocks.initial_num_blocks()
    num_reducers = num_mappers
    if key is None:
        num_reducers = 1
        boundaries = []
    else:
        boundaries = sort.sample_boundaries(
            blocks.get_blocks(), [(key, "ascending")] if isinstance(key, str) else key, num_reducers
        )
    DatasetContext.get_current()  # plan selection point (grouped_dataset.py:142-146)
    op = DeviceGroupbyOp(map_args=[boundaries, key, aggs], reduce_args=[key, aggs])
    return op.execute(blocks, num_reducers, clear_input_blocks)


class GroupedDataset:
    """python/ray/data/grouped_dataset.py:71-161."""

    def __init__(self, dataset, key: Optional[str]):
        self._dataset = dataset
        self._key = key

    def aggregate(self, *aggs: AggregateFn):
        from .dataset import Dataset

        for agg in aggs:
            agg._validate(self._dataset)
        out, stats = aggregate_impl(self._dataset._blocks.copy(), False, self._key, aggs)
        return Dataset(out, {"aggregate": stats})

    def count(self):
        return self.aggregate(Count())

    def sum(self, on: str, ignore_nulls: bool = True):
        return self.aggregate(Sum(on, ignore_nulls))

    def min(self, on: str, ignore_nulls: bool = True):
        return self.aggregate(Min(on, ignore_nulls))

    def max(self, on: str, ignore_nulls: bool = True):
        return self.aggregate(Max(on, ignore_nulls))

    def mean(self, on: str, ignore_nulls: bool = True):
        return self.aggregate(Mean(on, ignore_nulls))

    def std(self, on: str, ddof: int = 1, ignore_nulls: bool = True):
        return self.aggregate(Std(on, ddof, ignore_nulls))
