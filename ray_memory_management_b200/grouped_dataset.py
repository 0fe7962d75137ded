"""groupby().aggregate() on device blocks.

Mirrors python/ray/data/grouped_dataset.py:24-68 (the op) and :118-154 (the driver
lines of GroupedDataset.aggregate) of the reference: range partition by sampled
boundaries, per-partition combine on the map side, merge + finalize on the reduce side.
"""
from typing import List, Optional, Tuple, Union

import pyarrow as pa

from . import sort
from .aggregate import AggregateFn, Count, Max, Mean, Min, Std, Sum
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
        for j in range(1, len(edges)):  # cuts never go backwards, whatever the splitters are
            edges[j] = max(edges[j], edges[j - 1])
        return [result.slice(edges[j], edges[j + 1]) for j in range(output_num_blocks)], map_meta


def _group_ends(block, key: Optional[str]) -> List[int]:
    """End offsets of the runs of equal keys of a key-sorted block (get_key_boundaries,
    grouped_dataset.py:266-275), found on the device for every key type."""
    from .block import _segment_heads

    n = block.num_rows
    if key is None:
        return [n]
    col = block.column(key)
    if col.valid is not None and not bool(col.valid.all()):
        raise NotImplementedError("null groupby keys are not supported on device yet")
    starts, nseg = _segment_heads(col)
    return starts[1:nseg + 1].cpu().tolist()


def _batch_to_arrow(batch) -> pa.Table:
    """What BlockBuilder.add_block accepts from a map_groups UDF, as an Arrow table."""
    import pandas as pd

    if isinstance(batch, pa.Table):
        return batch
    if isinstance(batch, pd.DataFrame):
        return pa.Table.from_pandas(batch, preserve_index=False)
    raise TypeError(  # BlockAccessor.for_block's error for a non-block (block.py:359-375)
        "The map_groups UDF must return a pandas.DataFrame or a pyarrow.Table on the device "
        f"path, got: {type(batch)}"
    )


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
        builder = self._dataset._stats.child_builder("aggregate")
        out, stage_info = aggregate_impl(self._dataset._blocks.copy(), False, self._key, aggs)
        if not stage_info:  # empty input: nothing ran
            return Dataset(out, self._dataset._stats)
        return Dataset(out, builder.build_multistage(stage_info))

    def map_groups(self, fn, *, compute=None, batch_format: str = "native", **ray_remote_args):
        """python/ray/data/grouped_dataset.py:183-310: sort by the key (equal keys never
        straddle output blocks, because the range partition sends equal keys to one
        partition), cut every sorted block at its group boundaries and apply ``fn`` to each
        group in key order; the outputs are concatenated per block.

        The sort and the group boundaries are computed on the device; ``fn`` is a host UDF
        over pandas / pyarrow batches in the reference too, so each sorted block is exported
        once and the result imported once.  ``batch_format``: "native"/"pandas" hand ``fn`` a
        ``pandas.DataFrame`` (the reference promotes Arrow to pandas for "native"), "pyarrow" a
        ``pyarrow.Table``.  ``compute`` / ``ray_remote_args`` are accepted and unused."""
        from .dataset import Dataset

        if batch_format not in ("native", "pandas", "pyarrow"):
            raise ValueError(f"The batch format must be one of 'native', 'pandas', or 'pyarrow', got: {batch_format}")
        if self._key is not None:
            sorted_ds = self._dataset.sort(self._key)
        else:
            sorted_ds = self._dataset.repartition(1)
        builder = sorted_ds._stats.child_builder("map_batches")  # the reference's stage name
        out_blocks = []
        for block in sorted_ds.get_blocks():
            n = BlockAccessor.for_block(block).num_rows()
            if n == 0:
                continue
            ends = _group_ends(block, self._key)
            table = block.to_arrow()
            pieces, start = [], 0
            for end in ends:
                group = table.slice(start, end - start)
                applied = fn(group if batch_format == "pyarrow" else group.to_pandas())
                pieces.append(_batch_to_arrow(applied))
                start = end
            pieces = [p for p in pieces if p.num_columns > 0]
            if pieces:
                out_blocks.append(pa.concat_tables(pieces))
        out = BlockList.from_blocks(out_blocks)
        return Dataset(out, builder.build(out))

    def _aggregate_on(self, agg_cls: type, on, ignore_nulls: bool, *args, **kwargs):
        """grouped_dataset.py:163-181: ``on`` may be a column, a list of columns, or None for
        every column except the key."""
        aggs = self._dataset._build_multicolumn_aggs(
            agg_cls, on, ignore_nulls, *args, skip_cols=self._key, **kwargs
        )
        return self.aggregate(*aggs)

    def count(self):
        return self.aggregate(Count())

    def sum(self, on=None, ignore_nulls: bool = True):
        return self._aggregate_on(Sum, on, ignore_nulls)

    def min(self, on=None, ignore_nulls: bool = True):
        return self._aggregate_on(Min, on, ignore_nulls)

    def max(self, on=None, ignore_nulls: bool = True):
        return self._aggregate_on(Max, on, ignore_nulls)

    def mean(self, on=None, ignore_nulls: bool = True):
        return self._aggregate_on(Mean, on, ignore_nulls)

    def std(self, on=None, ddof: int = 1, ignore_nulls: bool = True):
        return self._aggregate_on(Std, on, ignore_nulls, ddof=ddof)
