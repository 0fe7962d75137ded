"""A minimal Dataset facade over a BlockList of device blocks.

Only the three operations on the hot path are provided, with the reference's call
shapes: Dataset.sort (python/ray/data/dataset.py:1714-1754), Dataset.random_shuffle
(:750-783) and Dataset.groupby (:1321-1353), plus the few inspection helpers the
reference's tests use (count, take, _block_num_rows, iter_rows, to_arrow).
"""
from typing import List, Optional, Sequence

import pyarrow as pa

from .aggregate import Max, Mean, Min, Std, Sum
from .block import BlockAccessor, BlockList, DeviceArrowBlock, _validate_key_fn
from .grouped_dataset import GroupedDataset
from .shuffle_and_partition import DeviceShufflePartitionOp
from .sort import sort_impl
from .stats import DatasetStats


class Dataset:
    def __init__(self, blocks: BlockList, stats: Optional[DatasetStats] = None, epoch: Optional[int] = None):
        self._blocks = blocks
        # the reference's source stage is "read" or "from_arrow_refs"; every later stage
        # hangs its DatasetStats off this one
        self._stats = stats or DatasetStats(stages={"from_blocks": blocks.get_metadata()}, parent=None)
        self._epoch = epoch  # set by repeat(); read by DatasetPipeline.iter_epochs

    # -- construction ----------------------------------------------------------------
    @staticmethod
    def from_arrow(tables: Sequence[pa.Table]) -> "Dataset":
        if isinstance(tables, pa.Table):
            tables = [tables]
        return Dataset(BlockList.from_blocks(list(tables)))

    @staticmethod
    def from_blocks(blocks: Sequence[DeviceArrowBlock]) -> "Dataset":
        return Dataset(BlockList.from_blocks(list(blocks)))

    # -- the hot path ------------------------------------------------------------------
    def sort(self, key: Optional[str] = None, descending: bool = False) -> "Dataset":
        if self._blocks.initial_num_blocks() == 0:
            return self
        if key is None or callable(key):
            raise ValueError("Table datasets need a column name as sort key")
        builder = self._stats.child_builder("sort")
        out, stage_info = sort_impl(self._blocks.copy(), False, key, descending)
        return Dataset(out, builder.build_multistage(stage_info), self._epoch)

    def random_shuffle(self, *, seed: Optional[int] = None, num_blocks: Optional[int] = None) -> "Dataset":
        if self._blocks.initial_num_blocks() == 0:
            return self
        n = num_blocks or self._blocks.initial_num_blocks()
        builder = self._stats.child_builder("random_shuffle")
        op = DeviceShufflePartitionOp(None, random_shuffle=True, random_seed=seed)
        out, stage_info = op.execute(self._blocks.copy(), n, False)
        return Dataset(out, builder.build_multistage(stage_info), self._epoch)

    def repartition(self, num_blocks: int, *, shuffle: bool = True) -> "Dataset":
        """Dataset.repartition(shuffle=True) (python/ray/data/_internal/stage_impl.py:27-69): the
        shuffle op without the random permutations -- every block is cut into ``num_blocks``
        slices and output block j concatenates slice j of every input block."""
        if not shuffle:
            raise NotImplementedError("repartition(shuffle=False) (fast_repartition) is not on the device path")
        if self._blocks.initial_num_blocks() == 0:
            return self
        builder = self._stats.child_builder("repartition")
        op = DeviceShufflePartitionOp(None, random_shuffle=False, random_seed=None)
        out, stage_info = op.execute(self._blocks.copy(), num_blocks, False)
        return Dataset(out, builder.build_multistage(stage_info), self._epoch)

    # -- pipelines: the per-window callers of sort / random_shuffle ----------------------
    def window(self, *, blocks_per_window: Optional[int] = None,
               bytes_per_window: Optional[int] = None) -> "DatasetPipeline":
        """Dataset.window (python/ray/data/dataset.py:3198-3330): one pipeline window per
        ``blocks_per_window`` blocks (default 10) or per ``bytes_per_window`` bytes."""
        from .dataset_pipeline import DatasetPipeline, split_block_list, split_block_list_by_bytes

        if blocks_per_window is not None and bytes_per_window is not None:
            raise ValueError("Only one windowing scheme can be specified.")
        if blocks_per_window is None:
            blocks_per_window = 10
        if bytes_per_window is not None:
            splits = split_block_list_by_bytes(self._blocks, bytes_per_window)
        else:
            splits = split_block_list(self._blocks, blocks_per_window)
        epoch = self._epoch
        return DatasetPipeline([lambda b=b: Dataset(b, epoch=epoch) for b in splits], length=len(splits))

    def repeat(self, times: Optional[int] = None) -> "DatasetPipeline":
        """Dataset.repeat (dataset.py:3112-3196): the same blocks once per epoch; stages added
        to the pipeline (a per-epoch random_shuffle, typically) run once per epoch."""
        from .dataset_pipeline import DatasetPipeline

        if times is not None and times < 1:
            raise ValueError("`times` must be >= 1, got {}".format(times))
        blocks = self._blocks

        class _Epochs:
            def __iter__(self_inner):
                i = 0
                while times is None or i < times:
                    yield (lambda e=i: Dataset(blocks.copy(), epoch=e))
                    i += 1

        return DatasetPipeline(_Epochs(), length=times or float("inf"))

    def groupby(self, key: Optional[str]) -> GroupedDataset:
        """dataset.py:1321-1353: the key is validated here (unknown column, callable key)."""
        if key is not None:
            _validate_key_fn(self, key)
        return GroupedDataset(self, key)

    def aggregate(self, *aggs):
        ret = self.groupby(None).aggregate(*aggs).take(1)
        return ret[0] if ret else None

    # -- global aggregations (python/ray/data/dataset.py:1386-1712) ------------------------
    # ``on``: a column name -> a scalar; a list of names -> a dict with one entry per column;
    # None -> every column.  None is returned for an empty dataset, an all-null column, or any
    # null with ignore_nulls=False.
    def sum(self, on=None, ignore_nulls: bool = True):
        return self._aggregate_result(self._aggregate_on(Sum, on, ignore_nulls))

    def min(self, on=None, ignore_nulls: bool = True):
        return self._aggregate_result(self._aggregate_on(Min, on, ignore_nulls))

    def max(self, on=None, ignore_nulls: bool = True):
        return self._aggregate_result(self._aggregate_on(Max, on, ignore_nulls))

    def mean(self, on=None, ignore_nulls: bool = True):
        return self._aggregate_result(self._aggregate_on(Mean, on, ignore_nulls))

    def std(self, on=None, ddof: int = 1, ignore_nulls: bool = True):
        return self._aggregate_result(self._aggregate_on(Std, on, ignore_nulls, ddof=ddof))

    def _aggregate_on(self, agg_cls: type, on, *args, **kwargs):
        """dataset.py:3585-3596."""
        aggs = self._build_multicolumn_aggs(agg_cls, on, *args, **kwargs)
        return self.aggregate(*aggs)

    def _build_multicolumn_aggs(self, agg_cls: type, on, ignore_nulls: bool, *args,
                                skip_cols: Optional[List[str]] = None, **kwargs):
        """dataset.py:3598-3629: one aggregation per column; ``on=None`` expands to every
        column of the schema except ``skip_cols`` (the groupby key)."""
        if on is None:
            schema = self.schema()
            if schema is not None and len(schema.names) > 0:
                skip = [skip_cols] if isinstance(skip_cols, str) else (skip_cols or [])
                on = [col for col in schema.names if col not in skip]
        if not isinstance(on, list):
            on = [on]
        return [agg_cls(on_, *args, ignore_nulls=ignore_nulls, **kwargs) for on_ in on]

    @staticmethod
    def _aggregate_result(result):
        """dataset.py:3631-3641: a one-column result row is unwrapped to its value."""
        if result is not None and len(result) == 1:
            return list(result.values())[0]
        return result

    # -- inspection --------------------------------------------------------------------
    def num_blocks(self) -> int:
        return self._blocks.initial_num_blocks()

    def _block_num_rows(self) -> List[int]:
        return [BlockAccessor.for_block(b).num_rows() for b in self._blocks.get_blocks()]

    def count(self) -> int:
        return sum(self._block_num_rows())

    def get_blocks(self) -> List[DeviceArrowBlock]:
        return self._blocks.get_blocks()

    def schema(self) -> Optional[pa.Schema]:
        """Schema of the first block that has columns; None for an empty dataset."""
        if self._blocks.is_cleared():
            return None
        for b in self._blocks.get_blocks():
            s = BlockAccessor.for_block(b).schema()
            if len(s.names) > 0:
                return s
        return None

    def to_pandas(self, limit: int = 100000):
        if self.count() > limit:
            raise ValueError(f"The dataset has more than the given limit of {limit} records.")
        return self.to_arrow_concat().to_pandas()

    def take_all(self, limit: int = 100000) -> List[dict]:
        if self.count() > limit:
            raise ValueError(f"The dataset has more than the given limit of {limit} records.")
        return list(self.iter_rows())

    def show(self, limit: int = 20) -> None:
        for row in self.take(limit):
            print(row)

    def to_arrow(self) -> List[pa.Table]:
        return [BlockAccessor.for_block(b).to_arrow() for b in self._blocks.get_blocks()]

    def to_arrow_concat(self) -> pa.Table:
        tabs = [t for t in self.to_arrow() if t.num_columns > 0]
        return pa.concat_tables(tabs) if tabs else pa.Table.from_pydict({})

    def iter_rows(self):
        for t in self.to_arrow():
            for row in t.to_pylist():
                yield row

    def take(self, limit: int = 20) -> List[dict]:
        out = []
        for row in self.iter_rows():
            out.append(row)
            if len(out) >= limit:
                break
        return out

    def stats(self) -> str:
        """The execution summary of this dataset and its ancestors as text
        (Dataset.stats, python/ray/data/dataset.py:3545-3553); substages are named
        ``<stage>_map`` / ``<stage>_reduce`` as in _internal/stats.py:60-75.  The structured
        form is ``self._stats`` (a ``DatasetStats``: ``.stages``, ``.parents``)."""
        return self._stats.summary_string()
