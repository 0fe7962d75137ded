"""B200-native Datasets shuffle/sort hot path (sort, random_shuffle, groupby-aggregate).

Host-side mirror of the reference's operator/block interface; all per-block compute
runs in the sm_100a CUDA library behind include/ds_shuffle.h (no CPU fallback).
"""
from . import _lib  # noqa: F401  (raises if the CUDA library has not been built)
from .aggregate import AbsMax, AggregateFn, Count, Max, Mean, Min, Std, Sum
from .block import (
    BlockAccessor,
    BlockExecStats,
    BlockList,
    BlockMetadata,
    DeviceArrowBlock,
    DeviceArrowBlockAccessor,
    DeviceColumn,
)
from .context import DatasetContext
from .dataset import Dataset
from .dataset_pipeline import DatasetPipeline
from .grouped_dataset import DeviceGroupbyOp, GroupedDataset, aggregate_impl
from .shuffle import DeviceShufflePlan, ShuffleOp
from .shuffle_and_partition import DeviceShufflePartitionOp
from .read_api import from_arrow, from_pandas, range_table, read_arrow_ipc, read_parquet, write_arrow_ipc
from .sort import DeviceSortOp, sample_boundaries, sort_impl
from .stats import DatasetStats

__all__ = [
    "AbsMax", "AggregateFn", "Count", "Max", "Mean", "Min", "Std", "Sum", "BlockAccessor", "BlockExecStats",
    "BlockList", "BlockMetadata", "DeviceArrowBlock", "DeviceArrowBlockAccessor", "DeviceColumn",
    "DatasetContext", "Dataset", "DeviceGroupbyOp", "GroupedDataset", "aggregate_impl",
    "DeviceShufflePlan", "ShuffleOp", "DeviceShufflePartitionOp", "DeviceSortOp",
    "sample_boundaries", "sort_impl", "DatasetPipeline", "DatasetStats", "from_arrow", "from_pandas",
    "range_table", "read_arrow_ipc", "read_parquet", "write_arrow_ipc",
]
