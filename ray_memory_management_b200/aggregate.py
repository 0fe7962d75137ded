"""Aggregations for groupby().aggregate() on device blocks.

Mirrors python/ray/data/aggregate.py:27-211 of the reference (AggregateFn, Count, Sum,
Min, Max, Mean) for the accumulator style init / accumulate_block / merge / finalize,
with the three phases expressed as device operations:

  accumulate_block -> one segmented reduction per accumulator part over a key-sorted
                      block (ArrowBlockAccessor.combine, arrow_block.py:396-467)
  merge            -> the same reduction over concatenated partials
                      (aggregate_combined_blocks, arrow_block.py:549-565)
  finalize         -> identity, or sum / count for Mean (aggregate.py:209)

Null handling (python/ray/data/_internal/null_aggregate.py) is not implemented yet:
device blocks reject columns with nulls at import, so ``ignore_nulls`` has no effect.
"""
from typing import Dict, Optional


class AggregateFn:
    """Base of the device aggregations; ``name`` is the output column name."""

    name: str = None

    def _validate(self, ds) -> None:
        pass

    # {part name: DeviceColumn with one value per run}
    def _device_accumulate(self, block, starts, nseg: int, n: int) -> Dict[str, "DeviceColumn"]:
        raise NotImplementedError

    # {part name: "sum"|"min"|"max"} used to merge partial accumulators
    def _device_merge_ops(self) -> Dict[str, str]:
        raise NotImplementedError

    def _device_finalize(self, parts):
        raise NotImplementedError


class _AggregateOnKeyBase(AggregateFn):
    def __init__(self, on: Optional[str], ignore_nulls: bool, kind: str):
        if on is not None and not isinstance(on, str):
            raise ValueError(
                f"on must be a string when aggregating on Arrow blocks, but got:{type(on)}."
            )
        self._key_fn = on
        self._ignore_nulls = ignore_nulls
        self._kind = kind
        self.name = f"{kind}({str(on)})"

    def _validate(self, ds) -> None:
        if self._key_fn is None:
            raise ValueError("aggregation on a table dataset needs a column name (on=...)")


class Count(AggregateFn):
    """aggregate.py:97-109 -- counts rows of the group, nulls included."""

    def __init__(self):
        self.name = "count()"

    def _device_accumulate(self, block, starts, nseg, n):
        from .block import _segment_counts

        return {"count": _segment_counts(starts, nseg)}

    def _device_merge_ops(self):
        return {"count": "sum"}

    def _device_finalize(self, parts):
        return parts["count"]


class _Reduce(_AggregateOnKeyBase):
    def _device_accumulate(self, block, starts, nseg, n):
        from .block import _segmented_reduce

        return {self._kind: _segmented_reduce(block.column(self._key_fn), None, starts, nseg, n, self._kind)}

    def _device_merge_ops(self):
        return {self._kind: self._kind}

    def _device_finalize(self, parts):
        return parts[self._kind]


class Sum(_Reduce):
    """aggregate.py:112-131."""

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True):
        super().__init__(on, ignore_nulls, "sum")


class Min(_Reduce):
    """aggregate.py:134-153."""

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True):
        super().__init__(on, ignore_nulls, "min")


class Max(_Reduce):
    """aggregate.py:156-175."""

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True):
        super().__init__(on, ignore_nulls, "max")


class Mean(_AggregateOnKeyBase):
    """aggregate.py:178-211 -- accumulator [sum, count], finalize sum / count."""

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True):
        super().__init__(on, ignore_nulls, "mean")

    def _device_accumulate(self, block, starts, nseg, n):
        from .block import _segment_counts, _segmented_reduce

        return {
            "sum": _segmented_reduce(block.column(self._key_fn), None, starts, nseg, n, "sum"),
            "count": _segment_counts(starts, nseg),
        }

    def _device_merge_ops(self):
        return {"sum": "sum", "count": "sum"}

    def _device_finalize(self, parts):
        from .block import _finalize_mean

        return _finalize_mean(parts["sum"], parts["count"])
