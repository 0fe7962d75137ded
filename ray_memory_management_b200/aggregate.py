"""Aggregations for groupby().aggregate() on device blocks.

Mirrors python/ray/data/aggregate.py:27-211 of the reference (AggregateFn, Count, Sum,
Min, Max, Mean) for the accumulator style init / accumulate_block / merge / finalize,
with the three phases expressed as device operations:

  accumulate_block -> one segmented reduction per accumulator part over a key-sorted
                      block (ArrowBlockAccessor.combine, arrow_block.py:396-467)
  merge            -> the same reduction over concatenated partials
                      (aggregate_combined_blocks, arrow_block.py:549-565)
  finalize         -> identity, or sum / count for Mean (aggregate.py:209)

Null handling follows python/ray/data/_internal/null_aggregate.py:12-26: with
``ignore_nulls=True`` nulls are skipped and an all-null group gives null; with
``ignore_nulls=False`` any null in the group gives null.  On device the ``has_data`` flag of
the reference's wrapped accumulators becomes two int64 parts per aggregate, ``nvalid`` and
``nrows``, which merge by addition like the values do.
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
        # a bad ``on`` is reported by _validate (TypeError for a non-key type, ValueError for a
        # callable or an unknown column), as in the reference (aggregate.py:93-94)
        self._key_fn = on
        self._ignore_nulls = ignore_nulls
        self._kind = kind
        self.name = f"{kind}({str(on)})"

    def _validate(self, ds) -> None:
        from .block import _validate_key_fn

        _validate_key_fn(ds, self._key_fn)  # aggregate.py:93-94


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


def _null_parts(block, on, starts, nseg, n):
    """{"nvalid", "nrows"}: non-null rows and all rows per run -- the SoA form of the
    reference's has_data bookkeeping (null_aggregate.py:29-65,191-225).  Without a validity
    mask both are the run lengths (one kernel, one shared column)."""
    from .block import _segment_counts, _segmented_count_valid

    nrows = _segment_counts(starts, nseg)
    col = block.column(on)
    nvalid = nrows if col.valid is None else _segmented_count_valid(col, None, starts, nseg, n)
    return {"nvalid": nvalid, "nrows": nrows}


def _with_validity(value, parts, ignore_nulls):
    """Attach the group validity of the null policy (null_aggregate.py:12-26,228-251)."""
    from .block import DeviceColumn, _finalize_validity

    if parts["nvalid"] is parts["nrows"]:
        return value  # no nulls anywhere: every group has a value
    return DeviceColumn(value.type, value.data, _finalize_validity(parts["nvalid"], parts["nrows"], ignore_nulls))


class _Reduce(_AggregateOnKeyBase):
    def _device_accumulate(self, block, starts, nseg, n):
        from .block import _segmented_reduce

        parts = {self._kind: _segmented_reduce(block.column(self._key_fn), None, starts, nseg, n, self._kind)}
        parts.update(_null_parts(block, self._key_fn, starts, nseg, n))
        return parts

    def _device_merge_ops(self):
        return {self._kind: "max" if self._kind == "absmax" else self._kind, "nvalid": "sum", "nrows": "sum"}

    def _device_finalize(self, parts):
        return _with_validity(parts[self._kind], parts, self._ignore_nulls)


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


class AbsMax(_Reduce):
    """aggregate.py:289-305 -- max of |x|."""

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True):
        super().__init__(on, ignore_nulls, "absmax")
        self.name = f"abs_max({str(on)})"


class Mean(_AggregateOnKeyBase):
    """aggregate.py:178-211 -- accumulator [sum, count of non-null rows], finalize sum / count."""

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True):
        super().__init__(on, ignore_nulls, "mean")

    def _device_accumulate(self, block, starts, nseg, n):
        from .block import _segmented_reduce

        parts = {"sum": _segmented_reduce(block.column(self._key_fn), None, starts, nseg, n, "sum")}
        parts.update(_null_parts(block, self._key_fn, starts, nseg, n))
        return parts

    def _device_merge_ops(self):
        return {"sum": "sum", "nvalid": "sum", "nrows": "sum"}

    def _device_finalize(self, parts):
        from .block import _finalize_mean

        return _with_validity(_finalize_mean(parts["sum"], parts["nvalid"]), parts, self._ignore_nulls)


class Std(_AggregateOnKeyBase):
    """aggregate.py:214-286 -- accumulator [M2, mean, count].

    Per block: count, sum -> mean, then M2 = sum (x - mean)^2 in a second segmented pass
    (sum_of_squared_diffs_from_mean, arrow_block.py:331-351).  Merging k partials of a key
    evaluates the parallel-variance formula (aggregate.py:236-249) against the merged mean
    in one pass, M2 = sum_i M2_i + n_i (mean_i - mean)^2, which equals the chain of pairwise
    merges up to rounding."""

    def __init__(self, on: Optional[str] = None, ddof: int = 1, ignore_nulls: bool = True):
        super().__init__(on, ignore_nulls, "std")
        self._ddof = ddof

    def _device_accumulate(self, block, starts, nseg, n):
        from .block import _finalize_mean, _segmented_reduce, _segmented_sumsqdiff

        col = block.column(self._key_fn)
        parts = _null_parts(block, self._key_fn, starts, nseg, n)
        total = _segmented_reduce(col, None, starts, nseg, n, "sum")
        mean = _finalize_mean(total, parts["nvalid"])
        parts["mean"] = mean
        parts["m2"] = _segmented_sumsqdiff(col, None, starts, nseg, n, mean)
        return parts

    def _device_merge_ops(self):
        # layout of the partial block only (the merge itself is _device_merge)
        return {"nvalid": "sum", "nrows": "sum", "mean": "sum", "m2": "sum"}

    def _device_merge(self, part, idx, starts, nseg, n):
        from .block import _finalize_mean, _scale, _segmented_reduce, _segmented_sumsqdiff

        nvalid = _segmented_reduce(part("nvalid"), idx, starts, nseg, n, "sum")
        nrows = _segmented_reduce(part("nrows"), idx, starts, nseg, n, "sum")
        total = _segmented_reduce(_scale(part("mean"), part("nvalid")), idx, starts, nseg, n, "sum")
        mean = _finalize_mean(total, nvalid)
        m2 = _segmented_sumsqdiff(part("mean"), idx, starts, nseg, n, mean,
                                  weights=part("nvalid"), addend=part("m2"))
        return {"nvalid": nvalid, "nrows": nrows, "mean": mean, "m2": m2}

    def _device_finalize(self, parts):
        from .block import _finalize_std

        return _with_validity(_finalize_std(parts["m2"], parts["nvalid"], self._ddof), parts, self._ignore_nulls)
