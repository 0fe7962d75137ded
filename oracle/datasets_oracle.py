"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference hot path.

This module restates, with pyarrow + numpy on the host, what the reference's
Datasets sort / random_shuffle / groupby path computes.  It exists to CHECK the
CUDA path (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline and
``--impl reference`` legs).  Nothing under ``ray_memory_management_b200/`` may
import it; the product path has no CPU fallback.

Parity status: PINNED.  ``oracle/gen_golden.py`` runs the reference's own source
files (under ``oracle/ref_shim.py``) in the build container and stores
input/output vectors in ``tests/golden/``; ``tests/test_oracle_golden.py`` checks
every function below against those vectors, and (container only)
``tests/test_oracle_vs_reference.py`` checks it against the live reference.

The arithmetic of the reference lives in third-party libraries that are not under
/root/reference: pyarrow (pinned >=6.0.1,<7.0.0 at python/setup.py:219; 24.0.0
here) and numpy (>=1.19, python/setup.py:209-214; 2.3 here).  The restatement
calls the same library entry points the reference calls (``sort_indices``,
``take``, ``searchsorted``, ``quantile(nearest)``, ``RandomState.permutation``) so
the oracle inherits their documented contracts: stable sort, signed int64 order,
memcmp order for binary, side="left" search.

All file:line citations are relative to /root/reference/python/ray/data/.
"""
import math
import random as _pyrandom
from typing import List, Optional, Sequence, Tuple

import numpy as np
import pyarrow as pa
import pyarrow.compute as pac


# ----------------------------------------------------------------------------
# sort
# ----------------------------------------------------------------------------
def _order(descending: bool) -> str:
    return "descending" if descending else "ascending"


def empty_table() -> pa.Table:
    # _internal/arrow_block.py:119-121 (ArrowBlockBuilder._empty_table)
    return pa.Table.from_pydict({})


def num_rows(table: pa.Table) -> int:
    # _internal/arrow_block.py:226-229
    return table.num_rows if table.num_columns > 0 else 0


def sort_table(table: pa.Table, col: str, descending: bool) -> pa.Table:
    """_internal/arrow_ops/transform_pyarrow.py:12-16: stable sort_indices + take."""
    idx = pac.sort_indices(table, sort_keys=[(col, _order(descending))])
    return table.take(idx)


def _key_values(table: pa.Table, col: str) -> np.ndarray:
    arr = table[col]
    arr = arr.combine_chunks() if arr.num_chunks else pa.array([], type=arr.type)
    return arr.to_numpy(zero_copy_only=False)


def partition_bounds(sorted_keys: np.ndarray, boundaries, descending: bool) -> np.ndarray:
    """_internal/arrow_block.py:373-386: row offsets that cut a sorted key column."""
    if descending:
        n = len(sorted_keys)
        return n - np.searchsorted(
            sorted_keys, boundaries, sorter=np.arange(n - 1, -1, -1)
        )
    return np.searchsorted(sorted_keys, boundaries)


def sort_and_partition(
    table: pa.Table, boundaries: Sequence, col: str, descending: bool
) -> List[pa.Table]:
    """_internal/arrow_block.py:353-394."""
    if table.num_rows == 0:
        return [empty_table() for _ in range(len(boundaries) + 1)]
    table = sort_table(table, col, descending)
    if len(boundaries) == 0:
        return [table]
    bounds = partition_bounds(_key_values(table, col), list(boundaries), descending)
    parts, last = [], 0
    for b in bounds:
        b = int(b)
        parts.append(table.slice(last, b - last).combine_chunks())
        last = b
    parts.append(table.slice(last).combine_chunks())
    return parts


def merge_sorted_blocks(blocks: Sequence[pa.Table], col: str, descending: bool) -> pa.Table:
    """_internal/arrow_block.py:473-486 -> transform_pyarrow.py:70-75.

    Note: a re-sort of the mapper-ordered concatenation, not a k-way merge."""
    blocks = [b for b in blocks if b.num_rows > 0]
    if not blocks:
        return empty_table()
    ret = pa.concat_tables(blocks)
    idx = pac.sort_indices(ret, sort_keys=[(col, _order(descending))])
    return ret.take(idx)


def nearest_rank_indices(num_samples: int, num_reducers: int) -> np.ndarray:
    """Positions np.quantile(sorted, q, method='nearest') selects, q = linspace(0,1,R)[1:].

    _internal/sort.py:110-115.  Derived by calling numpy on arange (so whatever
    tie rule numpy applies at exact .5 virtual indices is inherited)."""
    base = np.arange(num_samples, dtype=np.float64)
    qs = np.linspace(0, 1, num_reducers)
    return np.array(
        [int(np.quantile(base, q, method="nearest")) for q in qs][1:], dtype=np.int64
    )


def boundaries_from_samples(sample_items: np.ndarray, num_reducers: int) -> list:
    """_internal/sort.py:110-115: np.sort then nearest-rank quantiles, first dropped."""
    if len(sample_items) == 0:
        return [None] * (num_reducers - 1)
    s = np.sort(sample_items)
    return [s[i] for i in nearest_rank_indices(len(s), num_reducers)]


def sample_boundaries(
    tables: Sequence[pa.Table], col: str, num_reducers: int, rng=None
) -> list:
    """_internal/sort.py:80-115 with _internal/arrow_block.py:269-272 sampling.

    ``rng``: object with ``sample(range, k)``; the reference uses the process-global
    ``random`` module (unseeded), so pass ``random`` after ``random.seed`` to replay."""
    rng = rng or _pyrandom
    n_samples = int(num_reducers * 10 / len(tables))
    samples = []
    for t in tables:
        if num_rows(t) == 0:
            continue
        k = min(n_samples, t.num_rows)
        idx = rng.sample(range(t.num_rows), k)
        s = t.select([col]).take(idx)
        if s.num_rows > 0:
            samples.append(_key_values(s, col))
    if not samples:
        return [None] * (num_reducers - 1)
    return boundaries_from_samples(np.concatenate(samples), num_reducers)


def simple_shuffle(map_fn, reduce_fn, tables: Sequence[pa.Table], output_num_blocks: int):
    """_internal/shuffle.py:55-132 run in-process: M maps -> N reduces."""
    map_out = [map_fn(i, t, output_num_blocks) for i, t in enumerate(tables)]
    return [
        reduce_fn(*[map_out[i][j] for i in range(len(tables))])
        for j in range(output_num_blocks)
    ]


def sort_dataset(
    tables: Sequence[pa.Table], col: str, descending: bool = False, rng=None
) -> List[pa.Table]:
    """_internal/sort.py:120-154 (simple plan)."""
    if len(tables) == 0:
        return []
    r = len(tables)
    boundaries = sample_boundaries(tables, col, r, rng)
    if descending:
        boundaries.reverse()
    return simple_shuffle(
        lambda i, t, n: sort_and_partition(t, boundaries, col, descending),
        lambda *parts: merge_sorted_blocks(parts, col, descending),
        tables,
        r,
    )


def global_stable_sort(tables: Sequence[pa.Table], col: str, descending: bool) -> pa.Table:
    """The observable contract of Dataset.sort (SURVEY.md 8(c) golden property 1):
    concatenation of the output blocks == stable sort of the concatenated input."""
    return sort_table(pa.concat_tables(list(tables)), col, descending)


# ----------------------------------------------------------------------------
# random_shuffle
# ----------------------------------------------------------------------------
def random_shuffle_block(table: pa.Table, seed: Optional[int]) -> pa.Table:
    """_internal/arrow_block.py:184-186."""
    rs = np.random.RandomState(seed)
    return table.take(rs.permutation(num_rows(table)))


def shuffle_map(idx: int, table: pa.Table, output_num_blocks: int, seed: Optional[int]):
    """_internal/shuffle_and_partition.py:29-70 (block_udf=None, random_shuffle=True)."""
    seed_i = seed + idx if seed is not None else None
    table = random_shuffle_block(table, seed_i)
    slice_sz = max(1, math.ceil(num_rows(table) / output_num_blocks))
    slices = [
        table.slice(i * slice_sz, slice_sz) for i in range(output_num_blocks)
    ]
    np.random.RandomState(seed_i).shuffle(slices)
    return slices


def shuffle_reduce(seed: Optional[int], *parts: pa.Table) -> pa.Table:
    """_internal/shuffle_and_partition.py:73-97."""
    parts = [p for p in parts if p.num_columns > 0]
    t = pa.concat_tables(parts) if parts else empty_table()
    return random_shuffle_block(t, seed)


def random_shuffle_dataset(
    tables: Sequence[pa.Table], seed: Optional[int], num_blocks: Optional[int] = None
) -> List[pa.Table]:
    n = num_blocks or len(tables)
    return simple_shuffle(
        lambda i, t, nb: shuffle_map(i, t, nb, seed),
        lambda *parts: shuffle_reduce(seed, *parts),
        tables,
        n,
    )


def repartition_dataset(tables: Sequence[pa.Table], num_blocks: int) -> List[pa.Table]:
    """Dataset.repartition(shuffle=True): _ShufflePartitionOp with random_shuffle=False
    (_internal/stage_impl.py:27-69, _internal/shuffle_and_partition.py:29-97)."""

    def _map(i, t, nb):
        slice_sz = max(1, math.ceil(num_rows(t) / nb))
        return [t.slice(j * slice_sz, slice_sz) for j in range(nb)]

    def _reduce(*parts):
        parts = [p for p in parts if p.num_columns > 0]
        return pa.concat_tables(parts) if parts else empty_table()

    return simple_shuffle(_map, _reduce, tables, num_blocks)


# ----------------------------------------------------------------------------
# groupby / aggregate.  aggs: sequence of (kind, on) with kind in
# {"count","sum","mean","min","max"}; ``on`` is None for count.
# ----------------------------------------------------------------------------
def agg_name(kind: str, on: Optional[str]) -> str:
    # aggregate.py:108,130,152,174,210
    return "count()" if kind == "count" else f"{kind}({on})"


def resolved_names(aggs) -> List[str]:
    """_internal/arrow_block.py:456-463,469-471: duplicates become name_2, name_3."""
    seen, out = {}, []
    for kind, on in aggs:
        name = agg_name(kind, on)
        c = seen.get(name, 0)
        out.append(name if c == 0 else f"{name}_{c + 1}")
        seen[name] = c + 1
    return out


def _run_starts(sorted_keys: np.ndarray) -> np.ndarray:
    if len(sorted_keys) == 0:
        return np.zeros(0, dtype=np.int64)
    change = np.flatnonzero(sorted_keys[1:] != sorted_keys[:-1]) + 1
    return np.concatenate([[0], change]).astype(np.int64)


def _segment_reduce(kind: str, vals: np.ndarray, starts: np.ndarray, n: int):
    ends = np.concatenate([starts[1:], [n]])
    if kind == "count":
        return (ends - starts).astype(np.int64)
    if len(starts) <= 20000:
        # the reference reduces each group's slice with pyarrow (arrow_block.py:445-449)
        arr = pa.array(vals)
        fn = {"sum": pac.sum, "min": pac.min, "max": pac.max}[kind]
        out = [fn(arr.slice(int(s), int(e - s))).as_py() for s, e in zip(starts, ends)]
        return np.array(out, dtype=vals.dtype)
    uf = {"sum": np.add, "min": np.minimum, "max": np.maximum}[kind]
    return uf.reduceat(vals, starts)


class Partial:
    """SoA form of one combined block: keys + one accumulator tuple per agg.

    The reference stores accumulators as Arrow list columns carrying a trailing
    has_data flag (null_aggregate.py:29-46); without nulls the flag is always 1 and
    only the numeric parts matter, so they are kept as plain arrays here."""

    def __init__(self, keys: np.ndarray, accs: List[Tuple[np.ndarray, ...]]):
        self.keys, self.accs = keys, accs

    def __len__(self):
        return len(self.keys)


def combine(table: pa.Table, key: str, aggs) -> Partial:
    """_internal/arrow_block.py:396-467 on a key-sorted block."""
    n = num_rows(table)
    if n == 0:
        return Partial(np.zeros(0, dtype=np.int64), [() for _ in aggs])
    keys = _key_values(table, key)
    starts = _run_starts(keys)
    accs = []
    for kind, on in aggs:
        vals = _key_values(table, on) if on is not None else None
        if kind == "mean":  # aggregate.py:188-198: [sum, count]
            accs.append(
                (
                    _segment_reduce("sum", vals, starts, n),
                    _segment_reduce("count", vals, starts, n),
                )
            )
        else:
            accs.append((_segment_reduce(kind, vals, starts, n),))
    return Partial(keys[starts], accs)


def aggregate_combined_blocks(parts: Sequence[Partial], key: str, aggs, finalize: bool):
    """_internal/arrow_block.py:488-587: merge partials by key in block order.

    heapq.merge is stable across its inputs, so accumulators of one key are merged
    left-to-right in mapper order; a stable sort of the concatenation by key gives
    the same order."""
    parts = [p for p in parts if len(p) > 0]
    names = resolved_names(aggs)
    if not parts:
        return empty_table() if finalize else Partial(np.zeros(0, np.int64), [() for _ in aggs])
    keys = np.concatenate([p.keys for p in parts])
    order = np.argsort(keys, kind="stable")
    keys = keys[order]
    starts = _run_starts(keys)
    n = len(keys)
    merged = []
    for i, (kind, on) in enumerate(aggs):
        comps = []
        for c in range(len(parts[0].accs[i])):
            v = np.concatenate([p.accs[i][c] for p in parts])[order]
            # merge = a1 + a2 for sum/count/mean parts (aggregate.py:107,119,185-187);
            # min/max merge with min/max (aggregate.py:141,163)
            op = {"min": np.minimum, "max": np.maximum}.get(kind, np.add)
            comps.append(op.reduceat(v, starts) if n else v)
        merged.append(tuple(comps))
    out_keys = keys[starts]
    if not finalize:
        return Partial(out_keys, merged)
    cols = {key: out_keys}
    for name, (kind, on), acc in zip(names, aggs, merged):
        if kind == "mean":  # aggregate.py:209: a[0] / a[1]
            cols[name] = acc[0] / acc[1]
        else:
            cols[name] = acc[0]
    return pa.table(cols)


def groupby_dataset(tables: Sequence[pa.Table], key: str, aggs, rng=None) -> List[pa.Table]:
    """grouped_dataset.py:24-68,118-154 (simple plan)."""
    if len(tables) == 0:
        return []
    r = len(tables)
    boundaries = sample_boundaries(tables, key, r, rng)

    def _map(i, t, n):
        return [combine(p, key, aggs) for p in sort_and_partition(t, boundaries, key, False)]

    def _reduce(*parts):
        return aggregate_combined_blocks(parts, key, aggs, True)

    return simple_shuffle(_map, _reduce, tables, r)


def groupby_global(tables: Sequence[pa.Table], key: str, aggs) -> pa.Table:
    """Observable contract of groupby().aggregate(): one row per distinct key, key
    ascending, aggregates over all rows of the key (SURVEY.md 8(c) property 4)."""
    full = sort_table(pa.concat_tables(list(tables)), key, False)
    return aggregate_combined_blocks([combine(full, key, aggs)], key, aggs, True)


# ----------------------------------------------------------------------------
# GraySort-format synthetic records (experimental/raysort/constants.py:8: 100 B)
# ----------------------------------------------------------------------------
def graysort_table(n: int, seed: int, key_bytes: int = 10, payload_bytes: int = 90) -> pa.Table:
    """{key: fixed_size_binary(10), payload: fixed_size_binary(90)}; payload carries
    the little-endian row id in its first 8 bytes so stability is checkable."""
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 256, size=(n, key_bytes), dtype=np.uint8)
    payload = np.zeros((n, payload_bytes), dtype=np.uint8)
    payload[:, :8] = np.arange(n, dtype="<u8").view(np.uint8).reshape(n, 8)
    payload[:, 8:] = (np.arange(payload_bytes - 8, dtype=np.uint8) + 1)[None, :]
    return pa.table(
        {
            "key": fsb_from_numpy(keys),
            "payload": fsb_from_numpy(payload),
        }
    )


def fsb_from_numpy(a: np.ndarray) -> pa.Array:
    n, w = a.shape
    buf = pa.py_buffer(np.ascontiguousarray(a).tobytes())
    return pa.Array.from_buffers(pa.binary(w), n, [None, buf])


def fsb_to_numpy(arr) -> np.ndarray:
    if isinstance(arr, pa.ChunkedArray):
        arr = arr.combine_chunks()
    w = arr.type.byte_width
    buf = arr.buffers()[1]
    a = np.frombuffer(buf, dtype=np.uint8, count=(arr.offset + len(arr)) * w)
    return a.reshape(-1, w)[arr.offset :]
