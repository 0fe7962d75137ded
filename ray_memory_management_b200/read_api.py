"""Loaders: files and in-memory tables -> Datasets of device blocks.

Call shapes follow python/ray/data/read_api.py: ``range_table`` (:147-170), ``read_parquet``
(:322-378), ``from_pandas`` (:863-879), ``from_arrow`` (:999-1016).  ``read_arrow_ipc`` has no
counterpart there (the reference reads Arrow IPC only as its object-store wire format); it is
the cheapest on-disk format for this path because the column buffers are copied to HBM as
they lie in the file.  Decoding (Parquet pages, IPC framing) is pyarrow's on the host; only
fixed-width columns can be imported (see ``DeviceArrowBlock.from_arrow``).
"""
import glob
import os
from typing import List, Optional, Sequence, Union

import numpy as np
import pyarrow as pa

from .dataset import Dataset


def _expand(paths: Union[str, Sequence[str]], suffixes: Sequence[str]) -> List[str]:
    if isinstance(paths, (str, os.PathLike)):
        paths = [paths]
    out: List[str] = []
    for p in paths:
        p = os.fspath(p)
        if os.path.isdir(p):
            found = sorted(
                f for f in glob.glob(os.path.join(p, "**", "*"), recursive=True)
                if os.path.isfile(f) and (not suffixes or f.endswith(tuple(suffixes)))
            )
            out.extend(found)
        elif os.path.exists(p):
            out.append(p)
        else:
            raise FileNotFoundError(p)
    if not out:
        raise ValueError(f"no input files under {list(paths)}")
    return out


def _split_rows(table: pa.Table, parts: int) -> List[pa.Table]:
    n = table.num_rows
    edges = [n * i // parts for i in range(parts + 1)]
    return [table.slice(edges[i], edges[i + 1] - edges[i]) for i in range(parts)]


def range_table(n: int, *, parallelism: int = -1) -> Dataset:
    """{"value": 0..n-1} in ``parallelism`` blocks of near-equal length (default 1 block per
    ~50 M rows, at least 1; the reference's default of 200 blocks exists to feed 200 tasks)."""
    if parallelism is None or parallelism < 0:
        parallelism = max(1, -(-n // 50_000_000))
    parallelism = max(1, min(parallelism, max(n, 1)))
    edges = [n * i // parallelism for i in range(parallelism + 1)]
    tables = [pa.table({"value": np.arange(edges[i], edges[i + 1], dtype=np.int64)}) for i in range(parallelism)]
    return Dataset.from_arrow(tables)


def from_arrow(tables: Union[pa.Table, Sequence[pa.Table]]) -> Dataset:
    return Dataset.from_arrow(tables)


def from_pandas(dfs) -> Dataset:
    import pandas as pd

    if isinstance(dfs, pd.DataFrame):
        dfs = [dfs]
    return Dataset.from_arrow([pa.Table.from_pandas(df, preserve_index=False) for df in dfs])


def read_parquet(paths: Union[str, Sequence[str]], *, columns: Optional[List[str]] = None,
                 parallelism: int = -1) -> Dataset:
    """One block per row group (the reference's read tasks are built from the dataset's
    row-group fragments, parquet_datasource.py); ``parallelism`` > 0 regroups the rows into
    that many blocks instead."""
    import pyarrow.parquet as pq

    tables = []
    for path in _expand(paths, (".parquet", ".pq")):
        f = pq.ParquetFile(path)
        for g in range(f.num_row_groups):
            tables.append(f.read_row_group(g, columns=columns))
        if f.num_row_groups == 0:
            tables.append(f.schema_arrow.empty_table() if columns is None else f.schema_arrow.empty_table().select(columns))
    if parallelism and parallelism > 0:
        tables = _split_rows(pa.concat_tables(tables), parallelism)
    return Dataset.from_arrow(tables)


def read_arrow_ipc(paths: Union[str, Sequence[str]], *, columns: Optional[List[str]] = None) -> Dataset:
    """Arrow IPC files (the random-access "file" format, a.k.a. Feather v2) or streams: one
    block per record batch, memory-mapped so that the host copy made on the way to HBM is
    the only one."""
    tables = []
    for path in _expand(paths, (".arrow", ".feather", ".ipc")):
        source = pa.memory_map(path, "r")
        try:
            reader = pa.ipc.open_file(source)
            batches = [reader.get_batch(i) for i in range(reader.num_record_batches)]
            schema = reader.schema
        except pa.ArrowInvalid:
            source.seek(0)
            reader = pa.ipc.open_stream(source)
            batches = list(reader)
            schema = reader.schema
        if not batches:
            tables.append(schema.empty_table())
        for b in batches:
            t = pa.Table.from_batches([b])
            tables.append(t.select(columns) if columns is not None else t)
    return Dataset.from_arrow(tables)


def write_arrow_ipc(ds: Dataset, directory: str) -> List[str]:
    """One IPC file per block, named like the reference's ``write_*`` outputs
    (``{uuid}_{block_idx:06}.ext``, file_based_datasource.py) with a fixed stem."""
    os.makedirs(directory, exist_ok=True)
    out = []
    for i, table in enumerate(ds.to_arrow()):
        path = os.path.join(directory, f"block_{i:06}.arrow")
        with pa.OSFile(path, "wb") as sink, pa.ipc.new_file(sink, table.schema) as w:
            w.write_table(table)
        out.append(path)
    return out
