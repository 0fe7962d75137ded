"""DatasetPipeline: the per-window callers of the hot path.

The reference's ``DatasetPipeline`` (python/ray/data/dataset_pipeline.py) is a lazy
sequence of Dataset-producing callables plus a list of per-window stages.  Only what is
needed to reach ``Dataset.random_shuffle`` / ``Dataset.sort`` / ``Dataset.repartition`` once
per window is mirrored here:

* ``DatasetPipeline.from_iterable`` (dataset_pipeline.py:1112-1127)
* ``foreach_window`` (:1079-1098) and the three ``*_each_window`` wrappers that call it
  (``repartition_each_window`` :800-807, ``random_shuffle_each_window`` :809-819,
  ``sort_each_window`` :821-826)
* ``iter_datasets`` (:1062-1077), ``iter_epochs`` (:607-680), ``iter_rows`` / ``take`` /
  ``count`` for inspection, and the "cannot be read multiple times" rule.
* ``Dataset.window`` (dataset.py:3198-3330) and ``Dataset.repeat`` (:3112-3196) create one.

There is no pipelined executor behind it: a window's stages run when the consumer pulls
that window, as stream-ordered launches on this process's GPU (the reference overlaps
windows with Ray tasks; one GPU stream gives nothing to overlap with).
"""
import itertools
from typing import Callable, Iterable, Iterator, List, Optional

from .block import BlockList


class DatasetPipeline:
    def __init__(self, base_iterable: Iterable[Callable[[], "Dataset"]], stages=None,
                 length: Optional[int] = None, _executed: Optional[List[bool]] = None):
        self._base_iterable = base_iterable
        self._stages = list(stages or [])
        self._length = length
        # shared between a pipeline and everything derived from it (dataset_pipeline.py:84-92)
        self._executed = _executed or [False]

    # -- construction ------------------------------------------------------------------
    @staticmethod
    def from_iterable(iterable: Iterable[Callable[[], "Dataset"]]) -> "DatasetPipeline":
        length = len(iterable) if hasattr(iterable, "__len__") else None
        return DatasetPipeline(iterable, length=length)

    # -- per-window transforms -----------------------------------------------------------
    def foreach_window(self, fn: Callable[["Dataset"], "Dataset"]) -> "DatasetPipeline":
        if self._executed[0]:
            raise RuntimeError("Pipeline cannot be read multiple times.")
        return DatasetPipeline(self._base_iterable, self._stages + [fn], self._length,
                               _executed=self._executed)

    def repartition_each_window(self, num_blocks: int, *, shuffle: bool = False) -> "DatasetPipeline":
        return self.foreach_window(lambda ds: ds.repartition(num_blocks, shuffle=shuffle))

    def random_shuffle_each_window(self, *, seed: Optional[int] = None,
                                   num_blocks: Optional[int] = None) -> "DatasetPipeline":
        return self.foreach_window(lambda ds: ds.random_shuffle(seed=seed, num_blocks=num_blocks))

    def sort_each_window(self, key: Optional[str] = None, descending: bool = False) -> "DatasetPipeline":
        return self.foreach_window(lambda ds: ds.sort(key, descending))

    # -- re-shaping the window sequence ----------------------------------------------------
    def repeat(self, times: Optional[int] = None) -> "DatasetPipeline":
        """dataset_pipeline.py:442-523: the windows this pipeline produces are computed once
        (epoch 0) and replayed ``times - 1`` more times (forever if None) with increasing
        epoch numbers; stages added after repeat() run on every replay."""
        from .dataset import Dataset

        if self._length == float("inf"):
            raise ValueError("Cannot repeat a pipeline of infinite length.")
        if times is not None and times < 1:
            raise ValueError("`times` must be >= 1, got {}".format(times))
        source = self

        class _Replay:
            def __iter__(self_inner):
                cached = []
                for ds in source.iter_datasets():
                    ds._epoch = 0
                    cached.append(ds)
                    yield (lambda d=ds: d)
                epoch = 1
                while cached and (times is None or epoch < times):
                    for ds in cached:
                        yield (lambda d=ds, e=epoch: Dataset(d._blocks.copy(), d._stats, epoch=e))
                    epoch += 1

        if times is None:
            length = float("inf")
        else:
            length = None if self._length is None else self._length * times
        return DatasetPipeline(_Replay(), length=length)

    def rewindow(self, *, blocks_per_window: int, preserve_epoch: bool = True) -> "DatasetPipeline":
        """dataset_pipeline.py:365-440: re-cut the stream of blocks into windows of
        ``blocks_per_window`` blocks; with ``preserve_epoch`` a window never mixes epochs."""
        from .dataset import Dataset

        source = self

        class _Rewindow:
            def __iter__(self_inner):
                blocks, metas, epoch = [], [], None

                def flush():
                    bl = BlockList(list(blocks), list(metas))
                    blocks.clear()
                    metas.clear()
                    return (lambda b=bl, e=epoch: Dataset(b, epoch=e))

                for ds in source.iter_datasets():
                    if preserve_epoch and blocks and ds._epoch != epoch:
                        yield flush()
                    epoch = ds._epoch
                    for b, m in zip(ds._blocks.get_blocks(), ds._blocks.get_metadata()):
                        blocks.append(b)
                        metas.append(m)
                        if len(blocks) == blocks_per_window:
                            yield flush()
                if blocks:
                    yield flush()

        return DatasetPipeline(_Rewindow(), length=None)

    # -- consumption ---------------------------------------------------------------------
    def iter_datasets(self) -> Iterator["Dataset"]:
        if self._executed[0]:
            raise RuntimeError("Pipeline cannot be read multiple times.")
        self._executed[0] = True

        def gen():
            for make in self._base_iterable:
                ds = make()
                for stage in self._stages:
                    ds = stage(ds)
                yield ds

        return gen()

    def iter_epochs(self, max_epoch: int = -1) -> Iterator["DatasetPipeline"]:
        """Split the windows into sub-pipelines at epoch changes (``Dataset._epoch``, set by
        ``Dataset.repeat``); windows without an epoch all belong to epoch 0."""
        windows = self.iter_datasets()
        state = {"next": None, "done": False}

        def pull():
            if state["next"] is None and not state["done"]:
                try:
                    state["next"] = next(windows)
                except StopIteration:
                    state["done"] = True
            return state["next"]

        def epoch_windows(epoch):
            while True:
                ds = pull()
                if ds is None or (getattr(ds, "_epoch", None) or 0) != epoch:
                    return
                state["next"] = None
                yield (lambda d=ds: d)

        def gen():
            produced = 0
            while pull() is not None and (max_epoch < 0 or produced < max_epoch):
                epoch = getattr(state["next"], "_epoch", None) or 0
                sub = DatasetPipeline(epoch_windows(epoch))
                yield sub
                for _ in epoch_windows(epoch):  # drain what the consumer left unread
                    pass
                produced += 1

        return gen()

    def iter_rows(self):
        for ds in self.iter_datasets():
            yield from ds.iter_rows()

    def take(self, limit: int = 20) -> List[dict]:
        return list(itertools.islice(self.iter_rows(), limit))

    def take_all(self, limit: int = 100000) -> List[dict]:
        out = list(itertools.islice(self.iter_rows(), limit + 1))
        if len(out) > limit:
            raise ValueError(f"The pipeline has more than the given limit of {limit} records.")
        return out

    def count(self) -> int:
        return sum(ds.count() for ds in self.iter_datasets())

    def __repr__(self) -> str:
        return "DatasetPipeline(num_windows={}, num_stages={})".format(self._length, 1 + len(self._stages))


def split_block_list(blocks: BlockList, split_size: int) -> List[BlockList]:
    """BlockList.split(split_size) of the reference (block_list.py:59-76): ceil(len /
    split_size) windows of near-equal length, the way ``np.array_split`` cuts them (7 blocks
    with split_size 3 give windows of 3, 2 and 2 blocks, not 3, 3 and 1)."""
    bl, meta = blocks.get_blocks(), blocks.get_metadata()
    num_splits = -(-len(bl) // split_size)
    out, start = [], 0
    for s in range(num_splits):
        size = len(bl) // num_splits + (1 if s < len(bl) % num_splits else 0)
        out.append(BlockList(bl[start:start + size], meta[start:start + size],
                             owned_by_consumer=blocks._owned_by_consumer))
        start += size
    return out


def split_block_list_by_bytes(blocks: BlockList, bytes_per_split: int) -> List[BlockList]:
    """BlockList.split_by_bytes (block_list.py:78-113): greedy packing in block order; a
    window always holds at least one block, so the bound is soft for oversized blocks."""
    bl, meta = blocks.get_blocks(), blocks.get_metadata()
    out, cur_b, cur_m, cur_size = [], [], [], 0
    for b, m in zip(bl, meta):
        if m.size_bytes is None:
            raise RuntimeError("Block has unknown size, cannot use split_by_bytes()")
        size = m.size_bytes
        if cur_b and cur_size + size > bytes_per_split:
            out.append(BlockList(cur_b, cur_m, owned_by_consumer=blocks._owned_by_consumer))
            cur_b, cur_m, cur_size = [], [], 0
        cur_b.append(b)
        cur_m.append(m)
        cur_size += size
    if cur_b:
        out.append(BlockList(cur_b, cur_m, owned_by_consumer=blocks._owned_by_consumer))
    return out
