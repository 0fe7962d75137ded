"""DatasetContext flags read by the shuffle plans (python/ray/data/context.py:83-199)."""
import threading
from typing import Optional

_default_context: "Optional[DatasetContext]" = None
_context_lock = threading.Lock()


class DatasetContext:
    def __init__(self, use_push_based_shuffle: bool = False, use_device_shuffle: bool = True,
                 pipeline_push_based_shuffle_reduce_tasks: bool = True, use_polars: bool = False,
                 fuse_device_blocks: bool = True):
        # Plan selection happens where the reference reads use_push_based_shuffle
        # (sort.py:142-146, stage_impl.py:101-109, grouped_dataset.py:142-146).  This
        # package only ships the device plan; the flag is kept so that call sites and
        # tests read the same.
        self.use_push_based_shuffle = use_push_based_shuffle
        self.use_device_shuffle = use_device_shuffle
        self.pipeline_push_based_shuffle_reduce_tasks = pipeline_push_based_shuffle_reduce_tasks
        self.use_polars = use_polars
        # run map/reduce once over the concatenated blocks when they share a GPU (shuffle.py)
        self.fuse_device_blocks = fuse_device_blocks

    @staticmethod
    def get_current() -> "DatasetContext":
        global _default_context
        with _context_lock:
            if _default_context is None:
                _default_context = DatasetContext()
            return _default_context

    @staticmethod
    def _set_current(context: "DatasetContext") -> None:
        global _default_context
        _default_context = context
