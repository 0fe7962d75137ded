"""Per-stage execution stats behind ``Dataset.stats()``.

The reporting surface is the reference's (python/ray/data/_internal/stats.py): a chain of
``DatasetStats`` nodes, one per executed stage, built through ``DatasetStatsBuilder``;
all-to-all stages carry one substage per phase named ``<stage>_<phase>`` (``sort_map``,
``sort_reduce``, ...); ``summary_string()`` prints the text that the reference's tests
canonicalise, and this module's output is compared with the reference's byte for byte on the
same metadata (tests/test_host_callers.py).  The implementation is this repository's own: the
report is assembled from small line producers driven by one table of block statistics.

What the numbers mean here: "Remote wall time" of a block is the DEVICE time of the launches
that produced it (a CUDA-event pair around the map or reduce function of that block, see
``BlockExecStats``), "Remote cpu time" the host CPU time spent issuing them, "Tasks per node"
counts blocks per CUDA device.  There is no stats actor: nothing on this path is lazy.
"""
import time
from typing import Callable, Dict, Iterable, Iterator, List, Optional, Sequence, Set, Tuple, Union

from .block import BlockList, BlockMetadata

# (upper bound in seconds, multiplier, suffix): the first row whose bound exceeds the value wins
_TIME_UNITS: Tuple[Tuple[float, float, str], ...] = ((0.001, 1e6, "us"), (1.0, 1e3, "ms"))


def fmt(seconds: float) -> str:
    """2-decimal duration with the unit the reference prints (us up to 1 ms, ms up to 1 s)."""
    for bound, mult, suffix in _TIME_UNITS:
        if seconds <= bound:
            return f"{round(seconds * mult, 2)}{suffix}"
    return f"{round(seconds, 2)}s"


def _mean(values: Sequence[float]) -> float:
    return float(sum(values)) / len(values)


def _phase_names(stage_name: str, phases: Iterable[str]) -> Iterator[str]:
    """``a->b`` fused stage names keep their full name on the first phase only."""
    tail = stage_name.rsplit("->", 1)[-1]
    for i, phase in enumerate(phases):
        yield f"{stage_name if i == 0 else tail}_{phase}"


class DatasetStatsBuilder:
    """Times one stage on the host from construction to ``build*``."""

    def __init__(self, stage_name: str, parent: Optional["DatasetStats"]):
        self.stage_name = stage_name
        self.parent = parent
        self.start_time = time.perf_counter()

    def _finish(self, stages: Dict[str, List[BlockMetadata]], base_name: Optional[str]) -> "DatasetStats":
        node = DatasetStats(stages=stages, parent=self.parent, base_name=base_name)
        node.time_total_s = time.perf_counter() - self.start_time
        return node

    def build_multistage(self, stages: Dict[str, List[BlockMetadata]]) -> "DatasetStats":
        named = dict(zip(_phase_names(self.stage_name, stages), stages.values()))
        return self._finish(named, self.stage_name)

    def build(self, final_blocks: BlockList) -> "DatasetStats":
        return self._finish({self.stage_name: final_blocks.get_metadata()}, None)


class DatasetStats:
    _next_uuid = 0

    def __init__(self, *, stages: Dict[str, List[BlockMetadata]],
                 parent: Union[Optional["DatasetStats"], List["DatasetStats"]],
                 base_name: Optional[str] = None):
        self.stages = stages
        if parent is not None and not isinstance(parent, list):
            parent = [parent]
        self.parents: Optional[List["DatasetStats"]] = parent
        self.number = 1 + max(p.number for p in self.parents) if self.parents else 0
        self.base_name = base_name
        DatasetStats._next_uuid += 1
        self.dataset_uuid = f"ds{DatasetStats._next_uuid}"
        self.time_total_s = 0.0

    def child_builder(self, name: str) -> DatasetStatsBuilder:
        return DatasetStatsBuilder(name, self)

    # ------------------------------------------------------------------ report
    def summary_string(self, already_printed: Optional[Set[str]] = None) -> str:
        seen = set() if already_printed is None else already_printed
        chunks = []
        for p in self.parents or ():
            text = p.summary_string(seen)
            if text:
                chunks.append(text + "\n")
        chunks.extend(self._own_lines(seen))
        return "".join(chunks)

    def _elapsed(self):
        r = round(self.time_total_s, 2)
        return r if r > 0 else 0  # folds -0.0 and negative clock noise

    def _first_time(self, name: str, seen: Set[str]) -> bool:
        tag = self.dataset_uuid + name
        if tag in seen:
            return False
        seen.add(tag)
        return True

    def _own_lines(self, seen: Set[str]) -> Iterator[str]:
        if not self.stages:
            return
        if len(self.stages) == 1:
            (name, meta), = self.stages.items()
            yield f"Stage {self.number} {name}: "
            if self._first_time(name, seen):
                yield _head_line(meta, self._elapsed())
                yield from _block_lines(meta, "")
            else:
                yield "[execution cached]\n"
            return
        yield f"Stage {self.number} {self.base_name}: executed in {self._elapsed()}s\n"
        for i, (name, meta) in enumerate(self.stages.items()):
            yield f"\n\tSubstage {i} {name}: "
            if self._first_time(name, seen):
                yield f"{sum(m.exec_stats is not None for m in meta)}/{len(meta)} blocks executed\n"
                yield from _block_lines(meta, "\t")
            else:
                yield "\t[execution cached]\n"


def _head_line(meta: List[BlockMetadata], elapsed) -> str:
    """``a/b blocks executed in Ts[, c/b blocks split from parent]`` of a plain stage."""
    total = len(meta)
    ran = sum(m.exec_stats is not None for m in meta)
    parts = []
    if ran:
        parts.append(f"{ran}/{total} blocks executed in {elapsed}s")
    if ran < total:
        parts.append(f"{total - ran}/{total} blocks split from parent" + ("" if ran else f" in {elapsed}s"))
    return ", ".join(parts) + "\n"


# One row per statistic: (label, values of the blocks, how to print a number, print the total?)
_Row = Tuple[str, List, Callable[[float], str], bool]


def _rows(meta: List[BlockMetadata]) -> Iterator[_Row]:
    ex = [m.exec_stats for m in meta if m.exec_stats is not None]
    if ex:
        yield "Remote wall time", [e.wall_time_s for e in ex], fmt, True
        yield "Remote cpu time", [e.cpu_time_s for e in ex], fmt, True
        yield ("Peak heap memory usage (MiB)", [round(e.max_rss_bytes / (1024 * 1024), 2) for e in ex],
               str, False)
    for label, attr in (("Output num rows", "num_rows"), ("Output size bytes", "size_bytes")):
        vals = [getattr(m, attr) for m in meta if getattr(m, attr) is not None]
        if vals:
            yield label, vals, str, True


def _block_lines(meta: List[BlockMetadata], indent: str) -> Iterator[str]:
    for label, vals, show, with_total in _rows(meta):
        mean = _mean(vals)
        cells = [f"{show(min(vals))} min", f"{show(max(vals))} max",
                 f"{show(mean) if show is fmt else int(mean)} mean"]
        if with_total:
            cells.append(f"{show(sum(vals))} total")
        yield f"{indent}* {label}: {', '.join(cells)}\n"
    per_node: Dict[str, int] = {}
    for m in meta:
        if m.exec_stats is not None:
            per_node[m.exec_stats.node_id] = per_node.get(m.exec_stats.node_id, 0) + 1
    if per_node:
        c = list(per_node.values())
        yield (f"{indent}* Tasks per node: {min(c)} min, {max(c)} max, {int(_mean(c))} mean; "
               f"{len(c)} nodes used\n")
