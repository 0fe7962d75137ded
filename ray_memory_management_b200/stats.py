"""Per-stage execution stats with the reference's reporting surface.

``DatasetStats`` / ``DatasetStatsBuilder`` follow python/ray/data/_internal/stats.py:39-84
(builder, substage naming ``<stage>_map`` / ``<stage>_reduce``) and :226-380 (the summary
text that ``Dataset.stats()`` returns and the reference's test_stats.py canonicalises).
What differs by construction: "Remote wall/cpu time" are host times of the stream-ordered
launches of one block's map or reduce function (kernels run asynchronously; ``executed in``
is the host time of the whole stage including the final synchronisation when the caller
asked for it), "Tasks per node" counts blocks per CUDA device, and the "Peak heap memory"
line is the process RSS as in the reference.  No stats actor: nothing here is lazy.
"""
import collections
import time
from typing import Dict, List, Optional, Set, Union

import numpy as np

from .block import BlockList, BlockMetadata


def fmt(seconds: float) -> str:
    if seconds > 1:
        return str(round(seconds, 2)) + "s"
    elif seconds > 0.001:
        return str(round(seconds * 1000, 2)) + "ms"
    else:
        return str(round(seconds * 1000 * 1000, 2)) + "us"


class DatasetStatsBuilder:
    def __init__(self, stage_name: str, parent: Optional["DatasetStats"]):
        self.stage_name = stage_name
        self.parent = parent
        self.start_time = time.perf_counter()

    def build_multistage(self, stages: Dict[str, List[BlockMetadata]]) -> "DatasetStats":
        stage_infos = {}
        for i, (k, v) in enumerate(stages.items()):
            if i == 0:
                stage_infos[self.stage_name + "_" + k] = v
            else:
                stage_infos[self.stage_name.split("->")[-1] + "_" + k] = v
        stats = DatasetStats(stages=stage_infos, parent=self.parent, base_name=self.stage_name)
        stats.time_total_s = time.perf_counter() - self.start_time
        return stats

    def build(self, final_blocks: BlockList) -> "DatasetStats":
        stats = DatasetStats(stages={self.stage_name: final_blocks.get_metadata()}, parent=self.parent)
        stats.time_total_s = time.perf_counter() - self.start_time
        return stats


class DatasetStats:
    _next_uuid = 0

    def __init__(self, *, stages: Dict[str, List[BlockMetadata]],
                 parent: Union[Optional["DatasetStats"], List["DatasetStats"]],
                 base_name: Optional[str] = None):
        self.stages = stages
        if parent is not None and not isinstance(parent, list):
            parent = [parent]
        self.parents: Optional[List["DatasetStats"]] = parent
        self.number = 0 if not self.parents else max(p.number for p in self.parents) + 1
        self.base_name = base_name
        DatasetStats._next_uuid += 1
        self.dataset_uuid = f"ds{DatasetStats._next_uuid}"
        self.time_total_s = 0.0

    def _rounded_total(self):
        r = round(self.time_total_s, 2)
        return 0 if r <= 0 else r  # also folds -0.0

    def child_builder(self, name: str) -> DatasetStatsBuilder:
        return DatasetStatsBuilder(name, self)

    # -- python/ray/data/_internal/stats.py:226-282 ---------------------------------------
    def summary_string(self, already_printed: Optional[Set[str]] = None) -> str:
        if already_printed is None:
            already_printed = set()
        out = ""
        if self.parents:
            for p in self.parents:
                parent_sum = p.summary_string(already_printed)
                if parent_sum:
                    out += parent_sum
                    out += "\n"
        rounded_total = self._rounded_total()
        if len(self.stages) == 1:
            stage_name, metadata = next(iter(self.stages.items()))
            stage_uuid = self.dataset_uuid + stage_name
            out += "Stage {} {}: ".format(self.number, stage_name)
            if stage_uuid in already_printed:
                out += "[execution cached]\n"
            else:
                already_printed.add(stage_uuid)
                out += self._summarize_blocks(metadata, is_substage=False)
        elif len(self.stages) > 1:
            out += "Stage {} {}: executed in {}s\n".format(self.number, self.base_name, rounded_total)
            for n, (stage_name, metadata) in enumerate(self.stages.items()):
                stage_uuid = self.dataset_uuid + stage_name
                out += "\n"
                out += "\tSubstage {} {}: ".format(n, stage_name)
                if stage_uuid in already_printed:
                    out += "\t[execution cached]\n"
                else:
                    already_printed.add(stage_uuid)
                    out += self._summarize_blocks(metadata, is_substage=True)
        return out

    # -- stats.py:305-380 --------------------------------------------------------------------
    def _summarize_blocks(self, blocks: List[BlockMetadata], is_substage: bool) -> str:
        exec_stats = [m.exec_stats for m in blocks if m.exec_stats is not None]
        indent = "\t" if is_substage else ""
        rounded_total = self._rounded_total()
        if is_substage:
            out = "{}/{} blocks executed\n".format(len(exec_stats), len(blocks))
        else:
            out = ""
            if exec_stats:
                out = "{}/{} blocks executed in {}s".format(len(exec_stats), len(blocks), rounded_total)
            if len(exec_stats) < len(blocks):
                if exec_stats:
                    out += ", "
                out += "{}/{} blocks split from parent".format(len(blocks) - len(exec_stats), len(blocks))
                if not exec_stats:
                    out += " in {}s".format(rounded_total)
            out += "\n"

        def four(label, values, f=fmt):
            return indent + "* {}: {} min, {} max, {} mean, {} total\n".format(
                label, f(min(values)), f(max(values)), f(np.mean(values)), f(sum(values)))

        if exec_stats:
            out += four("Remote wall time", [e.wall_time_s for e in exec_stats])
            out += four("Remote cpu time", [e.cpu_time_s for e in exec_stats])
            mem = [round(e.max_rss_bytes / (1024 * 1024), 2) for e in exec_stats]
            out += indent + "* Peak heap memory usage (MiB): {} min, {} max, {} mean\n".format(
                min(mem), max(mem), int(np.mean(mem)))
        rows = [m.num_rows for m in blocks if m.num_rows is not None]
        if rows:
            out += indent + "* Output num rows: {} min, {} max, {} mean, {} total\n".format(
                min(rows), max(rows), int(np.mean(rows)), sum(rows))
        sizes = [m.size_bytes for m in blocks if m.size_bytes is not None]
        if sizes:
            out += indent + "* Output size bytes: {} min, {} max, {} mean, {} total\n".format(
                min(sizes), max(sizes), int(np.mean(sizes)), sum(sizes))
        if exec_stats:
            node_counts = collections.defaultdict(int)
            for s in exec_stats:
                node_counts[s.node_id] += 1
            out += indent + "* Tasks per node: {} min, {} max, {} mean; {} nodes used\n".format(
                min(node_counts.values()), max(node_counts.values()),
                int(np.mean(list(node_counts.values()))), len(node_counts))
        return out
