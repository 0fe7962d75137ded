"""ShuffleOp and the device shuffle plan.

``ShuffleOp`` has the reference's surface (python/ray/data/_internal/shuffle.py:9-52):
subclasses provide static ``map(idx, block, output_num_blocks, *map_args)`` returning
``[part_0, ..., part_{N-1}, BlockMetadata]`` and static ``reduce(*reduce_args,
*mapper_outputs, partial_reduce=False)`` returning ``(block, BlockMetadata)``.

``DeviceShufflePlan.execute`` has the signature and return value of
``SimpleShufflePlan.execute`` (shuffle.py:55-132) and of
``PushBasedShufflePlan.execute`` (push_based_shuffle.py:363-526).  Where the reference
submits M map tasks and N reduce tasks and lets the object store move M x N blocks, this
plan runs the map and reduce functions as stream-ordered kernel launches on the
process's GPU with every intermediate block resident in HBM; with more than one rank
(torch.distributed initialised and ``distributed=True``) the M x N exchange is an
all-to-all over NCCL (see distributed.py).
"""
from typing import Any, Dict, List, Optional, Tuple, Union

from . import _nvtx
from .block import Block, BlockAccessor, BlockExecStats, BlockList, BlockMetadata
from .context import DatasetContext


class ShuffleOp:
    def __init__(self, map_args: List[Any] = None, reduce_args: List[Any] = None):
        self._map_args = map_args or []
        self._reduce_args = reduce_args or []
        assert isinstance(self._map_args, list)
        assert isinstance(self._reduce_args, list)

    @staticmethod
    def map(idx: int, block: Block, output_num_blocks: int, *map_args: List[Any]) -> List[Union[BlockMetadata, Block]]:
        raise NotImplementedError

    @staticmethod
    def reduce(*mapper_outputs: List[Block], partial_reduce: bool = False) -> (Block, BlockMetadata):
        raise NotImplementedError


class DeviceShufflePlan(ShuffleOp):
    def execute(
        self,
        input_blocks: BlockList,
        output_num_blocks: int,
        clear_input_blocks: bool,
        *,
        map_ray_remote_args: Optional[Dict[str, Any]] = None,
        reduce_ray_remote_args: Optional[Dict[str, Any]] = None,
    ) -> Tuple[BlockList, Dict[str, List[BlockMetadata]]]:
        # *_ray_remote_args are accepted for signature compatibility; there are no
        # Ray tasks to configure on this path.
        input_blocks_list = input_blocks.get_blocks()
        input_num_blocks = len(input_blocks_list)
        in_blocks_owned_by_consumer = input_blocks._owned_by_consumer

        # All blocks of a single-process plan live in this GPU's HBM, so an op may run its
        # map and reduce functions once over the concatenated input instead of M + N times
        # over small blocks (same output blocks; see each op's ``fused`` for the argument).
        # With the reference's default of ~76 blocks for 10 M rows this is the difference
        # between ~100 ms of launch overhead and one 1 ms sort.
        fused = getattr(self, "fused", None)
        # (an op whose map + reduce over ONE block still does redundant work -- random_shuffle
        # permutes on both sides -- asks for the fused form from one block on)
        min_blocks = 1 if getattr(self, "fuse_single_block", False) else 2
        if fused is not None and DatasetContext.get_current().fuse_device_blocks and input_num_blocks >= min_blocks:
            stats = BlockExecStats.builder()
            with _nvtx.range(f"{type(self).__name__}.fused"):
                res = fused(input_blocks_list, output_num_blocks, *self._map_args)
            if res is not None:
                new_blocks, map_meta = res
                del input_blocks_list
                if clear_input_blocks:
                    input_blocks.clear()
                # one set of launches produced every output block: they share its host times
                fused_stats = stats.build()
                new_metadata = [
                    BlockAccessor.for_block(b).get_metadata(None, exec_stats=fused_stats) for b in new_blocks
                ]
                return (
                    BlockList(new_blocks, new_metadata, owned_by_consumer=in_blocks_owned_by_consumer),
                    {"map": map_meta, "reduce": new_metadata},
                )

        shuffle_map_out = []
        shuffle_map_metadata = []
        for i, block in enumerate(input_blocks_list):
            with _nvtx.range(f"{type(self).__name__}.map[{i}]"):
                refs = self.map(i, block, output_num_blocks, *self._map_args)
            assert len(refs) == 1 + output_num_blocks, (len(refs), output_num_blocks)
            shuffle_map_metadata.append(refs[-1])  # metadata is LAST (sort.py:58)
            shuffle_map_out.append(refs[:-1])
            input_blocks_list[i] = None

        # Release the inputs' HBM as early as the reference releases its object refs
        # (shuffle.py:95-101).
        del input_blocks_list, block
        if clear_input_blocks:
            input_blocks.clear()

        new_blocks, new_metadata = [], []
        for j in range(output_num_blocks):
            column = [shuffle_map_out[i][j] for i in range(input_num_blocks)]
            for i in range(input_num_blocks):
                shuffle_map_out[i][j] = None
            with _nvtx.range(f"{type(self).__name__}.reduce[{j}]"):
                blk, meta = self.reduce(*self._reduce_args, *column)
            del column
            new_blocks.append(blk)
            new_metadata.append(meta)
        del shuffle_map_out

        stats = {"map": shuffle_map_metadata, "reduce": new_metadata}
        return (
            BlockList(new_blocks, new_metadata, owned_by_consumer=in_blocks_owned_by_consumer),
            stats,
        )
