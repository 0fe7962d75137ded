"""random_shuffle / repartition op on device blocks.

Mirrors python/ray/data/_internal/shuffle_and_partition.py:12-105 of the reference:
map = permute the block with seed+idx, cut into ceil(n/N)-row slices, permute the list
of slices with a fresh RandomState(seed+idx); reduce = concat + permute with seed.
Row permutations come from the Philox kernel (ds_random_permutation), so results are
valid, per-seed-deterministic permutations but not bit-equal to MT19937 (SURVEY.md
8(c) property 3: the reference's own two plans already disagree with each other).
"""
import math
from typing import Callable, Iterable, List, Optional, Union

import numpy as np

from .block import Block, BlockAccessor, BlockExecStats, BlockMetadata, DeviceArrowBlock
from .shuffle import DeviceShufflePlan, ShuffleOp


class _ShufflePartitionOp(ShuffleOp):
    def __init__(self, block_udf=None, random_shuffle: bool = False, random_seed: Optional[int] = None):
        super().__init__(
            map_args=[block_udf, random_shuffle, random_seed],
            reduce_args=[random_shuffle, random_seed],
        )

    @staticmethod
    def map(idx: int, block: Block, output_num_blocks: int,
            block_udf: Optional[Callable[[Block], Iterable[Block]]], random_shuffle: bool,
            random_seed: Optional[int]) -> List[Union[BlockMetadata, Block]]:
        stats = BlockExecStats.builder()
        if block_udf:
            blocks = list(block_udf(block))
            block = DeviceArrowBlock.concat([BlockAccessor.for_block(b).to_block() for b in blocks])
        block = BlockAccessor.for_block(block)

        seed_i = None
        if random_shuffle:
            seed_i = random_seed + idx if random_seed is not None else None
            block = BlockAccessor.for_block(block.random_shuffle(seed_i))

        slice_sz = max(1, math.ceil(block.num_rows() / output_num_blocks))
        slices = [
            block.slice(i * slice_sz, (i + 1) * slice_sz, copy=False)
            for i in range(output_num_blocks)
        ]
        if random_shuffle:
            # order of the slices only: a host-side list shuffle, as in the reference
            np.random.RandomState(seed_i).shuffle(slices)

        num_rows = sum(s.num_rows for s in slices)
        assert num_rows == block.num_rows(), (num_rows, block.num_rows())
        metadata = block.get_metadata(input_files=None, exec_stats=stats.build())
        return slices + [metadata]

    @staticmethod
    def reduce(random_shuffle: bool, random_seed: Optional[int], *mapper_outputs: List[Block],
               partial_reduce: bool = False) -> (Block, BlockMetadata):
        stats = BlockExecStats.builder()
        new_block = DeviceArrowBlock.concat(
            [BlockAccessor.for_block(b).to_block() for b in mapper_outputs]
        )
        accessor = BlockAccessor.for_block(new_block)
        if random_shuffle:
            new_block = accessor.random_shuffle(random_seed if random_seed is not None else None)
            accessor = BlockAccessor.for_block(new_block)
        new_metadata = BlockMetadata(
            num_rows=accessor.num_rows(),
            size_bytes=accessor.size_bytes(),
            schema=accessor.schema(),
            input_files=None,
            exec_stats=stats.build(),
        )
        return new_block, new_metadata


class DeviceShufflePartitionOp(_ShufflePartitionOp, DeviceShufflePlan):
    pass
