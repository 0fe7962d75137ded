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
    # one block: the map side's and the reduce side's permutation compose to one permutation
    fuse_single_block = True

    @staticmethod
    def fused(blocks: List[Block], output_num_blocks: int, block_udf, random_shuffle: bool,
              random_seed: Optional[int]):
        """random_shuffle over blocks that share a GPU: ONE permutation of the concatenated
        rows, cut into the output block sizes the map/reduce plan would produce (those depend
        only on the block lengths and the per-mapper RandomState slice shuffles, replayed here
        on the host).  The contract kept is the reference's: a permutation of the rows,
        deterministic per seed.  Repartition (no permutation) and UDF-fused shuffles keep the
        map/reduce path, whose row placement is part of their result."""
        from . import _lib

        if not random_shuffle or block_udf is not None:
            return None
        stats = BlockExecStats.builder()
        accs = [BlockAccessor.for_block(b) for b in blocks]
        total = sum(a.num_rows() for a in accs)
        if total > _lib.DS_MAX_ROWS:
            return None
        sizes = [0] * output_num_blocks
        for i, a in enumerate(accs):  # shuffle_and_partition.py:56-65 on lengths only
            n = a.num_rows()
            slice_sz = max(1, math.ceil(n / output_num_blocks))
            lens = [max(0, min(n, (j + 1) * slice_sz) - min(n, j * slice_sz)) for j in range(output_num_blocks)]
            seed_i = random_seed + i if random_seed is not None else None
            np.random.RandomState(seed_i).shuffle(lens)
            sizes = [x + y for x, y in zip(sizes, lens)]
        map_meta = [a.get_metadata(input_files=None, exec_stats=stats.build()) for a in accs]
        live = [a.to_block() for a in accs if a.num_rows() > 0]
        if not live:
            return [DeviceArrowBlock.empty() for _ in range(output_num_blocks)], map_meta
        whole = BlockAccessor.for_block(DeviceArrowBlock.concat(live)).random_shuffle(random_seed)
        out, last = [], 0
        for sz in sizes:
            out.append(whole.slice(last, last + sz))
            last += sz
        return out, map_meta
