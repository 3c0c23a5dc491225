"""Host-side partitioning rules of the multi-GPU modes (the C++ loader applies the same rules,
csrc/model.cpp mat_row_shard / mat_col_shard).

* independent streams: stream s runs on rank s % world (weights replicated, no collective);
* tensor parallel (v4): output rows of key / value / receptance / ffn.key / ffn.receptance / head are
  split in equal contiguous slices (the last may be short), input columns of att.output / ffn.value
  in equal slices that must be multiples of the 32-weight block.
"""
from __future__ import annotations

from typing import List, Tuple


def streams_of_rank(n_streams: int, rank: int, world: int) -> List[int]:
    return [s for s in range(n_streams) if s % world == rank]


def row_shard(n_rows: int, rank: int, world: int) -> Tuple[int, int]:
    """(first row, row count) of `rank`."""
    per = (n_rows + world - 1) // world
    r0 = min(n_rows, per * rank)
    return r0, min(per, n_rows - r0)


def col_shard(n_cols: int, rank: int, world: int, block: int = 32) -> Tuple[int, int]:
    if n_cols % (world * block) != 0:
        raise ValueError(f"{n_cols} columns do not split over {world} ranks in blocks of {block}")
    per = n_cols // world
    return per * rank, per


def tp_supported(n_embed: int, ffn_dim: int, world: int) -> bool:
    return n_embed % (32 * world) == 0 and ffn_dim % (32 * world) == 0
