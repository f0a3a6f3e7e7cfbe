"""Process grid and partitions (reference: MPI_Dims_create/Cart_* grid, engine.cpp:40-63)."""
from dataclasses import dataclass

from ..ops import cpu


@dataclass
class Grid:
    """rows = data shards, cols = query shards; rank = row * cols + col."""
    rows: int
    cols: int

    @classmethod
    def for_world(cls, world, rows=0, cols=0):
        if rows > 0 and cols > 0:
            if rows * cols != world:
                raise ValueError(f"grid {rows}x{cols} does not match world size {world}")
            return cls(rows, cols)
        return cls(world, 1)  # data-shard everything: smallest operand traffic when Q*A << N*A

    @classmethod
    def like_reference(cls, world):
        return cls(*cpu.choose_grid(world))

    def coords(self, rank):
        return rank // self.cols, rank % self.cols

    def rank_of(self, row, col):
        return row * self.cols + col

    def column_ranks(self, col):
        return [self.rank_of(r, col) for r in range(self.rows)]

    def row_ranks(self, row):
        return [self.rank_of(row, c) for c in range(self.cols)]


def even_slice(total, parts, p):
    """Equal-size slices (last ones may be short or empty): [begin, end)."""
    size = -(-total // parts) if parts > 0 else total
    b = min(p * size, total)
    return b, min(b + size, total), size


def reference_block(total, parts, p):
    """The reference's partition: remainder to part 0 (engine.cpp:62-63)."""
    return cpu.block_range(total, parts, p)
