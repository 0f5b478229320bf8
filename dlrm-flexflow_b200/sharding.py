"""Shard maps of the embedding path (host logic, integer work — parity is bit-exact).

Table-wise placement is the reference's: table i whole on GPU i % G
(/root/reference/examples/cpp/DLRM/strategies/dlrm_strategy.cc:252-256).  The data-parallel
batch split is Legion's blocked restriction, ceil(B/G) samples per rank
(/root/reference/src/runtime/model.cc:876-895).  Row-wise sharding does not exist in the
reference (src/ops/embedding.cu:84-85); it is defined here with the same blocked extents on the
row dimension.
"""
from dataclasses import dataclass
from typing import List, Tuple


def table_owner(table: int, world: int) -> int:
    return table % world


def tables_of(rank: int, num_tables: int, world: int) -> List[int]:
    return [t for t in range(num_tables) if table_owner(t, world) == rank]


def batch_block(batch: int, world: int, rank: int) -> Tuple[int, int]:
    """[lo, hi) of the samples rank owns in the data-parallel part of the model."""
    ext = (batch + world - 1) // world
    lo = min(rank * ext, batch)
    hi = min(lo + ext, batch)
    return lo, hi


def rows_per_shard(rows: int, world: int) -> int:
    return (rows + world - 1) // world


def row_owner(row: int, rows: int, world: int) -> int:
    return row // rows_per_shard(rows, world)


def row_local(row: int, rows: int, world: int) -> int:
    return row % rows_per_shard(rows, world)


def shard_rows(rows: int, world: int, rank: int) -> int:
    """Number of rows of a row-wise sharded table that live on `rank`."""
    rps = rows_per_shard(rows, world)
    return max(0, min(rows, (rank + 1) * rps) - rank * rps)


@dataclass
class TableWisePlan:
    """Who owns what, and the all-to-all split sizes (in floats) of one exchange.

    Forward: rank g holds pooled [B][D] for each owned table and sends, to every rank h, the
    batch block of h of each owned table, table-major: [ntables(g)][B_h][D].  Rank h receives
    from g that block and scatters table t's slice into column block t of its interaction
    input [B_h][T*D] (the Concat layout, /root/reference/src/ops/concat.cu:211-251).
    Backward is the transpose.
    """
    num_tables: int
    world: int
    batch: int
    dim: int

    def owned(self, rank):
        return tables_of(rank, self.num_tables, self.world)

    def send_counts(self, rank):
        n = len(self.owned(rank))
        return [n * (batch_block(self.batch, self.world, h)[1] - batch_block(self.batch, self.world, h)[0])
                * self.dim for h in range(self.world)]

    def recv_counts(self, rank):
        lo, hi = batch_block(self.batch, self.world, rank)
        return [len(self.owned(g)) * (hi - lo) * self.dim for g in range(self.world)]
