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


def route_slots(idx, rows: int, world: int, owner: int):
    """Host statement of what ffemb_rowwise_route (include/ffemb.h) leaves in owner `owner`'s slot
    array for one table: idx [B][L] global row ids of the WHOLE batch -> int64 [B][L] with
    row - owner*rows_per_shard where `owner` holds the row and -1 (empty slot) elsewhere.  Slot
    positions are the global (sample, j) positions: nothing is compacted."""
    import numpy as np
    idx = np.asarray(idx, np.int64)
    rps = rows_per_shard(rows, world)
    mine = (idx // rps) == owner
    return np.where(mine, idx - owner * rps, -1).astype(np.int64)


def route_bytes(batch: int, bags, dim: int, world: int, owners_per_bag=None) -> int:
    """NVLink + local bytes ONE rank stores in a route of its ceil(batch/world) samples: 8 bytes per
    lookup and owner for the slots, plus one gradient row of dim floats per (bag, owner) pair —
    `owners_per_bag` distinct owners on average (default: min(L, world), the uniform-index bound)."""
    ext = (batch + world - 1) // world
    total = 0
    for L in bags:
        pairs = min(L, world) if owners_per_bag is None else owners_per_bag
        total += ext * L * 8 * world + int(ext * pairs * dim * 4)
    return total


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


# ---------------------------------------------------------------------------------------------
# Strategy files: the reference's TEXT format, as its loader parses it
# (/root/reference/src/runtime/strategy.cc:95-146) and its writer emits it (:147-189):
#   <number of ops>
#   per op:  <name>  <device_type: 0 GPU | 1 CPU>  <nDims>  <dim[0] .. dim[nDims-1]>
#            <number of device ids = prod(dims)>  <device ids ...>
# whitespace-separated.  dim[0] is the fastest (feature) dimension, dim[nDims-1] the sample
# dimension (/root/reference/src/runtime/model.cc:726-736).  A table placed whole on one GPU is
# "embedding<i>" with a [1,1] grid (examples/cpp/DLRM/strategies/dlrm_strategy.cc:97-117,252-256;
# the data loader keys on the same names, examples/cpp/DLRM/dlrm.cc:489).
# ---------------------------------------------------------------------------------------------
DEVICE_GPU, DEVICE_CPU = 0, 1


@dataclass
class ParallelConfig:
    """/root/reference/include/config.h:47-73."""
    device_type: int
    dims: List[int]
    device_ids: List[int]

    def num_parts(self):
        n = 1
        for d in self.dims:
            n *= d
        return n


def write_strategy(path, strategies):
    """strategies: {op name: ParallelConfig}.  Same token stream as save_strategies_to_file
    (ops in name order, like the std::map it iterates)."""
    with open(path, "w") as f:
        f.write(f"{len(strategies)}\n")
        for name in sorted(strategies):
            c = strategies[name]
            if c.device_type not in (DEVICE_GPU, DEVICE_CPU):
                raise ValueError("Unsupported Device Type")          # strategy.cc:172-174
            if len(c.device_ids) != c.num_parts():
                raise ValueError(f"{name}: {c.num_parts()} parts but {len(c.device_ids)} device ids")
            f.write(f"{name}\n{c.device_type}\n{len(c.dims)}\n")
            f.write("".join(f"{d}\t" for d in c.dims) + "\n")
            f.write(f"{c.num_parts()}\n")
            f.write("".join(f"{d}\t" for d in c.device_ids) + "\n")


def read_strategy(path):
    """-> {op name: ParallelConfig}; the checks of load_strategies_from_file (device type known,
    device id count == prod(dims) or 0, no op twice) raise ValueError here (asserts there)."""
    tok = open(path).read().split()
    pos = 0

    def nxt():
        nonlocal pos
        if pos >= len(tok):
            raise ValueError("strategy file ends early")
        pos += 1
        return tok[pos - 1]

    out = {}
    for _ in range(int(nxt())):
        name, dev, nd = nxt(), int(nxt()), int(nxt())
        if dev not in (DEVICE_GPU, DEVICE_CPU):
            raise ValueError("Unsupported Device Type")
        dims = [int(nxt()) for _ in range(nd)]
        n_ids = int(nxt())
        cfg = ParallelConfig(dev, dims, [])
        if n_ids not in (0, cfg.num_parts()):
            raise ValueError(f"{name}: {n_ids} device ids for {cfg.num_parts()} parts")
        cfg.device_ids = [int(nxt()) for _ in range(n_ids)]
        if name in out:
            raise ValueError(f"{name} listed twice")
        out[name] = cfg
    return out


def tablewise_strategy(owner, world):
    """The DLRM strategy for a table placement: embedding<i> whole on GPU owner[i] ([1,1] grid),
    concat data-parallel over all GPUs (dlrm_strategy.cc:252-262)."""
    s = {f"embedding{i}": ParallelConfig(DEVICE_GPU, [1, 1], [int(g)]) for i, g in enumerate(owner)}
    s["concat"] = ParallelConfig(DEVICE_GPU, [1, world], list(range(world)))
    return s


def owners_from_strategy(strategies, num_tables):
    """Table owners out of a strategy: every embedding<i> must be a [1,1] grid on a GPU — the
    operator forbids anything else (/root/reference/src/ops/embedding.cu:84-85)."""
    own = []
    for i in range(num_tables):
        c = strategies.get(f"embedding{i}")
        if c is None:
            raise KeyError(f"embedding{i} is not in the strategy")
        if c.device_type != DEVICE_GPU or c.num_parts() != 1 or len(c.device_ids) != 1:
            raise ValueError(f"embedding{i}: only a whole table on one GPU is a table-wise placement")
        own.append(c.device_ids[0])
    return own


def place_tables(rows, bags, dim, world, kind="balanced", hbm_bytes=150e9):
    """table -> rank.  modulo: the reference generator's i % G.  balanced: greedy longest-processing-
    time by a per-table cost in the classes measured on one B200 (profiles/r1_notes.md): rows from
    HBM plus a row read-modify-write per lookup (tables >> L2) > L2-resident tables through the
    sorted path > tables whose gradient fits on chip; plus a per-bag term for the pooled-row write /
    output-grad read; ties and capacity (hbm_bytes per GPU) by bytes."""
    T = len(rows)
    if kind == "modulo" or world == 1:
        return [t % world for t in range(T)]

    def cost(t):
        b = rows[t] * dim * 4
        per = 1.0 if b > (64 << 20) else (0.25 if rows[t] <= 16 or (b <= (80 << 10) and bags[t] >= 8) else 0.55)
        return bags[t] * per + 0.6

    load, mem, own = [0.0] * world, [0.0] * world, [0] * T
    for t in sorted(range(T), key=lambda t: -cost(t)):
        cands = sorted(range(world), key=lambda g: (load[g], mem[g]))
        g = next((g for g in cands if mem[g] + rows[t] * dim * 4 < hbm_bytes), cands[0])
        own[t], load[g], mem[g] = g, load[g] + cost(t), mem[g] + rows[t] * dim * 4
    return own


# ---------------------------------------------------------------------------------------------
# Shard plan with column splits and pipeline groups (table-wise sharding)
# ---------------------------------------------------------------------------------------------
@dataclass
class Part:
    """A placement unit: columns [col0, col0 + width) of table `table`, whole on GPU `rank`, processed
    in pipeline group `group` of that rank.  width == dim: the reference's unit, the whole table
    (/root/reference/examples/cpp/DLRM/strategies/dlrm_strategy.cc:252-256).  width < dim: a heavy table
    split by COLUMNS over several ranks — every rank of the split sees the same indices and pools /
    updates its own columns, so pooled rows and updated tables are bit-identical to the unsplit table
    and the exchange moves exactly the same bytes (each rank sends its columns of every sample)."""
    table: int
    col0: int
    width: int
    rank: int
    group: int = 0


def part_cost(rows, bag, dim, width):
    """Estimated time (ms at batch 65536) of one part on one B200: forward + index sort + segmented sum,
    in the three regimes measured per table in isolation (scripts/per_table_cost.py,
    profiles/r2_per_table_cost.md): rows streamed from HBM (table far larger than L2), L2-resident
    tables through the sorted path, tables whose gradient stays on chip.  A half-width part keeps the
    whole index sort and the per-lookup walk, so it costs more than half."""
    f = width / float(dim)
    half = 0.55 if f < 1.0 else 1.0          # measured: forward / level 1 of a D/2 part vs the full table
    scale = dim / 128.0
    table_bytes = rows * dim * 4
    if table_bytes > (64 << 20):
        per = 0.0065 * half + 0.0024 + 0.0145 * half
    elif rows <= 16 or (rows * dim * 4 <= (80 << 10) and bag >= 8):
        per = 0.0025 * half + 0.0023
    else:
        per = (0.0029 + 0.0037) * (0.65 if f < 1.0 else 1.0) + 0.0024
    return bag * per * max(scale, 0.25) + 0.01 * f * scale


def plan_shards(rows, bags, dim, world, kind="balanced", split=True, groups=1, hbm_bytes=150e9):
    """-> list of Part, one or more per table, in a deterministic order (every rank computes the same).

    kind "modulo"  : the reference generator's placement, table i whole on GPU i % world.
    kind "balanced": greedy longest-processing-time over part_cost; with split=True a table whose cost
                     exceeds 3/4 of the mean load per rank is first cut into two column halves (and a
                     half that still does into quarters), so that no single table sets the step.
    groups         : pipeline groups per rank (forward(g+1) overlaps the exchange of g, the push of g+1
                     overlaps the backward of g).  Parts of one group share one width (the kernels take
                     one row width per launch): a rank holding parts of two widths gets one group per
                     width; otherwise its parts are dealt into `groups` bins of equal cost."""
    T = len(rows)
    parts = [Part(t, 0, dim, t % world) for t in range(T)]
    if kind != "modulo" and world > 1:
        cost = lambda p: part_cost(rows[p.table], bags[p.table], dim, p.width)
        mean = sum(cost(p) for p in parts) / world
        if split:
            pieces = [1] * T  # a table is never cut into more parts than there are ranks to hold them
            for _ in range(2):  # halves, then quarters
                nxt = []
                for p in parts:
                    if (cost(p) > 0.75 * mean and p.width % 8 == 0 and p.width >= 32
                            and pieces[p.table] < world):
                        h = p.width // 2
                        nxt += [Part(p.table, p.col0, h, 0), Part(p.table, p.col0 + h, h, 0)]
                        pieces[p.table] += 1
                    else:
                        nxt.append(p)
                parts = nxt
        load, mem = [0.0] * world, [0.0] * world
        for p in sorted(parts, key=lambda p: (-cost(p), p.table, p.col0)):
            b = rows[p.table] * p.width * 4
            cands = sorted(range(world), key=lambda g: (load[g], mem[g], g))
            # two parts of one table never share a rank (they would be the unsplit table again)
            cands = [g for g in cands if not any(q.table == p.table and q.rank == g and q is not p and q.group == -1
                                                 for q in parts)] or cands
            g = next((g for g in cands if mem[g] + b < hbm_bytes), cands[0])
            p.rank, p.group = g, -1   # -1: placed, group not yet assigned
            load[g] += cost(p)
            mem[g] += b
        parts.sort(key=lambda p: (p.table, p.col0))
    # pipeline groups per rank
    for g in range(world):
        mine = [p for p in parts if p.rank == g]
        widths = sorted({p.width for p in mine}, reverse=True)
        if groups <= 1 and len(widths) <= 1:
            for p in mine:
                p.group = 0
        elif len(widths) > 1:
            for p in mine:
                p.group = widths.index(p.width)
        else:
            bins = [0.0] * groups
            for p in sorted(mine, key=lambda p: (-part_cost(rows[p.table], bags[p.table], dim, p.width), p.table)):
                k = min(range(groups), key=lambda k: (bins[k], k))
                p.group = k
                bins[k] += part_cost(rows[p.table], bags[p.table], dim, p.width)
    return parts


def num_groups(parts):
    return 1 + max((p.group for p in parts), default=0)


def plan_strategy(parts, world):
    """The plan in the reference's strategy-file vocabulary (write_strategy): a whole table is
    "embedding<i>" with a [1,1] grid on its GPU; a table split by columns is a [c,1] grid — dim[0] is the
    feature dimension (/root/reference/src/runtime/model.cc:726-736) — of c = dim / (narrowest part)
    equal cells whose device ids are the parts' ranks in column order, a part twice as wide as the
    narrowest one filling two neighbouring cells (a half next to two quarters: [4,1], ids a a b c).
    (The reference's Embedding refuses such a grid, src/ops/embedding.cu:84-85: the file is then only
    readable by this library's planner.)"""
    s = {}
    by_table = {}
    for p in parts:
        by_table.setdefault(p.table, []).append(p)
    for t, ps in by_table.items():
        ps = sorted(ps, key=lambda p: p.col0)
        cell = min(p.width for p in ps)
        if any(p.width % cell or p.col0 % cell for p in ps):
            raise ValueError(f"embedding{t}: column parts are not multiples of the narrowest one")
        ids = [int(p.rank) for p in ps for _ in range(p.width // cell)]
        s[f"embedding{t}"] = ParallelConfig(DEVICE_GPU, [len(ids), 1], ids)
    s["concat"] = ParallelConfig(DEVICE_GPU, [1, world], list(range(world)))
    return s


def parts_from_strategy(strategies, num_tables, dim):
    """Inverse of plan_strategy (groups are not part of the file: all 0): neighbouring cells of one
    table on the same GPU are one part."""
    parts = []
    for t in range(num_tables):
        c = strategies[f"embedding{t}"]
        n = c.dims[0] if c.dims else 0
        if c.device_type != DEVICE_GPU or len(c.dims) != 2 or c.dims[1] != 1 or n < 1 or dim % n != 0 \
                or len(c.device_ids) != n:
            raise ValueError(f"embedding{t}: expected a [c,1] grid on GPUs with c dividing the row width")
        w = dim // n
        i = 0
        while i < n:
            j = i
            while j + 1 < n and c.device_ids[j + 1] == c.device_ids[i]:
                j += 1
            parts.append(Part(t, i * w, (j - i + 1) * w, int(c.device_ids[i])))
            i = j + 1
    return parts
