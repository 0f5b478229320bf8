"""Synthetic workloads of BASELINE.md §3.4: the five configs as concrete, seeded inputs.

Generator: numpy PCG64, seed = 1234 + table_id.  Weights uniform
[-sqrt(1/R), sqrt(1/R)] as /root/reference/examples/cpp/DLRM/dlrm.cc:44-45; indices uniform in
[0, R) (the reference uses rand() % R, dlrm.cc:416) or Zipf over ranks through a fixed seeded
rank->row permutation; output-grad uniform [-1, 1].
"""
from dataclasses import dataclass, field
from typing import List

import numpy as np

# examples/cpp/DLRM/run_dlrm_kaggle_day1.sh:5 (the 10.1M-max cardinality list)
KAGGLE_ROWS = [1461, 584, 10131227, 2202608, 306, 24, 12518, 634, 4, 93146, 5684, 8351593, 3195,
               28, 14993, 5461306, 11, 5653, 2173, 4, 7046547, 18, 16, 286181, 105, 142572]
# public MLPerf DLRM Criteo-Terabyte cardinalities capped at 40M (not in the reference)
MLPERF_ROWS = [39884406, 39043, 17289, 7420, 20263, 3, 7120, 1543, 63, 38532951, 2953546, 403346,
               10, 2208, 11938, 155, 4, 976, 14, 39979771, 25641295, 39664984, 585935, 12972,
               108, 36]


@dataclass
class Workload:
    name: str
    rows: List[int]
    dim: int
    bags: List[int]          # bag width L per table
    batch: int
    aggr: int = 21           # AGGR_MODE_SUM
    dist: str = "uniform"    # or "zipf"
    zipf_alpha: float = 1.05
    lr: float = 0.01
    note: str = ""

    @property
    def num_tables(self):
        return len(self.rows)

    @property
    def lookups(self):
        return sum(self.batch * l for l in self.bags)

    def algorithmic_bytes(self, unique_rows):
        """total = 16 N + 4 D N + 8 D T B + 8 D U."""
        n, d = self.lookups, self.dim
        return 16 * n + 4 * d * n + 8 * d * self.num_tables * self.batch + 8 * d * unique_rows


def mlperf_bags(seed=1234):
    """cfg3: a per-table constant bag width in [1, 20], drawn once with seed 1234."""
    return [int(x) for x in np.random.Generator(np.random.PCG64(seed)).integers(1, 21, 26)]


def config(n, scale=1.0):
    """BASELINE config n (1-based).  scale < 1 shrinks batch (and rows for the big ones) for tests."""
    s = lambda x: max(1, int(round(x * scale)))
    if n == 1:
        return Workload("cfg1-dlrm-random-8x1Mx64", [s(1_000_000)] * 8, 64, [1] * 8, s(2048),
                        note="examples/cpp/DLRM/run_random.sh:9")
    if n == 2:
        return Workload("cfg2-criteo-kaggle-26xD16", [s(r) if r > 1000 else r for r in KAGGLE_ROWS],
                        16, [1] * 26, s(16384), note="run_dlrm_kaggle_day1.sh:5 cardinalities")
    if n == 3:
        return Workload("cfg3-mlperf-terabyte-26xD128", [s(r) if r > 1000 else r for r in MLPERF_ROWS],
                        128, mlperf_bags(), s(65536), note="multi-hot 1..20 per table")
    if n == 4:
        return Workload("cfg4-large-8x100Mx128", [s(100_000_000)] * 8, 128, [1] * 8, s(65536),
                        note="row-wise sharded over 8 GPUs")
    if n == 5:
        return Workload("cfg5-zipf-avg-bag64", [s(10_000_000)] * 8, 128, [64] * 8, s(262144),
                        aggr=22, dist="zipf", note="Zipf alpha=1.05, AVG")
    raise ValueError(n)


def table_seed(table_id):
    return 1234 + table_id


def gen_weights(rows, dim, table_id):
    rng = np.random.Generator(np.random.PCG64(table_seed(table_id)))
    a = np.sqrt(1.0 / rows)
    return rng.uniform(-a, a, (rows, dim)).astype(np.float32)


def _zipf_ranks(rng, rows, n, alpha):
    """n samples of rank in [0, rows) with P(rank k) ~ 1/(k+1)^alpha (inverse-CDF on a table
    for the head, continuous approximation for the tail)."""
    head = min(rows, 1 << 20)
    w = 1.0 / np.power(np.arange(1, head + 1, dtype=np.float64), alpha)
    if rows > head:  # integral of x^-alpha from head+0.5 to rows+0.5
        tail = ((head + 0.5) ** (1 - alpha) - (rows + 0.5) ** (1 - alpha)) / (alpha - 1)
    else:
        tail = 0.0
    total = w.sum() + tail
    u = rng.random(n) * total
    cdf = np.cumsum(w)
    out = np.searchsorted(cdf, u, side="right").astype(np.int64)
    in_tail = u >= cdf[-1]
    if tail > 0 and in_tail.any():
        r = u[in_tail] - cdf[-1]
        x = ((head + 0.5) ** (1 - alpha) - r * (alpha - 1)) ** (1.0 / (1 - alpha)) - 0.5
        out[in_tail] = np.clip(x.astype(np.int64), head, rows - 1)
    return np.minimum(out, rows - 1)


def gen_indices(wl, table_id, batch_no=0):
    """int64 [B][L] for one table and one batch of the pool."""
    rows, L = wl.rows[table_id], wl.bags[table_id]
    rng = np.random.Generator(np.random.PCG64([table_seed(table_id), 7919 + batch_no]))
    if wl.dist == "uniform":
        return rng.integers(0, rows, (wl.batch, L), dtype=np.int64)
    ranks = _zipf_ranks(rng, rows, wl.batch * L, wl.zipf_alpha)
    # fixed seeded rank->row map so hot rows are not address-adjacent: an affine bijection
    # mod rows (odd multiplier coprime to rows is enough for a permutation)
    prng = np.random.Generator(np.random.PCG64(table_seed(table_id) + 99991))
    a = int(prng.integers(1, rows)) | 1
    while np.gcd(a, rows) != 1:
        a += 2
    c = int(prng.integers(0, rows))
    rows_idx = (ranks.astype(object) * a + c) % rows if rows > (1 << 31) else (ranks * a + c) % rows
    return np.asarray(rows_idx, dtype=np.int64).reshape(wl.batch, L)


def gen_outgrad(wl, table_id, batch_no=0):
    rng = np.random.Generator(np.random.PCG64([table_seed(table_id), 104729 + batch_no]))
    return rng.uniform(-1.0, 1.0, (wl.batch, wl.dim)).astype(np.float32)
