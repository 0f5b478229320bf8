"""CPU tests of bench.py's accounting (no GPU): the algorithmic bytes it divides by are SURVEY §8d's
formula, U is counted exactly from the index tensors, a column-split table is credited once over the
ranks of the split, and the zipf generator is a permutation of ranks (hot rows not address-adjacent)."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


@pytest.fixture(scope="module")
def bench():
    return importlib.import_module("bench")


class _NoColl:
    """DeviceWorkload also builds the product's EmbeddingCollection, which takes CUDA tensors only (there
    is no CPU fallback); the accounting under test never touches it."""
    @staticmethod
    def EmbeddingCollection(*a, **k):
        return None


def _tiny(ffemb):
    return ffemb.synth.Workload("tiny", [50, 7, 1000], 16, [3, 1, 5], 64)


def test_algorithmic_bytes_are_the_survey_formula(bench, ffemb):
    S, wl = ffemb.sharding, _tiny(ffemb)
    parts = [S.Part(t, 0, wl.dim, 0, 0) for t in range(3)]
    dw = bench.DeviceWorkload(torch, _NoColl, wl, parts, torch.device("cpu"), pool=3, seed=1234)
    fwd, bwd, U = dw.bytes_per_step()
    B, D = wl.batch, wl.dim
    N = sum(B * l for l in wl.bags)
    T = len(wl.rows)
    u_exact = np.mean([[np.unique(dw.idx[t][k].numpy()).size for k in range(3)] for t in range(T)], axis=1).sum()
    assert U == pytest.approx(u_exact)
    # total = 16 N + 4 D N + 8 D T B + 8 D U
    assert fwd + bwd == pytest.approx(16 * N + 4 * D * N + 8 * D * T * B + 8 * D * u_exact)
    assert fwd == pytest.approx(8 * N + 4 * D * N + 4 * D * T * B)
    assert dw.lookups() == N
    # the cache-aware count never credits more than the §8d count
    cf, cb = dw.compulsory_bytes_per_step()
    assert cf <= fwd and cb == pytest.approx(bwd)


def test_a_column_split_table_is_credited_once_over_its_ranks(bench, ffemb):
    S, wl = ffemb.sharding, _tiny(ffemb)
    whole = bench.DeviceWorkload(torch, _NoColl, wl, [S.Part(2, 0, 16, 0, 0)], torch.device("cpu"), 2, 1234)
    halves = [bench.DeviceWorkload(torch, _NoColl, wl, [S.Part(2, c, 8, r, 0)], torch.device("cpu"), 2, 1234)
              for r, c in ((0, 0), (1, 8))]
    # the ranks of a split see the same index batches (seeded per table) ...
    assert all(torch.equal(a, b) for a, b in zip(halves[0].idx[0], halves[1].idx[0]))
    assert all(torch.equal(a, b) for a, b in zip(halves[0].idx[0], whole.idx[0]))
    # ... every rank walks all its lookups, but rows / distinct rows are credited by width
    fw, bw, Uw = whole.bytes_per_step()
    parts = [h.bytes_per_step() for h in halves]
    n = wl.batch * wl.bags[2]
    assert sum(p[2] for p in parts) == pytest.approx(Uw)
    assert sum(p[0] for p in parts) == pytest.approx(fw + 8 * n)   # the indices are read by both ranks
    assert sum(p[1] for p in parts) == pytest.approx(bw + 8 * n)
    assert sum(h.lookups() for h in halves) == 2 * whole.lookups()


def test_zipf_indices_follow_the_rank_law_through_a_fixed_permutation(bench):
    g = torch.Generator(device="cpu")
    g.manual_seed(5)
    rows = 10_000
    a = bench.zipf_indices(torch, g, rows, (20000, 4), 1.05, 3, torch.device("cpu"))
    assert a.dtype == torch.int64 and a.shape == (20000, 4) and int(a.min()) >= 0 and int(a.max()) < rows
    vals, counts = torch.unique(a, return_counts=True)
    order = torch.argsort(counts, descending=True)
    top = counts[order][:3].double() / a.numel()
    # P(rank k) ~ 1/(k+1)^1.05: the head row takes ~ 1/H of the mass, the next ~ 2^-1.05 of that
    H = float((1.0 / torch.arange(1, rows + 1, dtype=torch.float64) ** 1.05).sum())
    assert top[0] == pytest.approx(1 / H, rel=0.1)
    assert top[1] == pytest.approx(2 ** -1.05 / H, rel=0.15)
    # hot rows are not address-adjacent (the rank -> row map is a permutation, not the identity)
    hot = vals[order][:8]
    assert not torch.equal(hot.sort().values, torch.arange(8))
    g.manual_seed(5)
    b = bench.zipf_indices(torch, g, rows, (20000, 4), 1.05, 3, torch.device("cpu"))
    assert torch.equal(a, b)
