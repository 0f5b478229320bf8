"""GPU parity tests (run on the B200 box): the CUDA path, called through the C ABI, against the
oracle on the same seeded inputs, against the committed golden vectors of the reference's own
code, and — at full size — through size-independent properties.

Bars (north_star): gathered rows / forward sums are BIT-EXACT (sequential accumulation in j);
updated tables and gradients agree within 1e-6 relative error where rows repeat (summation
order), bit-exact where they do not; untouched rows are bit-identical.
"""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
from conftest import golden_path  # noqa: E402
from make_golden import CPU_CASES, GPU_CASES, make_inputs  # noqa: E402

pytestmark = pytest.mark.gpu
REL_TOL = 1e-6  # north_star: pooled outputs / updated tables within 1e-6 relative


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    torch.cuda.set_device(0)
    return torch


def dev(torch, x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


def rel_err(a, b):
    return np.max(np.abs(a.astype(np.float64) - b) / np.maximum(np.abs(b), 1e-30))


def table_tol(W, n_dup=1):
    """1e-6 relative on the updated table; absolute floor = 1e-6 * scale of the table."""
    return REL_TOL * np.maximum(np.abs(W), np.abs(W).max())


# ---------------------------------------------------------------------------------------------
# forward
# ---------------------------------------------------------------------------------------------
FWD_SHAPES = [  # R, D, L, B
    (1000, 16, 1, 4097), (1000, 64, 1, 2048), (1000, 128, 1, 513), (50, 4, 1, 100),
    (50, 8, 1, 100), (777, 32, 1, 1000), (300, 256, 1, 64), (300, 512, 1, 33),
    (1000, 16, 7, 1000), (1000, 64, 20, 257), (1000, 128, 20, 129), (1000, 128, 64, 65),
    (100, 48, 5, 99), (100, 20, 3, 77), (100, 100, 9, 41), (64, 256, 33, 17), (64, 512, 3, 9),
    (100, 18, 4, 55), (100, 7, 2, 31), (100, 1, 5, 31), (40, 130, 3, 20),  # generic (D % 4 != 0)
    (4, 16, 1, 5000), (3, 128, 20, 300),  # tiny tables: every row hot
]


@pytest.mark.parametrize("R,D,L,B", FWD_SHAPES)
@pytest.mark.parametrize("aggr", [21, 22, 122])
def test_forward_bit_exact_vs_oracle(ffemb, oracle, torch_cuda, R, D, L, B, aggr):
    torch = torch_cuda
    W, idx, _ = make_inputs(R, D, L, B, seed=R + D + L + B)
    W[0, 0] = -0.0  # out = 0 + w must give +0
    op = ffemb.Embedding(R, D, aggr)
    out = op.forward(dev(torch, idx), dev(torch, W))
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), oracle.forward(idx, W, aggr))


@pytest.mark.parametrize("L", [3, 7, 64])
@pytest.mark.parametrize("scale", [1e-38, 1e-33, 1e-28, 1e30])
def test_forward_avg_literal_division_edge_ranges(ffemb, oracle, torch_cuda, L, scale):
    """AGGR_MODE_AVG divides inside the bag loop (embedding.cu:186).  The kernel replaces the IEEE
    division by an exact-residual correction that is only valid for |a| >= 2^-100 and falls back
    to the division below that: denormal / tiny / huge accumulators must still be bit-exact."""
    torch = torch_cuda
    R, D, B = 200, 64, 300
    W, idx, _ = make_inputs(R, D, L, B, seed=L)
    W = (W.astype(np.float64) / np.abs(W).max() * scale).astype(np.float32)
    W[1] = 0.0
    W[2, ::2] = -0.0
    op = ffemb.Embedding(R, D, 22)
    out = op.forward(dev(torch, idx), dev(torch, W))
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().view(np.uint32), oracle.forward(idx, W, 22).view(np.uint32))


def test_forward_golden_reference_cpu(ffemb, torch_cuda):
    """Against the reference's own CPU tasks (tests/golden/ref_cpu.npz)."""
    torch = torch_cuda
    z = np.load(golden_path("ref_cpu.npz"))
    for name, R, D, L, B, aggr, seed in CPU_CASES:
        W, idx, _ = make_inputs(R, D, L, B, seed)
        out = ffemb.Embedding(R, D, aggr).forward(dev(torch, idx), dev(torch, W))
        assert np.array_equal(out.cpu().numpy(), z[name + ".fwd"]), name


@pytest.mark.skipif(not os.path.exists(golden_path("ref_gpu.npz")), reason="GPU goldens not generated yet")
def test_forward_and_step_golden_reference_gpu(ffemb, torch_cuda):
    """Against the reference's own GPU kernels run on a B200 (tests/golden/ref_gpu.npz)."""
    torch = torch_cuda
    z = np.load(golden_path("ref_gpu.npz"))
    for name, R, D, L, B, aggr, seed in GPU_CASES:
        W, idx, g = make_inputs(R, D, L, B, seed)
        op = ffemb.Embedding(R, D, aggr)
        dW = dev(torch, W)
        out = op.forward(dev(torch, idx), dW)
        assert np.array_equal(out.cpu().numpy(), z[name + ".fwd"]), name
        if D % 4 == 0:
            op.backward_sgd(dev(torch, idx), dev(torch, g), dW, 0.01)
            got, want = dW.cpu().numpy(), z[name + ".wnew"]
            cnt = np.bincount(idx.reshape(-1), minlength=R)
            assert np.array_equal(got[cnt <= 1], want[cnt <= 1]), name
            assert np.all(np.abs(got - want) <= table_tol(W)), name


def test_forward_batched_ragged_tables(ffemb, oracle, torch_cuda):
    torch = torch_cuda
    rows, bags, D, B = [5, 1000, 33, 20000, 4], [1, 3, 20, 2, 7], 32, 777
    Ws, idxs = [], []
    for t, (R, L) in enumerate(zip(rows, bags)):
        W, idx, _ = make_inputs(R, D, L, B, seed=100 + t)
        Ws.append(W), idxs.append(idx)
    for aggr in (21, 22):
        coll = ffemb.EmbeddingCollection([dev(torch, w) for w in Ws], bags, aggr)
        outs = coll.forward([dev(torch, i) for i in idxs])
        for t in range(len(rows)):
            assert np.array_equal(outs[t].cpu().numpy(), oracle.forward(idxs[t], Ws[t], aggr)), t


def test_forward_empty_and_unaligned(ffemb, oracle, torch_cuda):
    torch = torch_cuda
    W, idx, _ = make_inputs(100, 16, 2, 10, seed=1)
    op = ffemb.Embedding(100, 16, 21)
    out = op.forward(dev(torch, idx[:0]), dev(torch, W))  # empty batch
    assert out.shape == (0, 16)
    out = op.forward(dev(torch, np.zeros((6, 0), np.int64)), dev(torch, W))  # empty bags pool to 0
    assert not out.cpu().numpy().any()
    # a table view that is only 4-byte aligned takes the generic path and stays exact
    buf = torch.zeros(100 * 16 + 1, device="cuda")
    Wd = buf[1:].view(100, 16)
    Wd.copy_(dev(torch, W))
    out = ffemb.Embedding(100, 16, 21).forward(dev(torch, idx), Wd)
    assert np.array_equal(out.cpu().numpy(), oracle.forward(idx, W))


def test_error_behaviour(ffemb, torch_cuda):
    torch = torch_cuda
    W, idx, _ = make_inputs(10, 16, 1, 4, seed=2)
    with pytest.raises(ffemb.FFEmbError) as e:  # A.2: AGGR_MODE_NONE is invalid
        ffemb.Embedding(10, 16, ffemb.AGGR_MODE_NONE).forward(dev(torch, idx), dev(torch, W))
    assert e.value.code == 1
    with pytest.raises(ValueError):  # the task body's shape asserts
        ffemb.Embedding(10, 16).forward(dev(torch, idx), dev(torch, W), output=torch.empty(3, 16, device="cuda"))
    bad = torch.zeros(1, dtype=torch.int64, device="cuda")
    import ctypes as C
    i = dev(torch, np.array([0, 5, 10, -1, 9], np.int64))
    assert ffemb.lib().ffemb_check_indices(i.data_ptr(), 5, 10, bad.data_ptr(), None) == 0
    assert bad.item() == 2


# ---------------------------------------------------------------------------------------------
# backward
# ---------------------------------------------------------------------------------------------
BWD_SHAPES = [  # R, D, L, B   (D % 4 == 0 for the sorted path)
    (100000, 16, 1, 4097), (100000, 64, 1, 2048), (100000, 128, 1, 1000), (5000, 32, 1, 3000),
    (1000, 16, 7, 1000), (1000, 64, 20, 257), (1000, 128, 20, 300), (2000, 128, 64, 130),
    (4, 16, 1, 20000), (3, 128, 20, 700), (1, 64, 1, 5000), (11, 8, 3, 9999), (50, 4, 1, 100),
    (100, 48, 5, 99), (100, 20, 3, 77), (64, 256, 9, 70), (64, 512, 3, 40), (7, 16, 64, 3000),
    # tables whose whole gradient fits on chip (ffemb_backward_small.cu): tiny = register
    # accumulators (R <= 16), small = shared-memory accumulator; ragged tails, bag > lanes per slot
    (3, 128, 19, 3001), (4, 128, 16, 2049), (10, 64, 16, 4099), (14, 32, 20, 1777), (16, 128, 1, 20011),
    (1, 128, 3, 6000), (2, 96, 64, 515), (36, 128, 18, 2000), (63, 128, 3, 7001), (155, 128, 7, 4099),
    (108, 64, 18, 1500), (17, 32, 40, 999), (160, 128, 5, 5000), (300, 48, 2, 20000),
]


@pytest.mark.parametrize("R,D,L,B", BWD_SHAPES)
@pytest.mark.parametrize("aggr", [21, 22])
def test_backward_sgd_vs_reference_step(ffemb, oracle, torch_cuda, R, D, L, B, aggr):
    torch = torch_cuda
    W, idx, g = make_inputs(R, D, L, B, seed=7 * R + D + L + B)
    want = oracle.reference_step(idx, g, W, aggr, lr=0.01)  # zero-fill -> embed_backward -> sgd_update
    dW = dev(torch, W)
    ffemb.Embedding(R, D, aggr).backward_sgd(dev(torch, idx), dev(torch, g), dW, 0.01)
    torch.cuda.synchronize()
    got = dW.cpu().numpy()
    cnt = np.bincount(idx.reshape(-1), minlength=R)
    assert np.array_equal(got[cnt == 0], W[cnt == 0])            # untouched rows bit-identical
    assert np.array_equal(got[cnt == 1], want[cnt == 1])         # single lookups: same arithmetic
    # repeated rows: summation order differs; compare against the fp64 sum
    exact = W.astype(np.float64) - 0.01 * oracle.backward_dense_f64(idx, g, R, aggr)
    assert np.all(np.abs(got - exact) <= table_tol(W) * np.maximum(1, np.sqrt(cnt))[:, None])
    assert np.all(np.abs(got - want) <= 2 * table_tol(W) * np.maximum(1, np.sqrt(cnt))[:, None])


@pytest.mark.parametrize("R,D,L,B", [(100000, 16, 1, 4097), (1000, 64, 20, 257), (4, 16, 1, 20000), (40, 18, 3, 50)])
@pytest.mark.parametrize("aggr", [21, 22])
def test_backward_dense_variants(ffemb, oracle, torch_cuda, R, D, L, B, aggr):
    """API-exact unfused mode: accumulate into a dense grad (atomic and deterministic)."""
    torch = torch_cuda
    W, idx, g = make_inputs(R, D, L, B, seed=R + B)
    pre = np.random.default_rng(0).uniform(-1, 1, (R, D)).astype(np.float32)  # A.5: accumulated into
    exact = pre.astype(np.float64) + oracle.backward_dense_f64(idx, g, R, aggr)
    cnt = np.bincount(idx.reshape(-1), minlength=R)
    tol = REL_TOL * np.maximum(np.abs(exact), 1.0) * np.maximum(1, cnt)[:, None]
    op = ffemb.Embedding(R, D, aggr)
    variants = [False] if D % 4 else [False, True]
    for det in variants:
        dG = dev(torch, pre)
        op.backward(dev(torch, idx), dev(torch, g), dG, deterministic=det)
        got = dG.cpu().numpy()
        assert np.all(np.abs(got - exact) <= tol), det
        assert np.array_equal(got[cnt == 0], pre[cnt == 0])


@pytest.mark.parametrize("R,D,L,B", [(37, 64, 8, 20000), (5, 128, 9, 9000), (5000, 64, 8, 20000)])
def test_backward_is_deterministic_and_atomic_is_the_comparison(ffemb, torch_cuda, R, D, L, B):
    """Runs of the deterministic backward are bit-identical on every path (small: shared-memory
    accumulator, tiny: register accumulators, sorted) — the reference's atomics are not."""
    torch = torch_cuda
    W, idx, g = make_inputs(R, D, L, B, seed=5)  # heavy duplication
    outs = []
    for _ in range(3):
        dW = dev(torch, W)
        ffemb.Embedding(R, D, 21).backward_sgd(dev(torch, idx), dev(torch, g), dW, 0.01)
        outs.append(dW.cpu().numpy())
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])


def test_backward_path_selection(ffemb):
    """Which tables leave the sorted path (informational query of the C ABI)."""
    path = ffemb.lib().ffemb_backward_path
    assert path(3, 19, 128, 65536) == 1 and path(16, 1, 128, 65536) == 1
    assert path(36, 18, 128, 65536) == 2 and path(108, 18, 128, 65536) == 2
    assert path(155, 7, 128, 65536) == 0 and path(63, 3, 128, 65536) == 0   # short bags: little re-reading to save
    assert path(976, 20, 128, 65536) == 0 and path(40000000, 20, 128, 65536) == 0
    assert path(3, 19, 16, 65536) == 0 and path(3, 19, 256, 65536) == 0   # D outside [32, 128]
    assert path(3, 1, 128, 100) == 0                                       # too little work


@pytest.mark.parametrize("D,B", [(16, 2500), (64, 9000), (128, 17000)])
def test_backward_batched_ragged_tables(ffemb, oracle, torch_cuda, D, B):
    """One batched call mixing tiny, small and sorted tables (for D >= 32 and enough lookups)."""
    torch = torch_cuda
    rows, bags = [5, 100000, 33, 20000, 4, 1], [1, 3, 20, 2, 7, 1]
    Ws, idxs, gs = [], [], []
    for t, (R, L) in enumerate(zip(rows, bags)):
        W, idx, g = make_inputs(R, D, L, B, seed=200 + t)
        Ws.append(W), idxs.append(idx), gs.append(g)
    for aggr in (21, 22):
        dWs = [dev(torch, w) for w in Ws]
        coll = ffemb.EmbeddingCollection(dWs, bags, aggr)
        coll.backward_sgd([dev(torch, i) for i in idxs], [dev(torch, g) for g in gs], 0.01)
        for t, R in enumerate(rows):
            got = dWs[t].cpu().numpy()
            cnt = np.bincount(idxs[t].reshape(-1), minlength=R)
            want = oracle.reference_step(idxs[t], gs[t], Ws[t], aggr, lr=0.01)
            exact = Ws[t].astype(np.float64) - 0.01 * oracle.backward_dense_f64(idxs[t], gs[t], R, aggr)
            assert np.array_equal(got[cnt <= 1], want[cnt <= 1]), t
            assert np.all(np.abs(got - exact) <= table_tol(Ws[t]) * np.maximum(1, np.sqrt(cnt))[:, None]), t


def test_backward_streaming_and_resident_tables_interleaved(ffemb, oracle, torch_cuda):
    """Tables far larger than L2 (> 64 MB) are keyed first and their level-1 tiles are handed out
    interleaved with the L2-resident tables' (mix_cta): same results as table by table, and the
    batched call stays bit-identical to per-table calls on single-lookup rows."""
    torch = torch_cuda
    D, B = 128, 3000
    rows, bags = [2000, 140000, 977, 135000, 64], [5, 2, 9, 1, 3]   # 140000*512 B = 72 MB > 64 MB
    Ws, idxs, gs = [], [], []
    for t, (R, L) in enumerate(zip(rows, bags)):
        W, idx, g = make_inputs(R, D, L, B, seed=400 + t)
        Ws.append(W), idxs.append(idx), gs.append(g)
    dWs = [dev(torch, w) for w in Ws]
    coll = ffemb.EmbeddingCollection(dWs, bags, 21)
    coll.backward_sgd([dev(torch, i) for i in idxs], [dev(torch, g) for g in gs], 0.01)
    torch.cuda.synchronize()
    for t, R in enumerate(rows):
        got = dWs[t].cpu().numpy()
        cnt = np.bincount(idxs[t].reshape(-1), minlength=R)
        want = oracle.reference_step(idxs[t], gs[t], Ws[t], 21, lr=0.01)
        exact = Ws[t].astype(np.float64) - 0.01 * oracle.backward_dense_f64(idxs[t], gs[t], R, 21)
        assert np.array_equal(got[cnt <= 1], want[cnt <= 1]), t
        assert np.all(np.abs(got - exact) <= table_tol(Ws[t]) * np.maximum(1, np.sqrt(cnt))[:, None]), t
        single = [dev(torch, Ws[t])]
        ffemb.EmbeddingCollection(single, [bags[t]], 21).backward_sgd([dev(torch, idxs[t])], [dev(torch, gs[t])], 0.01)
        alone = single[0].cpu().numpy()
        assert np.array_equal(alone[cnt <= 1], got[cnt <= 1]), t
        assert np.all(np.abs(alone - got) <= REL_TOL * np.abs(got).max()), t   # repeated rows: tile cuts differ


@pytest.mark.parametrize("B,L", [(100, 1), (4096, 1), (4097, 1), (1024, 8), (8193, 1), (16384, 1), (819, 20),
                                 (16385, 1), (30000, 3)])
@pytest.mark.parametrize("R", [37, 100000, 3000000])
def test_backward_sort_is_the_stable_sort_by_row(ffemb, torch_cuda, B, L, R):
    """White box: after ffemb_backward_prepare the workspace holds the lookups sorted by row with the
    bags ascending inside a run of equal rows (the summation order of the segmented sum) — checked
    against numpy's stable sort, from a hundred lookups to beyond one radix-sort tile."""
    torch = torch_cuda
    D = 16
    if ffemb.lib().ffemb_backward_path(R, L, D, B) != 0:
        pytest.skip("this table takes the bag-major path: nothing is sorted")
    rng = np.random.default_rng(B * 31 + L + R)
    idx = rng.integers(0, R, size=(B, L), dtype=np.int64)
    coll = ffemb.EmbeddingCollection([torch.zeros((R, D), device="cuda")], [L])
    coll.backward_prepare([dev(torch, idx)])
    torch.cuda.synchronize()
    n = B * L
    seg = (n * 4 + 255) // 256 * 256  # workspace carve: keys_a, keys_b, vals_a, vals_b (256-byte aligned)
    ws = coll._workspaces(B)[0].cpu().numpy()
    keys = ws[seg:seg + 4 * n].view(np.uint32)
    vals = ws[3 * seg:3 * seg + 4 * n].view(np.uint32)
    order = np.argsort(idx.reshape(-1), kind="stable")
    assert np.array_equal(keys, idx.reshape(-1)[order].astype(np.uint32))
    assert np.array_equal(vals, (np.arange(n) // L)[order].astype(np.uint32))  # table 0: value = bag number


def test_backward_split_in_prepare_and_apply_is_bit_identical(ffemb, torch_cuda):
    """ffemb_backward_prepare (sort, indices only, here on a side stream) + ffemb_backward_sgd_prepared
    == ffemb_backward_sgd_batched, bit for bit."""
    torch = torch_cuda
    rows, bags, D, B = [5, 100000, 33, 20000, 4, 977], [1, 3, 20, 2, 7, 9], 64, 9000
    Ws, idxs, gs = [], [], []
    for t, (R, L) in enumerate(zip(rows, bags)):
        W, idx, g = make_inputs(R, D, L, B, seed=300 + t)
        Ws.append(W), idxs.append(dev(torch, idx)), gs.append(dev(torch, g))
    a = [dev(torch, w) for w in Ws]
    ffemb.EmbeddingCollection(a, bags).backward_sgd(idxs, gs, 0.01)
    b = [dev(torch, w) for w in Ws]
    coll = ffemb.EmbeddingCollection(b, bags)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    coll.backward_prepare(idxs, stream=side)
    torch.cuda.current_stream().wait_stream(side)
    coll.backward_sgd(idxs, gs, 0.01, prepared=True)
    torch.cuda.synchronize()
    for x, y in zip(a, b):
        assert torch.equal(x, y)


def test_step_captured_in_a_cuda_graph_replays_bit_exact(ffemb, oracle, torch_cuda):
    """The entry points are capturable (no synchronisation, no allocation, fork/join side stream):
    a forward + backward-with-SGD step captured once and replayed on new inputs gives the oracle's
    pooled rows and tables bit-identical to the eager calls', step after step (sorted path, small
    bag-major path and tiny tables together; the between-hook stands for the caller's own kernels)."""
    torch = torch_cuda
    rows, bags, D, B = [5, 100000, 33, 20000, 4, 977], [1, 3, 20, 2, 7, 9], 64, 3000
    Ws = [make_inputs(R, D, L, B, seed=700 + t)[0] for t, (R, L) in enumerate(zip(rows, bags))]
    a = [dev(torch, w) for w in Ws]
    b = [dev(torch, w) for w in Ws]
    eager = ffemb.EmbeddingCollection(a, bags)
    sg = ffemb.StepGraph(ffemb.EmbeddingCollection(b, bags), B, 0.01,
                         between=lambda s: [g.copy_(o).mul_(0.5) for g, o in zip(s.out_grads, s.outs)])
    for x, y in zip(a, b):
        assert torch.equal(x, y)  # capturing (and the lr = 0 warm-up step) changed no table
    for step in range(3):
        idxs = [make_inputs(R, D, L, B, seed=710 + 10 * step + t)[1] for t, (R, L) in enumerate(zip(rows, bags))]
        d_idx = [dev(torch, i) for i in idxs]
        outs = eager.forward(d_idx)
        eager.backward_sgd(d_idx, [o * 0.5 for o in outs], 0.01)
        for dst, src in zip(sg.idx, d_idx):
            dst.copy_(src)
        got = sg.replay()
        torch.cuda.synchronize()
        for t in range(len(rows)):
            assert torch.equal(got[t], outs[t]), (step, t)
            assert torch.equal(a[t], b[t]), (step, t)
            if step == 0:
                assert np.array_equal(got[t].cpu().numpy(), oracle.forward(idxs[t], Ws[t])), t


def test_host_step_matches_device_step(ffemb, oracle, torch_cuda):
    """HostStep (pinned host buffers, three streams) == the device-resident calls, bit for bit, and
    the pooled outputs that land on the host are the oracle's."""
    torch = torch_cuda
    rows, bags, D, B = [5, 50000, 33, 977], [1, 3, 20, 9], 64, 6000
    Ws, idxs, gs = [], [], []
    for t, (R, L) in enumerate(zip(rows, bags)):
        W, idx, g = make_inputs(R, D, L, B, seed=400 + t)
        Ws.append(W), idxs.append(idx), gs.append(g)
    a = [dev(torch, w) for w in Ws]
    ca = ffemb.EmbeddingCollection(a, bags)
    ca.backward_sgd([dev(torch, i) for i in idxs], [dev(torch, g) for g in gs], 0.01)
    b = [dev(torch, w) for w in Ws]
    hs = ffemb.HostStep(ffemb.EmbeddingCollection(b, bags), B)
    h_idx = [torch.from_numpy(i).pin_memory() for i in idxs]
    h_g = [torch.from_numpy(g).pin_memory() for g in gs]
    h_out = [torch.empty((B, D)).pin_memory() for _ in rows]
    hs.step(h_idx, h_g, h_out, 0.01)
    for t in range(len(rows)):
        assert np.array_equal(h_out[t].numpy(), oracle.forward(idxs[t], Ws[t])), t
        assert torch.equal(a[t], b[t]), t


def test_host_steps_in_flight_together_match_sequential_steps(ffemb, oracle, torch_cuda):
    """HostStep.submit / wait: six steps with different batches, two in flight at any time (device
    buffers double buffered) == the same six steps one after the other on the device, bit for bit;
    every step's pooled outputs are those of the table state it ran on.  Steps 2 and 3 replay the
    kernels from CUDA graphs (captured on the second use of a buffer set), steps 4 and 5 change the
    learning rate and run eagerly again."""
    torch = torch_cuda
    rows, bags, D, B, K = [7, 30000, 977], [2, 3, 9], 64, 5000, 6
    lrs = [0.01, 0.01, 0.01, 0.01, 0.02, 0.02]
    Ws, idxs, gs = [], [], []
    for t, (R, L) in enumerate(zip(rows, bags)):
        W = make_inputs(R, D, L, B, seed=500 + t)[0]
        Ws.append(W)
        idxs.append([make_inputs(R, D, L, B, seed=600 + 10 * k + t)[1] for k in range(K)])
        gs.append([make_inputs(R, D, L, B, seed=700 + 10 * k + t)[2] for k in range(K)])
    a = [dev(torch, w) for w in Ws]
    ca = ffemb.EmbeddingCollection(a, bags)
    seq_out = []
    for k in range(K):
        di = [dev(torch, idxs[t][k]) for t in range(len(rows))]
        seq_out.append([o.cpu().numpy() for o in ca.forward(di)])
        ca.backward_sgd(di, [dev(torch, gs[t][k]) for t in range(len(rows))], lrs[k])
    b = [dev(torch, w) for w in Ws]
    hs = ffemb.HostStep(ffemb.EmbeddingCollection(b, bags), B)
    assert hs.depth == 2 and hs.use_graph
    h_idx = [[torch.from_numpy(idxs[t][k]).pin_memory() for t in range(len(rows))] for k in range(K)]
    h_g = [[torch.from_numpy(gs[t][k]).pin_memory() for t in range(len(rows))] for k in range(K)]
    h_out = [[torch.empty((B, D)).pin_memory() for _ in rows] for k in range(K)]
    tickets = [hs.submit(h_idx[k], h_g[k], h_out[k], lrs[k]) for k in range(K)]
    for k in tickets:
        hs.wait(k)
    assert hs.graph_replays == 4  # two graphs per step, steps 2 and 3
    for k in range(K):
        for t in range(len(rows)):
            assert np.array_equal(h_out[k][t].numpy(), seq_out[k][t]), (k, t)
    for t in range(len(rows)):
        assert torch.equal(a[t], b[t]), t
    assert np.array_equal(seq_out[0][1], oracle.forward(idxs[1][0], Ws[1]))


def test_host_step_single_copy_layout(ffemb, oracle, torch_cuda):
    """HostStep.host_buffers(): pinned host tensors laid out like the device slabs move with ONE copy
    per kind (configs 1/2 are copy-launch-bound); same results as per-table host tensors."""
    torch = torch_cuda
    rows, bags, D, B = [7, 30000, 977], [2, 3, 9], 16, 1001   # odd sizes: views are padded to 16 bytes
    Ws, idxs, gs = [], [], []
    for t, (R, L) in enumerate(zip(rows, bags)):
        W, idx, g = make_inputs(R, D, L, B, seed=1200 + t)
        Ws.append(W), idxs.append(idx), gs.append(g)
    a = [dev(torch, w) for w in Ws]
    ffemb.EmbeddingCollection(a, bags).backward_sgd([dev(torch, i) for i in idxs], [dev(torch, g) for g in gs], 0.01)
    b = [dev(torch, w) for w in Ws]
    hs = ffemb.HostStep(ffemb.EmbeddingCollection(b, bags), B)
    hi, hg, ho = hs.host_buffers()
    assert all(x.is_pinned() for x in hi + hg + ho)
    assert hs._span(hi, hs.slab_idx[0]) is not None and hs._span(hg, hs.slab_g[0]) is not None
    assert hs._copy(hs.d_idx[0], hs.slab_idx[0], hi, None) == 1                      # one copy for all tables
    assert hs._copy(hs.d_idx[0], hs.slab_idx[0], [torch.from_numpy(i).pin_memory() for i in idxs], None) == 3
    for t in range(len(rows)):
        hi[t].copy_(torch.from_numpy(idxs[t]))
        hg[t].copy_(torch.from_numpy(gs[t]))
    hs.step(hi, hg, ho, 0.01)
    for t in range(len(rows)):
        assert np.array_equal(ho[t].numpy(), oracle.forward(idxs[t], Ws[t])), t
        assert torch.equal(a[t], b[t]), t


def test_linearity_and_roundtrip_at_full_size(ffemb, torch_cuda):
    """Size-independent properties at BASELINE config-2 scale (26 tables, D=16, B=16384) without
    the CPU oracle: (1) forward of bag-1 indices is a pure gather; (2) with lr = 1 and
    outgrad = forward output on distinct rows, the update zeroes exactly those rows
    (W - 1*W == 0) and leaves the checksum of the others unchanged."""
    torch = torch_cuda
    wl = ffemb.synth.config(2)
    g = torch.Generator(device="cuda").manual_seed(1234)
    Ws = [torch.rand((r, wl.dim), device="cuda", generator=g) + 0.5 for r in wl.rows]
    idx = [torch.from_numpy(ffemb.synth.gen_indices(wl, t)).cuda() for t in range(wl.num_tables)]
    coll = ffemb.EmbeddingCollection(Ws, wl.bags)
    outs = coll.forward(idx)
    for t in range(wl.num_tables):
        assert torch.equal(outs[t], Ws[t][idx[t][:, 0]])
    # distinct rows only: make the batch a permutation prefix per table where possible
    big = [t for t in range(wl.num_tables) if wl.rows[t] >= wl.batch]
    Wb = [Ws[t] for t in big]
    idxb = [torch.randperm(wl.rows[t], device="cuda", generator=g)[:wl.batch].view(-1, 1).contiguous() for t in big]
    sums_before = [w.double().sum().item() for w in Wb]
    collb = ffemb.EmbeddingCollection(Wb, [1] * len(big))
    outb = collb.forward(idxb)
    gathered = [o.double().sum().item() for o in outb]
    collb.backward_sgd(idxb, outb, 1.0)
    torch.cuda.synchronize()
    for k in range(len(big)):
        assert torch.count_nonzero(Wb[k][idxb[k][:, 0]]).item() == 0
        assert abs(Wb[k].double().sum().item() - (sums_before[k] - gathered[k])) <= 1e-6 * sums_before[k]


# ---------------------------------------------------------------------------------------------
# the drop-in boundary: the reference's Legion task frames driving OUR task bodies
# ---------------------------------------------------------------------------------------------
def _have_dropin():
    from oracle import oracle as O
    return O.ref_gpu_available() and O.dropin_gpu_available()


@pytest.mark.skipif(not _have_dropin(), reason="oracle/_ref not built (needs /root/reference at build time)")
@pytest.mark.parametrize("R,D,L,B,aggr", [(1000, 64, 1, 2048, 21), (5000, 128, 7, 700, 21), (300, 16, 5, 999, 22),
                                          (40, 18, 3, 50, 21), (100000, 128, 20, 513, 22)])
def test_dropin_task_bodies_match_reference_task_bodies(oracle, torch_cuda, R, D, L, B, aggr):
    """Embedding::forward_task / backward_task of csrc/flexflow/embedding_ffemb.cu, linked in place
    of the reference's embedding.cu behind the same (Task, PhysicalRegion) frames and the
    reference's own TensorAccessor objects, against the reference's task bodies run here."""
    torch = torch_cuda
    ref, ours = oracle.ref_gpu(), oracle.dropin_gpu()
    W, idx, g = make_inputs(R, D, L, B, seed=R + L)
    dW, didx, dg = dev(torch, W), dev(torch, idx), dev(torch, g)
    ws = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
    outs = []
    for lib in (ref, ours):
        out = torch.full((B, D), float("nan"), device="cuda")
        lib.ffref_gpu_forward_task(didx.data_ptr(), out.data_ptr(), dW.data_ptr(), L, D, B, R, aggr)
        torch.cuda.synchronize()
        outs.append(out.cpu().numpy())
    assert np.array_equal(outs[0], outs[1])  # bit-exact, AVG-in-loop quirk included
    cnt = np.bincount(idx.reshape(-1), minlength=R)
    exact = oracle.backward_dense_f64(idx, g, R, aggr)
    tol = REL_TOL * np.maximum(np.abs(exact), 1.0) * np.maximum(1, cnt)[:, None]
    grads = {}
    for name, lib, wsb in (("ref", ref, 0), ("ours-atomic", ours, 0), ("ours-sorted", ours, ws.numel())):
        lib.ffref_gpu_set_workspace(ws.data_ptr() if wsb else None, wsb)
        G = torch.zeros((R, D), device="cuda")  # Op::zero_grad ran first
        lib.ffref_gpu_backward_task(didx.data_ptr(), dg.data_ptr(), G.data_ptr(), L, D, B, R, aggr)
        torch.cuda.synchronize()
        grads[name] = G.cpu().numpy()
        assert np.all(np.abs(grads[name] - exact) <= tol), name
    single = cnt <= 1
    assert np.array_equal(grads["ref"][single], grads["ours-atomic"][single])
    if D % 4 == 0:
        assert np.array_equal(grads["ref"][single], grads["ours-sorted"][single])
        # fused leaf task == zero-fill -> backward_task -> ps_update_task of the reference
        Wref = dev(torch, W)
        ref.ffref_gpu_sgd_task(dev(torch, grads["ref"]).data_ptr(), Wref.data_ptr(), R, D, 0.01, 0.0)
        Wours = dev(torch, W)
        ours.ffref_gpu_set_workspace(ws.data_ptr(), ws.numel())
        ours.ffdropin_gpu_backward_sgd_task(didx.data_ptr(), dg.data_ptr(), Wours.data_ptr(), L, D, B, R, aggr, 0.01)
        torch.cuda.synchronize()
        a, b = Wours.cpu().numpy(), Wref.cpu().numpy()
        assert np.array_equal(a[single], b[single])
        assert np.all(np.abs(a - b) <= 2 * table_tol(W) * np.maximum(1, np.sqrt(cnt))[:, None])


@pytest.mark.skipif(not _have_dropin(), reason="oracle/_ref not built (needs /root/reference at build time)")
@pytest.mark.parametrize("aggr", [21, 22])
@pytest.mark.parametrize("prepared", [0, 1])
def test_dropin_group_leaf_tasks_match_reference_per_table_tasks(oracle, torch_cuda, aggr, prepared):
    """The per-GPU batched leaf tasks of csrc/flexflow/embedding_group_ffemb.cu (ONE Legion task listing
    the regions of all tables of a GPU: forward; index sort; backward + SGD on the tables themselves)
    driven through the (Task, PhysicalRegion) frames and the reference's own TensorAccessor objects,
    against the reference's per-table tasks run one by one: forward_task (bit-exact), then
    ZeroInitializer -> backward_task -> SGDOptimizer::ps_update_task per table."""
    import ctypes as C
    torch = torch_cuda
    ref, ours = oracle.ref_gpu(), oracle.dropin_gpu()
    shapes = [(1000, 1), (50000, 7), (7, 20), (300, 5), (120, 12)]   # sorted, tiny and small paths in one group
    D, B = 128, 2049
    Ws, idxs, gs = [], [], []
    for i, (R, L) in enumerate(shapes):
        W, idx, g = make_inputs(R, D, L, B, seed=31 * i + R)
        Ws.append(W); idxs.append(idx); gs.append(g)
    n = len(shapes)
    d_idx = [dev(torch, x) for x in idxs]
    d_g = [dev(torch, x) for x in gs]
    d_W = [dev(torch, x) for x in Ws]
    d_out = [torch.full((B, D), float("nan"), device="cuda") for _ in shapes]
    ws = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ours.ffref_gpu_set_workspace(ws.data_ptr(), ws.numel())
    arr = lambda ts: (C.c_void_p * n)(*[t.data_ptr() for t in ts])
    Ls = (C.c_int64 * n)(*[l for _, l in shapes])
    Rs = (C.c_int64 * n)(*[r for r, _ in shapes])
    ours.ffdropin_gpu_group_task(0, n, arr(d_idx), arr(d_out), arr(d_W), Ls, D, B, Rs, aggr, 0.0, 0)
    torch.cuda.synchronize()
    for i, (R, L) in enumerate(shapes):
        want = torch.full((B, D), float("nan"), device="cuda")
        ref.ffref_gpu_forward_task(d_idx[i].data_ptr(), want.data_ptr(), d_W[i].data_ptr(), L, D, B, R, aggr)
        torch.cuda.synchronize()
        assert torch.equal(d_out[i], want), ("group forward", i)
    if prepared:
        ours.ffdropin_gpu_group_task(1, n, arr(d_idx), arr(d_out), arr(d_W), Ls, D, B, Rs, aggr, 0.0, 0)
    ours.ffdropin_gpu_group_task(2, n, arr(d_idx), arr(d_g), arr(d_W), Ls, D, B, Rs, aggr, 0.01, prepared)
    torch.cuda.synchronize()
    for i, (R, L) in enumerate(shapes):
        G = torch.zeros((R, D), device="cuda")  # Op::zero_grad
        ref.ffref_gpu_backward_task(d_idx[i].data_ptr(), d_g[i].data_ptr(), G.data_ptr(), L, D, B, R, aggr)
        Wref = dev(torch, Ws[i])
        ref.ffref_gpu_sgd_task(G.data_ptr(), Wref.data_ptr(), R, D, 0.01, 0.0)
        torch.cuda.synchronize()
        a, b = d_W[i].cpu().numpy(), Wref.cpu().numpy()
        cnt = np.bincount(idxs[i].reshape(-1), minlength=R)
        assert np.array_equal(a[cnt <= 1], b[cnt <= 1]), ("group backward: single-lookup / untouched rows", i)
        assert np.all(np.abs(a - b) <= 2 * table_tol(Ws[i]) * np.maximum(1, np.sqrt(cnt))[:, None]), ("group backward", i)


# ---------------------------------------------------------------------------------------------
# sparse-input delivery: DataLoader::load_sparse_input, all tables in one launch
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("where", ["device", "zero_copy_host"])
def test_load_sparse_inputs_vs_oracle_ragged(ffemb, oracle, torch_cuda, where):
    """Ragged bag widths, gaps between column blocks, tables listed out of column order, explicit
    sample ids — bit-exact against the restated host loop of dlrm.cu:38-45."""
    torch = torch_cuda
    rng = np.random.default_rng(7)
    N, B = 5000, 1777
    bags, cols, row_len = [3, 1, 20, 1, 7], [40, 0, 5, 39, 26], 50   # columns 1..4, 25, 33..38, 43.. unused
    full = rng.integers(0, 1 << 40, (N, row_len), dtype=np.int64)
    tfull = torch.from_numpy(full).cuda() if where == "device" else torch.from_numpy(full).pin_memory()
    outs = ffemb.load_sparse_inputs(tfull, bags, B, first_sample=1234, cols=cols)
    torch.cuda.synchronize()
    for t in range(len(bags)):
        assert np.array_equal(outs[t].cpu().numpy(), oracle.load_sparse_input(full, cols[t], bags[t], B, 1234)), t
    ids = rng.permutation(N)[:B].astype(np.int32)
    outs = ffemb.load_sparse_inputs(tfull, bags, B, sample_ids=torch.from_numpy(ids).cuda(), cols=cols)
    torch.cuda.synchronize()
    for t in range(len(bags)):
        assert np.array_equal(outs[t].cpu().numpy(),
                              oracle.load_sparse_input(full, cols[t], bags[t], B, sample_ids=ids)), t
    assert ffemb.load_sparse_inputs(tfull, bags, 0, cols=cols)[0].shape == (0, 3)   # empty batch


def test_load_sparse_inputs_vs_reference_task(ffemb, oracle, torch_cuda):
    """Against the reference's own DataLoader::load_sparse_input (examples/cpp/DLRM/dlrm.cu, compiled
    unmodified into oracle/_ref/libffref_gpu.so) run once per table, as DataLoader::next_batch does."""
    if not oracle.ref_gpu_available() or not hasattr(oracle.ref_gpu(), "ffref_gpu_load_sparse_input"):
        pytest.skip("oracle/_ref/libffref_gpu.so without the data-loader task")
    torch = torch_cuda
    rng = np.random.default_rng(11)
    N, T, in_dim, B, first = 9000, 26, 4, 4096, 4000
    full = rng.integers(0, 10_000_000, (N, T * in_dim), dtype=np.int64)
    pinned = torch.from_numpy(full).pin_memory()
    ours = ffemb.load_sparse_inputs(pinned, [in_dim] * T, B, first_sample=first)
    torch.cuda.synchronize()
    for t in range(T):
        ref = torch.empty((B, in_dim), dtype=torch.int64, device="cuda")
        rc = oracle.ref_gpu().ffref_gpu_load_sparse_input(pinned.data_ptr(), N, T, in_dim, t, first, B, ref.data_ptr())
        assert rc == 0
        torch.cuda.synchronize()
        assert torch.equal(ours[t], ref), t


def test_load_sparse_inputs_errors(ffemb, torch_cuda):
    torch = torch_cuda
    full = torch.zeros((10, 8), dtype=torch.int64, device="cuda")
    with pytest.raises(ffemb.FFEmbError):   # overlapping column blocks
        ffemb.load_sparse_inputs(full, [4, 4], 5, cols=[0, 2])
    with pytest.raises(ffemb.FFEmbError):   # block past the end of a sample row
        ffemb.load_sparse_inputs(full, [4, 4], 5, cols=[0, 6])
    with pytest.raises(ValueError):          # pageable host memory is not device-accessible
        ffemb.load_sparse_inputs(torch.zeros((10, 8), dtype=torch.int64), [4], 5)


# ---------------------------------------------------------------------------------------------
# Concat-direct: pooled rows into, output-grads out of, the Concat tensors in place
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("aggr", [21, 22])
@pytest.mark.parametrize("D,B", [(64, 3000), (128, 9000), (16, 777)])
def test_concat_direct_matches_separate_buffers_and_oracle_layout(ffemb, oracle, torch_cuda, D, B, aggr):
    """forward_concat == oracle.concat_forward of the per-table pooled outputs (the layout of
    src/ops/concat.cu:211-251, the dense bottom-MLP block in front); backward_sgd_concat on the Concat
    gradient == backward_sgd on the column blocks copied out (what add_with_stride into zeroed
    regions produces, concat.cu:325-366): bit-identical tables.  Mix of tiny / small / sorted paths."""
    torch = torch_cuda
    rows, bags = [5, 100000, 33, 20000, 977], [1, 3, 20, 2, 9]
    T = len(rows)
    Ws, idxs = [], []
    for t, (R, L) in enumerate(zip(rows, bags)):
        W, idx, _ = make_inputs(R, D, L, B, seed=800 + t)
        Ws.append(W), idxs.append(idx)
    rng = np.random.default_rng(5)
    dense = rng.uniform(-1, 1, (B, D)).astype(np.float32)          # x(b): bottom-MLP output, column block 0
    stride = (1 + T) * D + 8                                        # + padding columns nobody owns
    cols = [(1 + t) * D for t in range(T)]
    dx = rng.uniform(-1, 1, (B, stride)).astype(np.float32)
    didx = [dev(torch, i) for i in idxs]
    # separate buffers (the reference's structure)
    a = [dev(torch, w) for w in Ws]
    ca = ffemb.EmbeddingCollection(a, bags, aggr)
    outs = ca.forward(didx)
    ca.backward_sgd(didx, [dev(torch, dx[:, c:c + D]) for c in cols], 0.01)
    # Concat-direct
    b = [dev(torch, w) for w in Ws]
    cb = ffemb.EmbeddingCollection(b, bags, aggr)
    x = torch.full((B, stride), 7.0, device="cuda")
    x[:, :D] = dev(torch, dense)
    cb.forward_concat(didx, x, cols)
    cb.backward_sgd_concat(didx, dev(torch, dx), 0.01, cols)
    torch.cuda.synchronize()
    want = oracle.concat_forward([dense] + [o.cpu().numpy() for o in outs])
    got = x.cpu().numpy()
    assert np.array_equal(got[:, :(1 + T) * D], want)
    assert np.all(got[:, (1 + T) * D:] == 7.0)                      # padding untouched
    assert np.array_equal(outs[1].cpu().numpy(), oracle.forward(idxs[1], Ws[1], aggr))
    for t in range(T):
        assert torch.equal(a[t], b[t]), t


def test_concat_direct_errors(ffemb, torch_cuda):
    torch = torch_cuda
    W = [torch.zeros((10, 16), device="cuda")]
    coll = ffemb.EmbeddingCollection(W, [1])
    idx = [torch.zeros((8, 1), dtype=torch.int64, device="cuda")]
    dx = torch.zeros((8, 32), device="cuda")
    with pytest.raises(ffemb.FFEmbError):
        coll.backward_sgd_concat(idx, dx, 0.01, cols=[20])      # block past the end of a row
    with pytest.raises(ffemb.FFEmbError):
        coll.backward_sgd_concat(idx, dx, 0.01, cols=[2])       # 16-byte fragments must stay aligned
    with pytest.raises(ffemb.FFEmbError):
        coll.forward_concat(idx, dx, cols=[20])


# ---------------------------------------------------------------------------------------------
# full sgd_update fused with the backward: weight decay / momentum / nesterov
# ---------------------------------------------------------------------------------------------
SGDX_CASES = [  # weight_decay, momentum, nesterov
    (1e-3, 0.0, False), (0.0, 0.9, False), (1e-3, 0.9, False), (5e-4, 0.8, True), (0.0, 0.0, False)]


@pytest.mark.parametrize("wd,mom,nesterov", SGDX_CASES)
@pytest.mark.parametrize("aggr", [21, 22])
def test_backward_sgdx_two_steps_vs_reference_optimizer(ffemb, oracle, torch_cuda, wd, mom, nesterov, aggr):
    """Two consecutive steps (the second sees the first one's momentum) against the oracle's
    zero-fill -> embed_backward -> sgd_update restatement, and — where oracle/_ref is built — against
    the reference's own SGDOptimizer::ps_update_task fed with the reference's own embed_backward:
    rows with at most one lookup and rows never touched are BIT-IDENTICAL (same FFMA contraction);
    repeated rows within 1e-6 (summation order of the gradient).  The touched bitmap returns zero."""
    torch = torch_cuda
    D, B, lr = 64, 3000, 0.05
    rows, bags = [5, 60000, 977, 33], [1, 2, 7, 20]
    T = len(rows)
    Ws = [make_inputs(R, D, L, B, seed=900 + t)[0] for t, (R, L) in enumerate(zip(rows, bags))]
    steps = [[make_inputs(R, D, L, B, seed=910 + 10 * k + t)[1:] for t, (R, L) in enumerate(zip(rows, bags))]
             for k in range(2)]
    dW = [dev(torch, w) for w in Ws]
    dV = [torch.zeros_like(w) for w in dW]
    coll = ffemb.EmbeddingCollection(dW, bags, aggr)
    oW = [w.copy() for w in Ws]
    oV = [np.zeros_like(w) for w in Ws]
    have_ref = oracle.ref_gpu_available() and hasattr(oracle.ref_gpu(), "ffref_gpu_sgd_task_full")
    rW = [dev(torch, w) for w in Ws]
    rV = [torch.zeros_like(w) for w in rW]
    multi = [np.zeros(R, bool) for R in rows]
    for k in range(2):
        idxs = [steps[k][t][0] for t in range(T)]
        gs = [steps[k][t][1] for t in range(T)]
        coll.backward_sgdx([dev(torch, i) for i in idxs], [dev(torch, g) for g in gs], lr, wd, mom, nesterov,
                           V=dV if mom > 0 else None)
        for t in range(T):
            grad = oracle.backward_dense(idxs[t], gs[t], rows[t], aggr)
            oW[t], oV[t] = oracle.sgd_update_full(oW[t], grad, oV[t], lr, wd, mom, nesterov)
            multi[t] |= np.bincount(idxs[t].reshape(-1), minlength=rows[t]) > 1
            if have_ref:
                rg = torch.zeros_like(rW[t])
                oracle.ref_gpu().ffref_gpu_backward_task(dev(torch, idxs[t]).data_ptr(), dev(torch, gs[t]).data_ptr(),
                                                         rg.data_ptr(), bags[t], D, B, rows[t], aggr)
                oracle.ref_gpu().ffref_gpu_sgd_task_full(rg.data_ptr(), rW[t].data_ptr(), rV[t].data_ptr(), D, rows[t],
                                                         lr, wd, mom, int(nesterov))
        torch.cuda.synchronize()
    for t in range(T):
        got, gotV = dW[t].cpu().numpy(), dV[t].cpu().numpy()
        once = ~multi[t]
        assert np.array_equal(got[once], oW[t][once]), t
        tol = REL_TOL * max(np.abs(oW[t]).max(), 1e-30) * 4
        assert np.all(np.abs(got - oW[t]) <= tol), t
        if mom > 0:
            assert np.array_equal(gotV[once], oV[t][once]), t
            assert np.all(np.abs(gotV - oV[t]) <= REL_TOL * np.abs(oV[t]).max() * 4), t
        if have_ref:
            ref = rW[t].cpu().numpy()
            assert np.array_equal(got[once], ref[once]), t
            assert np.all(np.abs(got - ref) <= tol), t
        assert int(coll._touched[t].abs().sum().item()) == 0, t


def test_backward_sgdx_plain_settings_match_the_fused_sgd_path(ffemb, torch_cuda):
    """weight_decay = momentum = 0: the general path == ffemb_backward_sgd_batched on single-lookup
    rows bit for bit (untouched rows get W -= lr*0, a bit-identical no-op)."""
    torch = torch_cuda
    D, B = 128, 2000
    rows, bags = [50000, 7], [1, 3]
    Ws, idxs, gs = [], [], []
    for t, (R, L) in enumerate(zip(rows, bags)):
        W, idx, g = make_inputs(R, D, L, B, seed=950 + t)
        Ws.append(W), idxs.append(dev(torch, idx)), gs.append(dev(torch, g))
    a, b = [dev(torch, w) for w in Ws], [dev(torch, w) for w in Ws]
    ffemb.EmbeddingCollection(a, bags).backward_sgd(idxs, gs, 0.01)
    ffemb.EmbeddingCollection(b, bags).backward_sgdx(idxs, gs, 0.01)
    torch.cuda.synchronize()
    cnt = np.bincount(idxs[0].cpu().numpy().reshape(-1), minlength=rows[0])
    assert np.array_equal(a[0].cpu().numpy()[cnt <= 1], b[0].cpu().numpy()[cnt <= 1])
    assert torch.allclose(a[1], b[1], rtol=1e-6, atol=1e-8)


@pytest.mark.parametrize("wd", [0.0, 1e-3])
@pytest.mark.parametrize("aggr", [21, 22])
def test_backward_adam_two_steps_vs_reference_optimizer(ffemb, oracle, torch_cuda, wd, aggr):
    """adam_update fused with the backward: two consecutive steps (alpha_t advanced like
    AdamOptimizer::next) against the oracle's restatement and the reference's own
    AdamOptimizer::ps_update_task fed with the reference's own embed_backward.  Rows with at most one
    lookup per step and rows never touched are BIT-IDENTICAL in W, M and V; repeated rows within 1e-6
    of the step size."""
    torch = torch_cuda
    D, B, alpha, b1, b2, eps = 64, 3000, 0.001, 0.9, 0.999, 1e-8
    rows, bags = [5, 60000, 977], [1, 2, 7]
    T = len(rows)
    Ws = [make_inputs(R, D, L, B, seed=1000 + t)[0] for t, (R, L) in enumerate(zip(rows, bags))]
    steps = [[make_inputs(R, D, L, B, seed=1010 + 10 * k + t)[1:] for t, (R, L) in enumerate(zip(rows, bags))]
             for k in range(2)]
    dW = [dev(torch, w) for w in Ws]
    dM, dV = [torch.zeros_like(w) for w in dW], [torch.zeros_like(w) for w in dW]
    coll = ffemb.EmbeddingCollection(dW, bags, aggr)
    oW, oM, oV = [w.copy() for w in Ws], [np.zeros_like(w) for w in Ws], [np.zeros_like(w) for w in Ws]
    have_ref = oracle.ref_gpu_available() and hasattr(oracle.ref_gpu(), "ffref_gpu_adam_task")
    rW = [dev(torch, w) for w in Ws]
    rM, rV = [torch.zeros_like(w) for w in rW], [torch.zeros_like(w) for w in rW]
    multi = [np.zeros(R, bool) for R in rows]
    for k in range(2):
        a_t = np.float32(oracle.adam_alpha_t(alpha, b1, b2, k + 1))
        idxs = [steps[k][t][0] for t in range(T)]
        gs = [steps[k][t][1] for t in range(T)]
        coll.backward_adam([dev(torch, i) for i in idxs], [dev(torch, g) for g in gs], a_t, dM, dV, b1, b2, wd, eps)
        for t in range(T):
            grad = oracle.backward_dense(idxs[t], gs[t], rows[t], aggr)
            oW[t], oM[t], oV[t] = oracle.adam_update(oW[t], grad, oM[t], oV[t], a_t, b1, b2, wd, eps)
            multi[t] |= np.bincount(idxs[t].reshape(-1), minlength=rows[t]) > 1
            if have_ref:
                rg = torch.zeros_like(rW[t])
                oracle.ref_gpu().ffref_gpu_backward_task(dev(torch, idxs[t]).data_ptr(), dev(torch, gs[t]).data_ptr(),
                                                         rg.data_ptr(), bags[t], D, B, rows[t], aggr)
                oracle.ref_gpu().ffref_gpu_adam_task(rg.data_ptr(), rW[t].data_ptr(), rV[t].data_ptr(), rM[t].data_ptr(),
                                                     D, rows[t], float(a_t), b1, b2, wd, eps)
        torch.cuda.synchronize()
    for t in range(T):
        got, gM, gV = dW[t].cpu().numpy(), dM[t].cpu().numpy(), dV[t].cpu().numpy()
        once = ~multi[t]
        if have_ref:   # the reference's own kernel is the authority for the rounding
            assert np.array_equal(got[once], rW[t].cpu().numpy()[once]), t
            assert np.array_equal(gM[once], rM[t].cpu().numpy()[once]), t
            assert np.array_equal(gV[once], rV[t].cpu().numpy()[once]), t
            # repeated rows: the reference's atomics sum in a different (and varying) order, and Adam's
            # normalised step alpha_t*M/(sqrt(V)+eps) is not Lipschitz where a gradient sum cancels to
            # ~0 (its sign decides a full step): every element stays within two steps, and all but a
            # vanishing fraction within 1e-6-class distance
            dev_ = np.abs(got - rW[t].cpu().numpy())
            assert dev_.max() <= 2 * 2 * alpha * 3.2, t
            assert (dev_ > 1e-5).mean() < 1e-3, t
        assert np.array_equal(got[once], oW[t][once]), t
        assert np.array_equal(gM[once], oM[t][once]) and np.array_equal(gV[once], oV[t][once]), t
        assert np.all(np.abs(gM - oM[t]) <= REL_TOL * max(np.abs(oM[t]).max(), 1e-30) * 4), t
        assert int(coll._touched[t].abs().sum().item()) == 0, t


def test_general_optimizers_on_an_empty_batch_still_update_every_row(ffemb, oracle, torch_cuda):
    """The reference's sgd_update / adam_update run over the whole table whatever the batch holds
    (optimizer_kernel.cu:23-41,206-226): with no lookups at all every row still decays."""
    torch = torch_cuda
    R, D = 1000, 32
    W = make_inputs(R, D, 1, 8, seed=1100)[0]
    idx = [torch.zeros((0, 3), dtype=torch.int64, device="cuda")]
    g = [torch.zeros((0, D), device="cuda")]
    dW, dV = [dev(torch, W)], [torch.full((R, D), 0.25, device="cuda")]
    ffemb.EmbeddingCollection(dW, [3]).backward_sgdx(idx, g, 0.1, 1e-2, 0.9, True, V=dV)
    wantW, wantV = oracle.sgd_update_full(W, np.zeros_like(W), np.full_like(W, 0.25), 0.1, 1e-2, 0.9, True)
    assert np.array_equal(dW[0].cpu().numpy(), wantW) and np.array_equal(dV[0].cpu().numpy(), wantV)
    dW, dM, dV = [dev(torch, W)], [torch.full((R, D), 0.5, device="cuda")], [torch.full((R, D), 0.125, device="cuda")]
    ffemb.EmbeddingCollection(dW, [3]).backward_adam(idx, g, 1e-3, dM, dV, 0.9, 0.999, 1e-2, 1e-8)
    wW, wM, wV = oracle.adam_update(W, np.zeros_like(W), np.full_like(W, 0.5), np.full_like(W, 0.125), 1e-3, 0.9, 0.999, 1e-2, 1e-8)
    assert np.array_equal(dW[0].cpu().numpy(), wW)
    assert np.array_equal(dM[0].cpu().numpy(), wM) and np.array_equal(dV[0].cpu().numpy(), wV)


# ---------------------------------------------------------------------------------------------
# row-wise backward: route + the owners' padded backward, G ranks emulated on ONE GPU
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("G", [2, 8])
@pytest.mark.parametrize("aggr", [21, 22])
def test_rowwise_route_and_padded_backward_emulated_ranks(ffemb, oracle, torch_cuda, G, aggr):
    """Every "rank" h runs ffemb_rowwise_route for ITS batch block into G local owner buffers (all peer
    pointers are plain local memory here), then every "owner" runs ffemb_backward_sgd_padded on its shard:
      * slot arrays == the oracle's stable bucketing by owner (oracle_route_rows: owner = row / ceil(R/G),
        local row = row % ceil(R/G)), bit-exact, -1 everywhere else;
      * routed output-grad rows == the senders' rows on every owner of >= 1 lookup of the bag;
      * shards put back together == the unsharded step: rows with one lookup bit-identical to the
        single-GPU call, all rows within the summation-order bound of the oracle's reference step."""
    import ctypes as C
    torch = torch_cuda
    B_, D = 1001, 64   # ragged last block
    rows, bags = [10007, 300, 50001], [1, 4, 3]
    T, sumL = len(rows), sum(bags)
    ext = (B_ + G - 1) // G
    Bp = G * ext
    rng = np.random.default_rng(5 + G + aggr)
    Ws = [rng.uniform(-0.5, 0.5, (r, D)).astype(np.float32) for r in rows]
    idx = [rng.integers(0, r, (B_, l), dtype=np.int64) for r, l in zip(rows, bags)]
    idx[1][:, 0] = 7                      # one hot row, met by every sample of every rank
    dx = rng.uniform(-1, 1, (B_, T * D)).astype(np.float32)
    d_idx, d_dx = [dev(torch, i) for i in idx], dev(torch, dx)
    slots = [torch.full((Bp * sumL,), -1, dtype=torch.int64, device="cuda") for _ in range(G)]
    grows = [torch.zeros((T * Bp, D), dtype=torch.float32, device="cuda") for _ in range(G)]
    L_ = ffemb.lib()
    for h in range(G):
        lo, hi = oracle.batch_block(B_, G, h)
        rt = ffemb.binding.ffemb_route_t()
        for g in range(G):
            rt.peer_slots[g], rt.peer_rows[g] = slots[g].data_ptr(), grows[g].data_ptr()
        rt.num_peers, rt.rank, rt.samples_per_peer = G, h, ext
        arr = (ffemb.binding.ffemb_table_t * T)()
        keep = [d_idx[t][lo:hi].contiguous() for t in range(T)]
        for t in range(T):
            arr[t].idx, arr[t].R, arr[t].L, arr[t].reserved = keep[t].data_ptr(), rows[t], bags[t], t * D
        mine = d_dx[lo:hi].contiguous()
        ffemb.binding.check(L_.ffemb_rowwise_route(arr, T, D, hi - lo, mine.data_ptr(), T * D, C.byref(rt), None))
        torch.cuda.synchronize()
    # slots and routed rows against the oracle's bucketing
    views, off = [], 0
    for t in range(T):
        views.append([s[off:off + Bp * bags[t]].view(Bp, bags[t]) for s in slots])
        off += Bp * bags[t]
    for t in range(T):
        offsets, perm, local = oracle.route_rows(idx[t].reshape(-1), rows[t], G)
        for g in range(G):
            want = np.full(Bp * bags[t], -1, np.int64)
            want[perm[offsets[g]:offsets[g + 1]]] = local[offsets[g]:offsets[g + 1]]
            assert np.array_equal(views[t][g].cpu().numpy().reshape(-1), want), (t, g, "routed slots")
            bags_here = (want.reshape(Bp, bags[t])[:B_] >= 0).any(axis=1)
            got = grows[g][t * Bp:t * Bp + B_].cpu().numpy()
            assert np.array_equal(got[bags_here], dx[bags_here, t * D:(t + 1) * D]), (t, g, "routed rows")
    # the owners' backward, then put the shards back together
    lr = 0.01
    new = [np.empty_like(w) for w in Ws]
    for g in range(G):
        shards = []
        for t in range(T):
            rps = oracle.rows_per_shard(rows[t], G)
            sh = np.zeros((rps, D), np.float32)
            part = Ws[t][g * rps:(g + 1) * rps]
            sh[:len(part)] = part
            shards.append(dev(torch, sh))
        coll = ffemb.EmbeddingCollection(shards, bags, aggr)
        coll.backward_sgd_padded([views[t][g] for t in range(T)], [grows[g][t * Bp:(t + 1) * Bp] for t in range(T)], lr)
        torch.cuda.synchronize()
        for t in range(T):
            rps = oracle.rows_per_shard(rows[t], G)
            n = len(Ws[t][g * rps:(g + 1) * rps])
            new[t][g * rps:g * rps + n] = shards[t].cpu().numpy()[:n]
    one = [dev(torch, w) for w in Ws]
    ffemb.EmbeddingCollection(one, bags, aggr).backward_sgd(d_idx, [d_dx[:, t * D:(t + 1) * D].contiguous() for t in range(T)], lr)
    torch.cuda.synchronize()
    for t in range(T):
        cnt = np.bincount(idx[t].reshape(-1), minlength=rows[t])
        single = one[t].cpu().numpy()
        assert np.array_equal(new[t][cnt <= 1], single[cnt <= 1]), (t, "single-lookup / untouched rows")
        want = oracle.reference_step(idx[t], dx[:, t * D:(t + 1) * D].copy(), Ws[t].copy(), aggr, lr)
        tol = table_tol(Ws[t]) * np.sqrt(np.maximum(cnt, 1))[:, None] * 2
        assert np.all(np.abs(new[t].astype(np.float64) - want) <= tol), (t, "updated table")


@pytest.mark.parametrize("aggr", [22, 122])
@pytest.mark.parametrize("D", [16, 128, 256])
def test_avg_backward_division_once_per_bag_equals_division_per_lookup(ffemb, oracle, torch_cuda, aggr, D):
    """AGGR_MODE_AVG backward: gradient = output_grad / in_dim (embedding.cu:206-211).  With a workspace
    sized by ffemb_backward_workspace_avg the sorted backward divides once per bag into scratch and sums
    the quotients; with the plain workspace it divides inside the walk.  Same IEEE quotients, same order
    of adds: the updated tables must be bit-identical — denormal, huge and zero gradients included."""
    import ctypes as C
    torch = torch_cuda
    rows, bags, B = [5000, 70000, 900], [9, 1, 20], 3001
    T = len(rows)
    rng = np.random.default_rng(D + aggr)
    Ws = [rng.uniform(-0.5, 0.5, (r, D)).astype(np.float32) for r in rows]
    idx = [rng.integers(0, r, (B, l), dtype=np.int64) for r, l in zip(rows, bags)]
    gs = [rng.uniform(-1, 1, (B, D)).astype(np.float32) for _ in rows]
    gs[0][::7] *= 1e-39   # denormal quotients
    gs[0][3::11] *= 1e37  # huge
    gs[2][5] = 0.0
    n, total = sum(B * l for l in bags), sum(rows)
    L_ = ffemb.lib()
    small = L_.ffemb_backward_workspace(n, total, D)
    big = L_.ffemb_backward_workspace_avg(n, total, D, B, T)
    assert big >= small + T * B * D * 4
    d_idx, d_g = [dev(torch, x) for x in idx], [dev(torch, x) for x in gs]
    res = []
    for nbytes in (small, big):
        dW = [dev(torch, w) for w in Ws]
        ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        arr = (ffemb.binding.ffemb_table_t * T)()
        for t in range(T):
            arr[t].idx, arr[t].out, arr[t].table = d_idx[t].data_ptr(), d_g[t].data_ptr(), dW[t].data_ptr()
            arr[t].R, arr[t].L = rows[t], bags[t]
        ffemb.binding.check(L_.ffemb_backward_sgd_batched(arr, T, D, B, aggr, 0.01, ws.data_ptr(), nbytes, None))
        torch.cuda.synchronize()
        res.append([w.cpu().numpy() for w in dW])
    for t in range(T):
        assert np.array_equal(res[0][t].view(np.uint32), res[1][t].view(np.uint32)), t
        if aggr == 22:
            want = oracle.reference_step(idx[t], gs[t].copy(), Ws[t].copy(), aggr, 0.01)
            cnt = np.bincount(idx[t].reshape(-1), minlength=rows[t])
            assert np.array_equal(res[1][t][cnt <= 1], want[cnt <= 1]), (t, "single-lookup rows vs the reference step")


@pytest.mark.parametrize("aggr", [21, 22])
def test_backward_aligned_key_space_per_table_sorts(ffemb, oracle, torch_cuda, aggr):
    """Few large batches of lookups into tables whose own row bits need fewer radix passes than the global
    key (here 2 x 60000 rows: 16 bits each, 17 together, 4.2 M lookups per table) take the ALIGNED key
    space — table t's keys are (t << 16) + row and every table's stretch is sorted on its own.  Same
    contract as any batched backward: untouched and single-lookup rows bit-identical to the reference
    step, repeated rows within the summation bound."""
    torch = torch_cuda
    rows, bags, D, B = [60000, 60000], [64, 64], 16, 65536
    rng = np.random.default_rng(17 + aggr)
    Ws = [rng.uniform(-0.5, 0.5, (r, D)).astype(np.float32) for r in rows]
    idx = [rng.integers(0, r, (B, l), dtype=np.int64) for r, l in zip(rows, bags)]
    idx[0][:, 0] = 59999                       # a hot last row right below the next table's key range
    idx[1][::3, 1] = 0                         # and a hot first row of the next table
    gs = [rng.uniform(-1, 1, (B, D)).astype(np.float32) for _ in rows]
    d_idx, d_g = [dev(torch, x) for x in idx], [dev(torch, x) for x in gs]
    together = [dev(torch, w) for w in Ws]
    ffemb.EmbeddingCollection(together, bags, aggr).backward_sgd(d_idx, d_g, 0.01)
    torch.cuda.synchronize()
    for t in range(len(rows)):
        got = together[t].cpu().numpy()
        want = oracle.reference_step(idx[t], gs[t].copy(), Ws[t].copy(), aggr, 0.01)
        cnt = np.bincount(idx[t].reshape(-1), minlength=rows[t])
        assert np.array_equal(got[cnt <= 1], want[cnt <= 1]), (t, "single-lookup / untouched rows")
        tol = table_tol(Ws[t]) * np.sqrt(np.maximum(cnt, 1))[:, None] * 2
        assert np.all(np.abs(got.astype(np.float64) - want) <= tol), (t, "updated table")


@pytest.mark.parametrize("aggr,D", [(21, 64), (22, 128), (21, 16)])
def test_backward_hot_rows_under_skew(ffemb, oracle, torch_cuda, aggr, D):
    """Skewed indices through the default path (level 1 + combine + the tree for chains of > 128 partials):
    one row with 630 K lookups (several per bag), a hot LAST row of a table and a hot FIRST row of the next
    one, rows of 7000-9500 lookups.  Untouched and single-lookup rows bit-identical to the reference step,
    repeated rows within the summation bound of the fp64 sum, and two runs on the same inputs bit-identical
    (the reference's atomics are not).  test_backward_hot_rows_opt_in runs the same data through the opt-in
    bag-major hot-row kernels."""
    torch = torch_cuda
    rows, bags, B = [100000, 50000], [32, 32], 65536      # 4.2 M sorted lookups: the hot path's minimum
    rng = np.random.default_rng(3 + aggr + D)
    Ws = [rng.uniform(-0.5, 0.5, (r, D)).astype(np.float32) for r in rows]
    idx = [rng.integers(0, r, (B, l), dtype=np.int64) for r, l in zip(rows, bags)]
    flat0 = idx[0].reshape(-1)
    sel = rng.random(flat0.size)
    flat0[sel < 0.30] = 5                       # one very hot row (~630 K lookups, several per bag)
    flat0[(sel >= 0.30) & (sel < 0.31)] = 99999  # the table's last row, hot (~21 K)
    for k in range(6):                          # rows around the threshold: 7000 .. 9500 lookups
        m = (sel >= 0.32 + 0.006 * k) & (sel < 0.32 + 0.006 * k + (7000 + 500 * k) / flat0.size)
        flat0[m] = 1000 + k
    flat1 = idx[1].reshape(-1)
    flat1[rng.random(flat1.size) < 0.05] = 0    # the next table's first row, hot
    gs = [rng.uniform(-1, 1, (B, D)).astype(np.float32) for _ in rows]
    d_idx, d_g = [dev(torch, x) for x in idx], [dev(torch, x) for x in gs]
    res = []
    for rep in range(2):
        dW = [dev(torch, w) for w in Ws]
        ffemb.EmbeddingCollection(dW, bags, aggr).backward_sgd(d_idx, d_g, 0.01)
        torch.cuda.synchronize()
        res.append([w.cpu().numpy() for w in dW])
    for t in range(len(rows)):
        assert np.array_equal(res[0][t].view(np.uint32), res[1][t].view(np.uint32)), (t, "not reproducible")
        want = oracle.reference_step(idx[t], gs[t].copy(), Ws[t].copy(), aggr, 0.01)
        cnt = np.bincount(idx[t].reshape(-1), minlength=rows[t])
        got = res[0][t]
        assert np.array_equal(got[cnt <= 1], want[cnt <= 1]), (t, "single-lookup / untouched rows")
        exact = Ws[t].astype(np.float64) - 0.01 * oracle.backward_dense_f64(idx[t], gs[t], rows[t], aggr)
        tol = table_tol(Ws[t]) * np.sqrt(np.maximum(cnt, 1))[:, None]
        assert np.all(np.abs(got - exact) <= tol), (t, "updated table vs the fp64 sum")


def test_backward_hot_rows_opt_in(torch_cuda):
    """FFEMB_HOT=1 (a process of its own: the tuning switches are read once): rows with >= 8192 lookups
    leave level 1 and are summed bag-major (hot_select / hot_mark / hot_sum / hot_apply) — same bars."""
    import subprocess
    env = dict(os.environ, FFEMB_HOT="1")
    r = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), "hot_rows_check.py")], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "hot_rows_check ok" in r.stdout, r.stdout[-3000:]
