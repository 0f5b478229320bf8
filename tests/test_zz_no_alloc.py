"""ffemb_backward_sgd_static: the fused backward + SGD with its scratch in a file-scope device array of
the library (no cudaMalloc anywhere in the call) against ffemb_backward_sgd with a caller's workspace.
No counterpart in the reference.

Collected last on purpose; passes alone in a fresh process and after
the other tests — the static array keeps whatever the previous call left in it, which is what the
second call of the GPU test runs into.
"""
import ctypes as C

import numpy as np
import pytest

REL_TOL = 1e-6  # the bar of tests/test_gpu_parity.py for updated tables


def _step_both_ways(ffemb, torch, R, D, L, B, aggr, lr, seed):
    """One step through each entry point from the same table; -> (static, workspace) tables on the host."""
    L_ = ffemb.binding.lib()
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, R, (B, L), dtype=np.int64)
    idx[: B // 8] = idx[0]  # rows that repeat a lot: the segmented sum has real work
    W = rng.standard_normal((R, D)).astype(np.float32)
    g = rng.standard_normal((B, D)).astype(np.float32)
    d_idx, d_g = torch.from_numpy(idx).cuda(), torch.from_numpy(g).cuda()
    Wa, Wb = torch.from_numpy(W).cuda(), torch.from_numpy(W).cuda()
    need = int(L_.ffemb_backward_workspace(B * L, R, D))
    assert 0 < need <= int(L_.ffemb_static_scratch_bytes())
    ws = torch.empty(need + 256, dtype=torch.uint8, device="cuda")
    ws_ptr = (ws.data_ptr() + 255) & ~255
    stream = torch.cuda.current_stream().cuda_stream
    ffemb.binding.check(L_.ffemb_backward_sgd_static(d_idx.data_ptr(), d_g.data_ptr(), Wa.data_ptr(), L, D, B, R,
                                                     aggr, lr, stream))
    ffemb.binding.check(L_.ffemb_backward_sgd(d_idx.data_ptr(), d_g.data_ptr(), Wb.data_ptr(), L, D, B, R, aggr, lr,
                                              ws_ptr, need, stream))
    torch.cuda.synchronize()
    return Wa.cpu().numpy(), Wb.cpu().numpy(), W


@pytest.mark.gpu
def test_static_scratch_backward_matches_the_workspace_entry_point_twice_in_a_row(ffemb):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    torch.cuda.set_device(0)
    # two calls in a row on the shared array, the second smaller than the first (it finds the first
    # call's sort buffers and flags behind its own span), then the first shape again
    shapes = [(200_000, 64, 8, 8192, 21), (5_000, 128, 3, 1000, 22), (200_000, 64, 8, 8192, 21)]
    for k, (R, D, L, B, aggr) in enumerate(shapes):
        got, want, W0 = _step_both_ways(ffemb, torch, R, D, L, B, aggr, 0.05, seed=100 + k)
        assert np.isfinite(got).all()
        assert not np.array_equal(want, W0), "the step must have changed the table"
        tol = REL_TOL * np.maximum(np.abs(want), np.abs(want).max())
        err = np.abs(got.astype(np.float64) - want)
        assert np.all(err <= tol), f"call {k}: worst error {err.max():.3e}"
        untouched = np.all(want == W0, axis=1)
        assert np.array_equal(got[untouched].view(np.uint32), W0[untouched].view(np.uint32))


def test_static_scratch_refuses_sizes_beyond_the_array(ffemb):
    """Host-side argument check only (returns before any CUDA call): runs without a GPU."""
    L_ = ffemb.binding.lib()
    cap = int(L_.ffemb_static_scratch_bytes())
    assert cap == 64 << 20
    B, L, D, R = 1 << 20, 64, 128, 10_000_000
    assert int(L_.ffemb_backward_workspace(B * L, R, D)) > cap
    rc = L_.ffemb_backward_sgd_static(None, None, None, L, D, B, R, 21, C.c_float(0.1), None)
    assert rc == ffemb.binding.FFEMB_EWORKSPACE
    msg = L_.ffemb_last_error().decode()
    assert "static" in msg and str(cap) in msg
    with pytest.raises(ffemb.binding.FFEmbError):
        ffemb.binding.check(rc)
    # nothing to do: no scratch needed, no CUDA call made
    assert L_.ffemb_backward_sgd_static(None, None, None, 4, 64, 0, 100, 21, C.c_float(0.1), None) == 0
