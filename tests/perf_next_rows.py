"""Measurements for the SURVEY 8(f) rows (run under gpurun, one GPU; lives under tests/ because it
times the reference's own code from oracle/_ref next to ours — only tests/ may execute oracle/):
  f1  Concat-direct forward/backward vs separate [B][D] buffers + Concat's strided copy / add passes
  f2  sparse-input delivery: ffemb_load_sparse_inputs from zero-copy host memory and from device
      memory, next to the reference's own DataLoader::load_sparse_input (oracle/_ref) per table
  f3  full sgd_update (weight decay + momentum) fused with the backward vs the reference's own
      pipeline on the same B200: ZeroInitializer::init_task + Embedding::backward_task +
      SGDOptimizer::ps_update_task (oracle/_ref/libffref_gpu.so)
CUDA events on the current stream, warm-up 3, median of 10.  Prints one JSON line per row."""
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("dlrm-flexflow_b200")
from oracle import oracle as O  # noqa: E402  (the reference arm of each comparison)


def timed(fn, warm=3, reps=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return float(np.median(ms))


def row_f1(wl_no=3, rows_cap=4_000_000):
    wl = pkg.synth.config(wl_no)
    rows = [min(r, rows_cap) for r in wl.rows]
    T, D, B = wl.num_tables, wl.dim, wl.batch
    g = torch.Generator(device="cuda").manual_seed(1)
    W = [torch.empty((r, D), device="cuda").uniform_(-0.01, 0.01, generator=g) for r in rows]
    idx = [torch.randint(0, r, (B, l), device="cuda", generator=g) for r, l in zip(rows, wl.bags)]
    coll = pkg.EmbeddingCollection(W, wl.bags, wl.aggr)
    stride = (1 + T) * D
    cols = [(1 + t) * D for t in range(T)]
    x = torch.zeros((B, stride), device="cuda")
    dx = torch.empty((B, stride), device="cuda").uniform_(-1, 1, generator=g)
    outs = [torch.empty((B, D), device="cuda") for _ in range(T)]
    ogs = [torch.zeros((B, D), device="cuda") for _ in range(T)]
    xv, dxv = x.view(B, 1 + T, D), dx.view(B, 1 + T, D)

    def separate():  # the reference's structure: per-table buffers + Concat's per-input passes
        coll.forward(idx, outs)
        for t in range(T):
            xv[:, 1 + t, :].copy_(outs[t])          # copy_with_stride, concat.cu:242-248
        for t in range(T):
            ogs[t].zero_()                           # Op::zero_grad on the output-grad region
            ogs[t].add_(dxv[:, 1 + t, :])            # add_with_stride, concat.cu:356-360
        coll.backward_sgd(idx, ogs, 0.0)

    def direct():
        coll.forward_concat(idx, x, cols)
        coll.backward_sgd_concat(idx, dx, 0.0, cols)

    a, b = timed(separate), timed(direct)
    return {"row": "f1 concat-direct", "workload": wl.name + f"-rowscap{rows_cap}", "separate_buffers_ms": round(a, 3),
            "concat_direct_ms": round(b, 3), "speedup": round(a / b, 3),
            "bytes_not_moved": int(4 * T * B * D * 4 + T * B * D * 4)}


def row_f2(T=26, in_dim=4, B=65536, N=200_000):
    rng = np.random.default_rng(0)
    full = torch.from_numpy(rng.integers(0, 10_000_000, (N, T * in_dim), dtype=np.int64))
    pinned = full.pin_memory()
    devfull = full.cuda()
    outs = [torch.empty((B, in_dim), dtype=torch.int64, device="cuda") for _ in range(T)]
    bytes_ = B * T * in_dim * 8
    zc = timed(lambda: pkg.load_sparse_inputs(pinned, [in_dim] * T, B, first_sample=1000, outs=outs))
    dv = timed(lambda: pkg.load_sparse_inputs(devfull, [in_dim] * T, B, first_sample=1000, outs=outs))
    res = {"row": "f2 sparse-input delivery", "tables": T, "in_dim": in_dim, "batch": B, "bytes": bytes_,
           "ours_zero_copy_host_ms": round(zc, 3), "ours_zero_copy_host_gbs": round(bytes_ / zc / 1e6, 1),
           "ours_device_ms": round(dv, 3), "ours_device_gbs": round(bytes_ / dv / 1e6, 1)}
    if O.ref_gpu_available() and hasattr(O.ref_gpu(), "ffref_gpu_load_sparse_input"):
        import time
        L = O.ref_gpu()

        def ref():
            for t in range(T):
                L.ffref_gpu_load_sparse_input(pinned.data_ptr(), N, T, in_dim, t, 1000, B, outs[t].data_ptr())
        ref()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            ref()      # host loop + cudaMemcpy + cudaDeviceSynchronize per table: wall clock is the cost
        torch.cuda.synchronize()
        rm = (time.perf_counter() - t0) / 3 * 1e3
        res.update({"reference_26_tasks_ms": round(rm, 3), "speedup_vs_reference": round(rm / zc, 1)})
    return res


def row_f3(T=8, R=1_000_000, D=64, B=2048, L=1):
    g = torch.Generator(device="cuda").manual_seed(2)
    W = [torch.empty((R, D), device="cuda").uniform_(-0.001, 0.001, generator=g) for _ in range(T)]
    V = [torch.zeros_like(w) for w in W]
    idx = [torch.randint(0, R, (B, L), device="cuda", generator=g) for _ in range(T)]
    og = [torch.empty((B, D), device="cuda").uniform_(-1, 1, generator=g) for _ in range(T)]
    coll = pkg.EmbeddingCollection(W, [L] * T, 21)
    ours = timed(lambda: coll.backward_sgdx(idx, og, 0.01, 1e-4, 0.9, False, V=V))
    res = {"row": "f3 full sgd_update fused", "workload": f"cfg1 shape: {T} x {R} x {D}, batch {B}, wd 1e-4, momentum 0.9",
           "ours_ms": round(ours, 3), "table_bytes": T * R * D * 4,
           "ours_gbs": round(16 * T * R * D / ours / 1e6, 1)}
    if O.ref_gpu_available() and hasattr(O.ref_gpu(), "ffref_gpu_sgd_task_full"):
        Lr = O.ref_gpu()
        G = [torch.zeros_like(w) for w in W]

        def ref():
            for t in range(T):
                Lr.ffref_gpu_zero_task(G[t].data_ptr(), D, R)
                Lr.ffref_gpu_backward_task(idx[t].data_ptr(), og[t].data_ptr(), G[t].data_ptr(), L, D, B, R, 21)
                Lr.ffref_gpu_sgd_task_full(G[t].data_ptr(), W[t].data_ptr(), V[t].data_ptr(), D, R, 0.01, 1e-4, 0.9, 0)
        rm = timed(ref)
        res.update({"reference_kernels_on_b200_ms": round(rm, 3), "speedup_vs_reference_kernels": round(rm / ours, 2)})
    return res


def row_f3_adam(T=8, R=1_000_000, D=64, B=2048, L=1):
    g = torch.Generator(device="cuda").manual_seed(3)
    W = [torch.empty((R, D), device="cuda").uniform_(-0.001, 0.001, generator=g) for _ in range(T)]
    M, V = [torch.zeros_like(w) for w in W], [torch.zeros_like(w) for w in W]
    idx = [torch.randint(0, R, (B, L), device="cuda", generator=g) for _ in range(T)]
    og = [torch.empty((B, D), device="cuda").uniform_(-1, 1, generator=g) for _ in range(T)]
    coll = pkg.EmbeddingCollection(W, [L] * T, 21)
    ours = timed(lambda: coll.backward_adam(idx, og, 1e-3, M, V, 0.9, 0.999, 1e-4, 1e-8))
    res = {"row": "f3 adam_update fused", "workload": f"cfg1 shape: {T} x {R} x {D}, batch {B}, weight decay 1e-4",
           "ours_ms": round(ours, 3), "ours_gbs": round(24 * T * R * D / ours / 1e6, 1)}
    if O.ref_gpu_available() and hasattr(O.ref_gpu(), "ffref_gpu_adam_task"):
        Lr = O.ref_gpu()
        G = [torch.zeros_like(w) for w in W]

        def ref():
            for t in range(T):
                Lr.ffref_gpu_zero_task(G[t].data_ptr(), D, R)
                Lr.ffref_gpu_backward_task(idx[t].data_ptr(), og[t].data_ptr(), G[t].data_ptr(), L, D, B, R, 21)
                Lr.ffref_gpu_adam_task(G[t].data_ptr(), W[t].data_ptr(), V[t].data_ptr(), M[t].data_ptr(), D, R,
                                       1e-3, 0.9, 0.999, 1e-4, 1e-8)
        rm = timed(ref)
        res.update({"reference_kernels_on_b200_ms": round(rm, 3), "speedup_vs_reference_kernels": round(rm / ours, 2)})
    return res


if __name__ == "__main__":
    torch.cuda.set_device(0)
    for fn in (row_f1, row_f2, row_f3, row_f3_adam):
        print(json.dumps(fn()), flush=True)
