"""Generate golden vectors from the REFERENCE'S OWN CODE (oracle/_ref, compiled unmodified
from /root/reference by oracle/Makefile).  The reference holds no fixtures for the embedding
op, so these runs are the pin.

  python tests/golden/make_golden.py cpu   # here: src/ops/embedding.cc through the shim
  python tests/golden/make_golden.py gpu   # on a B200 (gpurun): src/ops/embedding.cu,
                                           # ZeroInitializer::init_task, SGDOptimizer::ps_update_task

Inputs are regenerated from the seeds stored in each case, so the .npz files hold only the
case parameters and the reference outputs.  Output: tests/golden/ref_cpu.npz / ref_gpu.npz
(gpu mode writes to gpurun_out/ref_gpu.npz when run on the box; copy it here and commit).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

# (name, R, D, L, B, aggr, seed)
CPU_CASES = [
    ("task_d64", 5000, 64, 1, 257, 21, 11),     # forward_task_cpu/backward_task_cpu: D=64 unrolled path
    ("task_d128", 3000, 128, 1, 130, 21, 12),   # D=128 unrolled path
    ("task_d32", 1000, 32, 1, 64, 21, 13),      # D=32 unrolled path
    ("task_d16", 999, 16, 1, 300, 21, 14),      # generic path (cfg2's D)
    ("task_d20", 50, 20, 1, 77, 21, 15),        # generic path with a scalar tail
    ("task_dups", 7, 64, 1, 200, 21, 16),       # heavy duplicates (backward accumulates)
    ("bag_sum_d128", 4000, 128, 20, 65, 21, 17),   # EmbeddingLookup direct, fixed bags
    ("bag_sum_d16", 4000, 16, 7, 129, 21, 18),
    ("bag_mean_d64", 4000, 64, 5, 100, 122, 19),   # normalize_by_lengths -> true mean
    ("bag_mean_d48", 300, 48, 64, 33, 122, 20),
]
GPU_CASES = [
    ("sum_l1_d64", 5000, 64, 1, 257, 21, 31),
    ("sum_l1_d16", 999, 16, 1, 1000, 21, 32),
    ("sum_l20_d128", 4000, 128, 20, 65, 21, 33),
    ("avg_l4_d64", 100, 64, 4, 50, 22, 34),        # AVG-in-loop quirk
    ("avg_l64_d128", 1000, 128, 64, 40, 22, 35),
    ("avg_l3_d20", 64, 20, 3, 31, 22, 36),
    ("sum_dups_d16", 4, 16, 1, 512, 21, 37),       # R=4 like the smallest Kaggle tables
    ("sum_l7_d100", 321, 100, 7, 45, 21, 38),
]


def make_inputs(R, D, L, B, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    a = np.sqrt(1.0 / R)
    W = rng.uniform(-a, a, (R, D)).astype(np.float32)
    idx = rng.integers(0, R, (B, L), dtype=np.int64)
    g = rng.uniform(-1, 1, (B, D)).astype(np.float32)
    return W, idx, g


def gen_cpu():
    out = {}
    for name, R, D, L, B, aggr, seed in CPU_CASES:
        W, idx, g = make_inputs(R, D, L, B, seed)
        out[name + ".params"] = np.array([R, D, L, B, aggr, seed], np.int64)
        if name.startswith("task_"):
            out[name + ".fwd"] = O.ref_cpu_forward_task(idx, W)
            out[name + ".wgrad"] = O.ref_cpu_backward_task(idx, g, np.zeros((R, D), np.float32))
        else:
            out[name + ".fwd"] = O.ref_cpu_lookup(idx, W, normalize=(aggr == 122))
    path = os.path.join(ROOT, "tests", "golden", "ref_cpu.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


def gen_gpu():
    import torch
    L_ = O.ref_gpu()
    dev = torch.device("cuda:0")
    out = {}
    for name, R, D, L, B, aggr, seed in GPU_CASES:
        W, idx, g = make_inputs(R, D, L, B, seed)
        dW, didx, dg = (torch.from_numpy(x).to(dev) for x in (W, idx, g))
        dout = torch.full((B, D), float("nan"), device=dev)
        L_.ffref_gpu_forward_task(didx.data_ptr(), dout.data_ptr(), dW.data_ptr(), L, D, B, R, aggr)
        dgrad = torch.full((R, D), float("nan"), device=dev)
        L_.ffref_gpu_zero_task(dgrad.data_ptr(), R, D)
        L_.ffref_gpu_backward_task(didx.data_ptr(), dg.data_ptr(), dgrad.data_ptr(), L, D, B, R, aggr)
        torch.cuda.synchronize()
        wgrad = dgrad.cpu().numpy()
        L_.ffref_gpu_sgd_task(dgrad.data_ptr(), dW.data_ptr(), R, D, 0.01, 0.0)
        torch.cuda.synchronize()
        out[name + ".params"] = np.array([R, D, L, B, aggr, seed], np.int64)
        out[name + ".fwd"] = dout.cpu().numpy()
        out[name + ".wgrad"] = wgrad
        out[name + ".wnew"] = dW.cpu().numpy()
    d = os.path.join(ROOT, "gpurun_out")
    os.makedirs(d, exist_ok=True)
    path = os.path.join(d, "ref_gpu.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    {"cpu": gen_cpu, "gpu": gen_gpu}[sys.argv[1]]()
