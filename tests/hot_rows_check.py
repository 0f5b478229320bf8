"""The opt-in hot-row path of the sorted backward (FFEMB_HOT=1; csrc/ffemb_backward.cu: hot_select / hot_mark /
hot_sum / hot_apply) on skewed indices, against the oracle.  Run as a script in a process of its own — the
tuning switches are read once per process — by tests/test_gpu_parity.py::test_backward_hot_rows_opt_in."""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ffemb = importlib.import_module("dlrm-flexflow_b200")
from oracle import oracle  # noqa: E402


def main():
    assert os.environ.get("FFEMB_HOT") == "1"
    for aggr, D in ((21, 64), (22, 128), (21, 16)):
        rows, bags, B = [100000, 50000], [32, 32], 65536      # 4.2 M sorted lookups: the hot path's minimum
        rng = np.random.default_rng(3 + aggr + D)
        Ws = [rng.uniform(-0.5, 0.5, (r, D)).astype(np.float32) for r in rows]
        idx = [rng.integers(0, r, (B, l), dtype=np.int64) for r, l in zip(rows, bags)]
        flat0 = idx[0].reshape(-1)
        sel = rng.random(flat0.size)
        flat0[sel < 0.30] = 5                        # one very hot row (~630 K lookups, several per bag)
        flat0[(sel >= 0.30) & (sel < 0.31)] = 99999  # the table's last row, hot (~21 K)
        for k in range(6):                           # rows around the threshold: 7000 .. 9500 lookups
            m = (sel >= 0.32 + 0.006 * k) & (sel < 0.32 + 0.006 * k + (7000 + 500 * k) / flat0.size)
            flat0[m] = 1000 + k
        flat1 = idx[1].reshape(-1)
        flat1[rng.random(flat1.size) < 0.05] = 0     # the next table's first row, hot
        gs = [rng.uniform(-1, 1, (B, D)).astype(np.float32) for _ in rows]
        dev = lambda x: torch.from_numpy(np.ascontiguousarray(x)).cuda()
        d_idx, d_g = [dev(x) for x in idx], [dev(x) for x in gs]
        res = []
        for rep in range(2):
            dW = [dev(w) for w in Ws]
            ffemb.EmbeddingCollection(dW, bags, aggr).backward_sgd(d_idx, d_g, 0.01)
            torch.cuda.synchronize()
            res.append([w.cpu().numpy() for w in dW])
        for t in range(len(rows)):
            assert np.array_equal(res[0][t].view(np.uint32), res[1][t].view(np.uint32)), (aggr, D, t, "not reproducible")
            want = oracle.reference_step(idx[t], gs[t].copy(), Ws[t].copy(), aggr, 0.01)
            cnt = np.bincount(idx[t].reshape(-1), minlength=rows[t])
            got = res[0][t]
            assert np.array_equal(got[cnt <= 1], want[cnt <= 1]), (aggr, D, t, "single-lookup / untouched rows")
            exact = Ws[t].astype(np.float64) - 0.01 * oracle.backward_dense_f64(idx[t], gs[t], rows[t], aggr)
            tol = 1e-6 * np.abs(Ws[t]).max() * np.sqrt(np.maximum(cnt, 1))[:, None]
            assert np.all(np.abs(got - exact) <= tol), (aggr, D, t, "updated table vs the fp64 sum")
    print("hot_rows_check ok")


if __name__ == "__main__":
    main()
