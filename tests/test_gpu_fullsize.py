"""Full-size parity (GPU): BASELINE configs 1, 2 and one table of config 5 at their REAL sizes against
the reference's own GPU kernels (oracle/_ref/libffref_gpu.so = /root/reference/src/ops/embedding.cu,
src/runtime/initializer_kernel.cu, src/runtime/optimizer_kernel.cu compiled unmodified for sm_100a):

    embed_forward                         src/ops/embedding.cu:166-190
    zero-fill -> embed_backward (atomics) src/runtime/model.cc:466-490, src/ops/embedding.cu:192-217
    -> sgd_update over every row          src/runtime/optimizer_kernel.cu:23-41

Bars:
  * pooled outputs BIT-EXACT (SUM and AVG as the reference computes it, divide inside the bag loop);
  * rows no lookup touches and rows with exactly one lookup BIT-IDENTICAL to the reference step;
  * repeated rows: the reference's atomics add in no fixed order, so its own result moves from run to
    run; both results are therefore held against the exact (fp64) sum of the same addends with the
    backward-error bound of a floating-point sum, elementwise,
        |W - W_exact| <= 1e-6 * (|W0| + lr * sum_i |g_i|)                      (the stated norm)
    (1e-6 relative to the magnitude of the operands of the update — the north star's "1e-6 relative
    error, which covers summation-order differences").  The reference's own run-to-run spread on the
    hottest row is recorded beside our error in gpurun_out/r2_parity_fullsize.json.
config 5's hottest row collects ~1.4 M addends of the 16.8 M lookups of a table (Zipf alpha = 1.05).
"""
import json
import os

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TOL = 1e-6


@pytest.fixture(scope="module")
def env(ffemb):
    import torch
    from oracle import oracle as O
    assert torch.cuda.is_available()
    if not O.ref_gpu_available():
        pytest.skip("oracle/_ref/libffref_gpu.so not built (needs /root/reference at build time)")
    torch.cuda.set_device(0)
    return torch, O.ref_gpu()


def _record(name, rec):
    d = os.path.join(ROOT, "gpurun_out")
    try:
        os.makedirs(d, exist_ok=True)
        p = os.path.join(d, "r2_parity_fullsize.json")
        cur = json.load(open(p)) if os.path.exists(p) else {}
        cur[name] = rec
        json.dump(cur, open(p, "w"), indent=1)
    except Exception:
        pass
    print(name, json.dumps(rec))


def _ref_step(torch, ref, idx, og, W, L, D, B, aggr, lr):
    """zero-fill -> backward_kernel -> ps_update_task of the reference on a copy of W."""
    R = int(W.shape[0])
    Wr, Wg = W.clone(), torch.empty_like(W)
    ref.ffref_gpu_zero_task(Wg.data_ptr(), R, D)
    ref.ffref_gpu_backward_kernel(idx.data_ptr(), og.data_ptr(), Wg.data_ptr(), L, D, B, aggr, None)
    ref.ffref_gpu_sgd_task(Wg.data_ptr(), Wr.data_ptr(), R, D, float(lr), 0.0)
    torch.cuda.synchronize()
    return Wr


def _exact(torch, idx, og, W0, L, aggr, lr):
    """fp64: W0 - lr * sum of the addends, and lr * sum of their magnitudes (the bound's scale)."""
    R, D = W0.shape
    g = og.double()
    if aggr != 21:
        g = (og / float(L)).double()  # each addend is the fp32 quotient the kernels add (embedding.cu:210)
    acc = torch.zeros((R, D), dtype=torch.float64, device=W0.device)
    mag = torch.zeros((R, D), dtype=torch.float64, device=W0.device)
    ga = g.abs()
    for j in range(L):
        acc.index_add_(0, idx[:, j], g)
        mag.index_add_(0, idx[:, j], ga)
    acc.mul_(-float(lr)).add_(W0.double())
    return acc, mag.mul_(float(lr))


def _check_table(torch, name, W_ours, W_ref, W_ref2, W0, idx, og, L, aggr, lr):
    R = int(W0.shape[0])
    cnt = torch.bincount(idx.reshape(-1), minlength=R)
    single = cnt <= 1
    assert torch.equal(W_ours[single], W_ref[single]), f"{name}: untouched / single-lookup rows differ"
    exact, mag = _exact(torch, idx, og, W0, L, aggr, lr)
    bound = mag.add_(W0.double().abs()).mul_(TOL).clamp_min_(1e-300)
    err_ours = (W_ours.double() - exact).abs_()
    err_ref = (W_ref.double() - exact).abs_()
    del exact
    worst = (err_ours / bound).max().item()
    assert worst <= 1.0, f"{name}: |W - W_exact| exceeds 1e-6 * (|W0| + lr*sum|g|) by {worst:.3f}x"
    hot = int(torch.argmax(cnt).item())
    rec = {"rows": R, "max_lookups_on_one_row": int(cnt[hot].item()),
           "rows_repeated": int((cnt > 1).sum().item()),
           "ours_worst_err_over_bound": worst,
           "ref_worst_err_over_bound": (err_ref / bound).max().item(),
           "hottest_row_abs_err_ours": err_ours[hot].max().item(),
           "hottest_row_abs_err_ref": err_ref[hot].max().item(),
           "hottest_row_ref_run_to_run_spread": (W_ref[hot] - W_ref2[hot]).abs().max().item(),
           "ref_run_to_run_max_spread_any_row": (W_ref - W_ref2).abs().max().item()}
    return rec


@pytest.mark.parametrize("cfg", [1, 2])
def test_config_at_full_size_vs_reference_gpu_kernels(ffemb, env, cfg):
    torch, ref = env
    wl = ffemb.synth.config(cfg)
    D, B, lr = wl.dim, wl.batch, wl.lr
    Ws = [torch.from_numpy(ffemb.synth.gen_weights(r, D, t)).cuda() for t, r in enumerate(wl.rows)]
    idx = [torch.from_numpy(ffemb.synth.gen_indices(wl, t)).cuda() for t in range(wl.num_tables)]
    og = [torch.from_numpy(ffemb.synth.gen_outgrad(wl, t)).cuda() for t in range(wl.num_tables)]
    W0 = [w.clone() for w in Ws]
    coll = ffemb.EmbeddingCollection(Ws, wl.bags, wl.aggr)
    outs = coll.forward(idx)
    coll.backward_sgd(idx, og, lr)
    torch.cuda.synchronize()
    recs = {}
    for t in range(wl.num_tables):
        L = wl.bags[t]
        out_ref = torch.full((B, D), float("nan"), device="cuda")
        ref.ffref_gpu_forward_kernel(idx[t].data_ptr(), out_ref.data_ptr(), W0[t].data_ptr(), L, D, B, wl.aggr, None)
        torch.cuda.synchronize()
        assert torch.equal(outs[t], out_ref), f"cfg{cfg} table {t}: pooled output differs"
        Wr = _ref_step(torch, ref, idx[t], og[t], W0[t], L, D, B, wl.aggr, lr)
        Wr2 = _ref_step(torch, ref, idx[t], og[t], W0[t], L, D, B, wl.aggr, lr)
        recs[f"table{t}"] = _check_table(torch, f"cfg{cfg} table {t}", Ws[t], Wr, Wr2, W0[t], idx[t], og[t], L,
                                         wl.aggr, lr)
    worst = max(recs.values(), key=lambda r: r["max_lookups_on_one_row"])
    _record(f"cfg{cfg}", {"tables": wl.num_tables, "batch": B, "dim": D, "most_repeated_table": worst,
                          "ours_worst_err_over_bound": max(r["ours_worst_err_over_bound"] for r in recs.values()),
                          "ref_worst_err_over_bound": max(r["ref_worst_err_over_bound"] for r in recs.values())})


@pytest.mark.parametrize("aggr", [22, 122])
def test_config5_one_table_at_full_size_vs_reference_gpu_kernels(ffemb, env, aggr):
    """T = 1 of config 5: 10 M rows x 128, bag 64, batch 262144, Zipf(1.05): 16.8 M lookups, the hottest
    row takes ~8 % of them.  aggr 22 = AVG as the reference computes it (its kernels are the oracle for
    forward and step); 122 = the opt-in true mean: the forward is held against the exact mean within
    1e-6 relative (there is no reference GPU kernel for it; its CPU definition is checked at small sizes
    in test_gpu_parity.py), the backward is the same g / L scatter and is checked like AVG."""
    torch, ref = env
    wl = ffemb.synth.config(5)
    t, D, B, L, lr = 0, wl.dim, wl.batch, wl.bags[0], wl.lr
    R = wl.rows[0]
    a = (1.0 / R) ** 0.5  # examples/cpp/DLRM/dlrm.cc:44-45; generated on the device (5 GB)
    W = torch.empty((R, D), device="cuda").uniform_(-a, a, generator=torch.Generator(device="cuda").manual_seed(1234))
    idx = torch.from_numpy(ffemb.synth.gen_indices(wl, t)).cuda()
    og = torch.from_numpy(ffemb.synth.gen_outgrad(wl, t)).cuda()
    W0 = W.clone()
    coll = ffemb.EmbeddingCollection([W], [L], aggr)
    out = coll.forward([idx])[0]
    coll.backward_sgd([idx], [og], lr)
    torch.cuda.synchronize()
    if aggr == 22:
        out_ref = torch.full((B, D), float("nan"), device="cuda")
        ref.ffref_gpu_forward_kernel(idx.data_ptr(), out_ref.data_ptr(), W0.data_ptr(), L, D, B, 22, None)
        torch.cuda.synchronize()
        assert torch.equal(out, out_ref), "cfg5: pooled output (AVG, divide-in-loop) differs"
    else:
        mean = torch.zeros((B, D), dtype=torch.float64, device="cuda")
        for j in range(L):
            mean += W0[idx[:, j]].double()
        mean /= L
        scale = torch.zeros((B, D), dtype=torch.float64, device="cuda")
        for j in range(L):
            scale += W0[idx[:, j]].double().abs()
        assert ((out.double() - mean).abs() <= TOL * (scale / L) + 1e-30).all(), "cfg5: true mean off"
        del mean, scale
    Wr = _ref_step(torch, ref, idx, og, W0, L, D, B, 22, lr)
    Wr2 = _ref_step(torch, ref, idx, og, W0, L, D, B, 22, lr)
    rec = _check_table(torch, f"cfg5 aggr {aggr}", W, Wr, Wr2, W0, idx, og, L, 22, lr)
    # determinism of our path at this size: a second run from the same state gives the same bits
    W2 = W0.clone()
    coll2 = ffemb.EmbeddingCollection([W2], [L], aggr)
    coll2.backward_sgd([idx], [og], lr)
    torch.cuda.synchronize()
    assert torch.equal(W, W2), "cfg5: our update is not bit-reproducible run to run"
    rec["ours_bit_reproducible"] = True
    _record(f"cfg5_T1_aggr{aggr}", rec)
