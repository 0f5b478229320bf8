"""Tuning aid: forward time per table of cfg3 in isolation (B=65536, D=128), CUDA events."""
import importlib, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("dlrm-flexflow_b200")
wl = pkg.synth.config(3)
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(1)
cap = int(os.environ.get("ROWS_CAP", "8000000"))
res = []
for t in range(wl.num_tables):
    R, L = min(wl.rows[t], cap), wl.bags[t]
    W = torch.rand((R, wl.dim), device=dev, generator=g)
    idx = [torch.randint(0, R, (wl.batch, L), device=dev, generator=g) for _ in range(3)]
    coll = pkg.EmbeddingCollection([W], [L])
    out = [torch.empty((wl.batch, wl.dim), device=dev)]
    for k in range(3): coll.forward([idx[k]], out)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for k in range(9): coll.forward([idx[k % 3]], out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 9
    n = wl.batch * L
    res.append((t, R, L, ms, n * 512 / ms / 1e9))
    del W, idx
for r in res: print("t=%2d R=%9d L=%2d  %.4f ms  %.2f TB/s(row bytes)" % (r[0], r[1], r[2], r[3], r[4] / 1e3))
print("sum ms", sum(r[3] for r in res))
