"""Tuning aid: backward (sort + reduce + SGD) time per table of cfg3 in isolation, per phase."""
import importlib, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("dlrm-flexflow_b200")
wl = pkg.synth.config(int(os.environ.get("WL", "3")))
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(1)
cap = int(os.environ.get("ROWS_CAP", "8000000"))
tot = {}
for t in range(wl.num_tables):
    R, L = min(wl.rows[t], cap), wl.bags[t]
    W = torch.rand((R, wl.dim), device=dev, generator=g)
    idx = [torch.randint(0, R, (wl.batch, L), device=dev, generator=g) for _ in range(3)]
    coll = pkg.EmbeddingCollection([W], [L])
    og = [torch.rand((wl.batch, wl.dim), device=dev, generator=g)]
    for k in range(3): coll.backward_sgd([idx[k]], og, 0.01)
    pkg.profile_enable(True); pkg.profile_collect()
    for k in range(9): coll.backward_sgd([idx[k % 3]], og, 0.01)
    ms, _ = pkg.profile_collect(); pkg.profile_enable(False)
    ms = {k: v / 9 for k, v in ms.items() if v > 0}
    for k, v in ms.items(): tot[k] = tot.get(k, 0) + v
    print("t=%2d R=%9d L=%2d " % (t, R, L), {k: round(v, 4) for k, v in ms.items()})
    del W, idx
print("sum", {k: round(v, 4) for k, v in tot.items()})
