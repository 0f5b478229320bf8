"""Tuning aid (GPU): cost of every cfg3 table in isolation — forward, sort, level 1, upper levels, small
path — at full width and at half width (a column-split part), B = 65536.  Feeds sharding.part_cost."""
import importlib, json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("dlrm-flexflow_b200")
wl = pkg.synth.config(3)
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(1)
cap = int(os.environ.get("ROWS_CAP", "8000000"))
res = []
for t in range(wl.num_tables):
    for D in (wl.dim, wl.dim // 2):
        R, L = min(wl.rows[t], cap), wl.bags[t]
        W = torch.rand((R, D), device=dev, generator=g)
        idx = [torch.randint(0, R, (wl.batch, L), device=dev, generator=g) for _ in range(3)]
        coll = pkg.EmbeddingCollection([W], [L])
        out = [torch.empty((wl.batch, D), device=dev)]
        og = [torch.rand((wl.batch, D), device=dev, generator=g)]
        for k in range(3):
            coll.forward([idx[k]], out); coll.backward_sgd([idx[k]], og, 0.01)
        pkg.profile_enable(True); pkg.profile_collect()
        for k in range(9):
            coll.forward([idx[k % 3]], out); coll.backward_sgd([idx[k % 3]], og, 0.01)
        ms, _ = pkg.profile_collect(); pkg.profile_enable(False)
        ms = {k: round(v / 9, 4) for k, v in ms.items() if v > 0}
        res.append({"t": t, "R": wl.rows[t], "L": L, "D": D, **ms})
        print(res[-1], flush=True)
        del W, idx, coll
json.dump(res, open("gpurun_out/r2_per_table_cost.json", "w"), indent=0)
