import importlib, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("dlrm-flexflow_b200")
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(7)
rows = [5, 100000, 33, 20000, 4, 1, 977, 3000000, 14]
bags = [1, 3, 20, 2, 7, 1, 20, 4, 16]
T, D, B = len(rows), 128, 20001
Ws = [torch.rand((r, D), device=dev, generator=g) - 0.5 for r in rows]
idx = [torch.randint(0, r, (B, l), device=dev, generator=g) for r, l in zip(rows, bags)]
full = pkg.EmbeddingCollection(Ws, bags, 21)
want = torch.cat(full.forward(idx), dim=1)
for tabs in ([1, 7], [1], [7]):
    for c0 in (0, 64):
        x = torch.zeros((B, T * D), device=dev)
        Wp = [Ws[t][:, c0:c0 + 64].contiguous() for t in tabs]
        coll = pkg.EmbeddingCollection(Wp, [bags[t] for t in tabs], 21)
        coll.forward_concat([idx[t] for t in tabs], x, cols=[t * D + c0 for t in tabs])
        torch.cuda.synchronize()
        for t in tabs:
            a, b = x[:, t * D + c0:t * D + c0 + 64], want[:, t * D + c0:t * D + c0 + 64]
            print(tabs, c0, t, torch.equal(a, b), (a != b).sum().item(), (a - b).abs().max().item())
        plain = coll.forward([idx[t] for t in tabs])
        for i, t in enumerate(tabs):
            print(' plain', t, torch.equal(plain[i], want[:, t * D + c0:t * D + c0 + 64]))
