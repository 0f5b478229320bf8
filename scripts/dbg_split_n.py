import importlib, os, sys, torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("dlrm-flexflow_b200")
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
g = torch.Generator(device=dev).manual_seed(7)
rows = [5, 100000, 33, 20000, 4, 1, 977, 3000000, 14]
bags = [1, 3, 20, 2, 7, 1, 20, 4, 16]
T, D, B = len(rows), 128, 20001
Ws = [torch.rand((r, D), device=dev, generator=g) - 0.5 for r in rows]
idx = [torch.randint(0, r, (B, l), device=dev, generator=g) for r, l in zip(rows, bags)]
S = pkg.sharding
lo, hi = S.batch_block(B, world, rank)
want = torch.cat(pkg.EmbeddingCollection(Ws, bags, 21).forward(idx), dim=1)[lo:hi]
parts = []
for t in range(T):
    if t in (1, 7):
        parts += [S.Part(t, 0, D // 2, t % world), S.Part(t, D // 2, D // 2, (t + 1) % world)]
    else:
        parts.append(S.Part(t, 0, D, t % world))
for gk in range(world):
    mine_g = [p for p in parts if p.rank == gk]
    widths = sorted({p.width for p in mine_g}, reverse=True)
    for i, p in enumerate(mine_g):
        p.group = widths.index(p.width) if len(widths) > 1 else i % 2
fx = pkg.exchange.FusedTableWiseExchange(T, None, B, D, rank, world, dev, parts=parts)
K, mine = fx.K, fx.mine
Wp = [[Ws[p.table][:, p.col0:p.col0 + p.width].contiguous() for p in mine[k]] for k in range(K)]
colls = [pkg.EmbeddingCollection(Wp[k], [bags[p.table] for p in mine[k]], 21) if mine[k] else None for k in range(K)]
idxg = [[idx[p.table] for p in mine[k]] for k in range(K)]
x = fx.forward(colls, idxg)[:hi - lo]
torch.cuda.synchronize(); dist.barrier()
for r in range(world):
    if r == rank:
        print("rank", rank, "K", K, [[(p.table, p.col0, p.width) for p in m] for m in mine], flush=True)
        for p in fx.parts:
            c0 = p.table * D + p.col0
            a, b = x[:, c0:c0 + p.width], want[:, c0:c0 + p.width]
            bad = (a != b)
            print("  part", (p.table, p.col0, p.width, p.rank, p.group), "equal", torch.equal(a, b), "bad rows", bad.any(dim=1).sum().item(),
                  "first bad row", (bad.any(dim=1).nonzero()[0].item() if bad.any() else -1), "zeros", (a == 0).all(dim=1).sum().item(), flush=True)
    dist.barrier()
dist.destroy_process_group()
