"""Multi-GPU parity of the table-wise exchange (run under torchrun, one rank per GPU):
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/multigpu_check.py
Every rank also holds ALL tables and computes the single-GPU result locally; the sharded step
(fused peer-memory exchange and the NCCL all-to-all form) must reproduce it:
  * interaction input x_h [B_h][T*D] bit-exact (pooled rows are routed, not recomputed),
  * output-grads on the owners bit-exact copies of the dx column blocks,
  * owned tables after backward+SGD bit-identical to a single-GPU call on the same group of tables
    (same kernels, same summation tree) and within 1e-6-class tolerance of the all-tables call
    (the tree's tile boundaries move with the set of tables batched together).
Invoked by tests/test_multigpu.py (pytest -m gpu, skipped with fewer than 2 GPUs)."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("dlrm-flexflow_b200")


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rows = [5, 100000, 33, 20000, 4, 1, 977, 3000000, 14]
    bags = [1, 3, 20, 2, 7, 1, 20, 4, 16]
    T, D, B = len(rows), 128, 20001  # ragged last block: ceil(B / world) split
    for aggr in (21, 22):
        g = torch.Generator(device=dev).manual_seed(7)  # same inputs on every rank
        Ws = [torch.rand((r, D), device=dev, generator=g) - 0.5 for r in rows]
        idx = [torch.randint(0, r, (B, l), device=dev, generator=g) for r, l in zip(rows, bags)]
        dx_full = torch.rand((B, T * D), device=dev, generator=g) - 0.5
        # single-GPU truth
        ref_W = [w.clone() for w in Ws]
        full = pkg.EmbeddingCollection(ref_W, bags, aggr)
        outs = full.forward(idx)
        x_full = torch.cat(outs, dim=1)
        og = [dx_full[:, t * D:(t + 1) * D].contiguous() for t in range(T)]
        full.backward_sgd(idx, og, 0.01)
        owner = [t % world for t in range(T)]
        owned = [t for t in range(T) if owner[t] == rank]
        lo, hi = pkg.sharding.batch_block(B, world, rank)
        for kind in ("fused", "nccl"):
            myW = [Ws[t].clone() for t in owned]
            coll = pkg.EmbeddingCollection(myW, [bags[t] for t in owned], aggr) if owned else None
            myidx = [idx[t] for t in owned]
            if kind == "fused":
                fx = pkg.exchange.FusedTableWiseExchange(T, owner, B, D, rank, world, dev)
                fx.dx.copy_(dx_full[lo:hi])
                x = fx.forward(coll, myidx)[:hi - lo]
                outgrads = fx.backward()
            else:
                ex = pkg.exchange.TableWiseExchange(T, owner, B, D, rank, world, dev)
                po = coll.forward(myidx) if coll else []
                x = ex.forward(po)
                ex.dx.copy_(dx_full[lo:hi])
                outgrads = ex.backward([torch.empty((B, D), device=dev) for _ in owned])
            torch.cuda.synchronize()
            assert torch.equal(x, x_full[lo:hi]), (kind, aggr, "interaction input differs")
            for i, t in enumerate(owned):
                assert torch.equal(outgrads[i], og[t]), (kind, aggr, t, "routed output-grad differs")
            if coll:
                coll.backward_sgd(myidx, outgrads, 0.01)
                same_group = [Ws[t].clone() for t in owned]
                pkg.EmbeddingCollection(same_group, [bags[t] for t in owned], aggr).backward_sgd(
                    myidx, [og[t] for t in owned], 0.01)
                torch.cuda.synchronize()
                for i, t in enumerate(owned):
                    assert torch.equal(myW[i], same_group[i]), (kind, aggr, t, "updated table differs")
                    dup = B * bags[t] / rows[t] + 1
                    assert torch.allclose(myW[i], ref_W[t], rtol=0, atol=1e-6 * 0.01 * dup ** 0.5 * 8 + 1e-7), \
                        (kind, aggr, t, "updated table far from the all-tables call")
            dist.barrier()
        # ---- the pipelined, column-split form (ShardedStep): heavy tables cut into column parts, parts
        # dealt into 2 groups per rank, push under the neighbouring group's compute, flags instead of
        # barriers.  Three steps in a row with fresh gradients: the next step's forward may overtake a
        # peer still reading the previous x (x is double-buffered by step parity). ----
        S = pkg.sharding
        for split_tables in ((), (1, 7)):
            parts = []
            for t in range(T):
                if t in split_tables:
                    parts += [S.Part(t, 0, D // 2, t % world), S.Part(t, D // 2, D // 2, (t + 1) % world)]
                else:
                    parts.append(S.Part(t, 0, D, t % world))
            for g in range(world):   # groups: by width where a rank holds two widths, else alternate
                mine_g = [p for p in parts if p.rank == g]
                widths = sorted({p.width for p in mine_g}, reverse=True)
                for i, p in enumerate(mine_g):
                    p.group = widths.index(p.width) if len(widths) > 1 else i % 2
            fx = pkg.exchange.FusedTableWiseExchange(T, None, B, D, rank, world, dev, parts=parts)
            K = fx.K
            mine = fx.mine
            Wp = [[Ws[p.table][:, p.col0:p.col0 + p.width].clone() for p in mine[k]] for k in range(K)]  # (a full-width slice is a view)
            colls = [pkg.EmbeddingCollection(Wp[k], [bags[p.table] for p in mine[k]], aggr) if mine[k] else None
                     for k in range(K)]
            sstep = pkg.exchange.ShardedStep(fx, colls, dev)
            truth = [w.clone() for w in Ws]
            tcoll = pkg.EmbeddingCollection(truth, bags, aggr)
            for it in range(3):
                # fresh indices and gradients every step; the tables move only in the last one (lr = 0 leaves
                # them bit-identical), so that every step's x can be held against the unsharded tables bit for bit
                lr = 0.01 if it == 2 else 0.0
                idx_it = [(i + it) % r for i, r in zip(idx, rows)]
                idxg = [[idx_it[p.table] for p in mine[k]] for k in range(K)]
                dxs = torch.rand((B, T * D), device=dev, generator=torch.Generator(device=dev).manual_seed(100 + it)) - 0.5
                fx.dx.copy_(dxs[lo:hi])
                want_x = torch.cat(tcoll.forward(idx_it), dim=1)
                x = sstep.run(idxg, lr)[:hi - lo]
                # no synchronisation with the peers here on purpose: compare, then go straight into the next step
                assert torch.equal(x, want_x[lo:hi]), ("pipelined", aggr, split_tables, it, "interaction input differs")
                for k in range(K):
                    for j, p in enumerate(mine[k]):
                        c0 = p.table * D + p.col0
                        assert torch.equal(fx.outgrads_by_group[k][j], dxs[:, c0:c0 + p.width]), \
                            ("pipelined", aggr, split_tables, it, p, "routed output-grad differs")
                tcoll.backward_sgd(idx_it, [dxs[:, t * D:(t + 1) * D].contiguous() for t in range(T)], lr)
                torch.cuda.synchronize()
                for k in range(K):
                    for j, p in enumerate(mine[k]):
                        want = truth[p.table][:, p.col0:p.col0 + p.width]
                        cnt = torch.bincount(idx_it[p.table].reshape(-1), minlength=rows[p.table])
                        assert torch.equal(Wp[k][j][cnt <= 1], want[cnt <= 1]), ("pipelined", it, p, "single-lookup rows differ")
                        dup = B * bags[p.table] / rows[p.table] + 1
                        assert torch.allclose(Wp[k][j], want, rtol=0, atol=1e-6 * 0.01 * dup ** 0.5 * 8 + 1e-7), \
                            ("pipelined", aggr, split_tables, it, p, "updated part far from the unsharded table")
            # the same step replayed from a CUDA graph (ShardedStepGraph) on the SAME state: eager step, tables put
            # back, graph step — interaction input, routed output-grads and updated parts must be bit-identical
            for it in (3, 4):
                idx_it = [(i + it) % r for i, r in zip(idx, rows)]
                idxg = [[idx_it[p.table] for p in mine[k]] for k in range(K)]
                dxs = torch.rand((B, T * D), device=dev, generator=torch.Generator(device=dev).manual_seed(100 + it)) - 0.5
                fx.dx.copy_(dxs[lo:hi])
                saved = [[w.clone() for w in Wp[k]] for k in range(K)]
                x_e = sstep.run(idxg, 0.01)[:hi - lo].clone()
                torch.cuda.synchronize()
                dist.barrier()
                after_e = [[w.clone() for w in Wp[k]] for k in range(K)]
                og_e = [[g.clone() for g in fx.outgrads_by_group[k]] for k in range(K)]
                for k in range(K):
                    for j in range(len(Wp[k])):
                        Wp[k][j].copy_(saved[k][j])
                torch.cuda.synchronize()
                dist.barrier()
                gr = sstep.capture(idxg, 0.01, parity=fx.parity ^ 1)
                x_g = gr.replay()[:hi - lo]
                torch.cuda.synchronize()
                assert torch.equal(x_g, x_e), ("graph replay", aggr, split_tables, it, "interaction input differs from the eager step")
                for k in range(K):
                    for j, p in enumerate(mine[k]):
                        assert torch.equal(fx.outgrads_by_group[k][j], og_e[k][j]), ("graph replay", it, p, "routed output-grad")
                        assert torch.equal(Wp[k][j], after_e[k][j]), ("graph replay", it, p, "updated part differs from the eager step")
                del gr
                dist.barrier()
        # ---- forward-only loop across ranks (no backward between the steps): x of step k must survive
        # until this rank has compared it although the peers are already storing step k + 1 ----
        fx = pkg.exchange.FusedTableWiseExchange(T, owner, B, D, rank, world, dev)
        myW = [Ws[t].clone() for t in owned]
        coll = pkg.EmbeddingCollection(myW, [bags[t] for t in owned], aggr) if owned else None
        shifted = [[(i + s) % r for i, r in zip(idx, rows)] for s in range(4)]
        wants = [torch.cat(pkg.EmbeddingCollection(Ws, bags, aggr).forward(sh), dim=1)[lo:hi].clone() for sh in shifted]
        for s4 in range(4):
            x = fx.forward(coll, [shifted[s4][t] for t in owned])[:hi - lo]
            assert torch.equal(x, wants[s4]), ("forward-only", aggr, s4)
        dist.barrier()
    # ---- row-wise sharded table: remote gathers in the forward, all-gather + local backward ----
    for (R, L, aggr) in ((100003, 1, 21), (5001, 5, 21), (70001, 3, 22)):
        g = torch.Generator(device=dev).manual_seed(11 + L)
        W = torch.rand((R, D), device=dev, generator=g) - 0.5
        idx = torch.randint(0, R, (B, L), device=dev, generator=g)
        og = torch.rand((B, D), device=dev, generator=g) - 0.5
        refW = W.clone()
        one = pkg.EmbeddingCollection([refW], [L], aggr)
        ref_out = one.forward([idx])[0]
        one.backward_sgd([idx], [og], 0.01)
        lo, hi = pkg.sharding.batch_block(B, world, rank)
        rw = pkg.exchange.RowWiseTable(R, D, L, B, rank, world, dev, aggr)
        rw.load_rows(W)
        out = rw.forward(idx[lo:hi].contiguous())
        torch.cuda.synchronize()
        assert torch.equal(out, ref_out[lo:hi]), ("row-wise forward differs", R, L, aggr)
        rw.backward_sgd(idx[lo:hi].contiguous(), og[lo:hi].contiguous(), 0.01)
        torch.cuda.synchronize()
        r0 = rank * rw.rps
        mine, want = rw.shard[:rw.local_rows], refW[r0:r0 + rw.local_rows]
        cnt = torch.bincount(idx.reshape(-1), minlength=R)[r0:r0 + rw.local_rows]
        assert torch.equal(mine[cnt <= 1], want[cnt <= 1]), ("row-wise update: single-lookup rows differ", R, L)
        assert torch.allclose(mine, want, rtol=0, atol=1e-6), ("row-wise update differs", R, L)
        out2 = rw.forward(idx[lo:hi].contiguous())  # next step reads the updated shards of every rank
        torch.cuda.synchronize()
        assert torch.allclose(out2, one.forward([idx])[0][lo:hi], rtol=0, atol=1e-5)
        dist.barrier()
    # ---- several row-wise tables in one step (RowWiseCollection).  routed: route kernel + flags + the owners'
    # padded backward, two steps in a row (the second forward reads every rank's updated shards);
    # legacy: round 1's two all-gathers + whole-batch backward ----
    for aggr in (21, 22):
        rows3, bags3 = [100003, 5001, 70001], [1, 5, 3]
        g = torch.Generator(device=dev).manual_seed(23)
        Ws = [torch.rand((r, D), device=dev, generator=g) - 0.5 for r in rows3]
        lo, hi = pkg.sharding.batch_block(B, world, rank)
        for legacy in (False, True):
            refW = [w.clone() for w in Ws]
            one = pkg.EmbeddingCollection(refW, bags3, aggr)
            rc = pkg.exchange.RowWiseCollection(rows3, D, bags3, B, rank, world, dev, aggr, legacy=legacy)
            rc.load_rows(Ws)
            for it in range(2):
                gi = torch.Generator(device=dev).manual_seed(31 + it)
                idx = [torch.randint(0, r, (B, l), device=dev, generator=gi) for r, l in zip(rows3, bags3)]
                dx = torch.rand((B, len(rows3) * D), device=dev, generator=gi) - 0.5
                ref_out = one.forward(idx)
                mine_idx = [i[lo:hi].contiguous() for i in idx]
                outs = rc.forward(mine_idx)
                torch.cuda.synchronize()
                for t in range(len(rows3)):
                    if it == 0:
                        assert torch.equal(outs[t], ref_out[t][lo:hi]), ("row-wise collection forward differs", t, legacy)
                    else:   # the tables of step 2 carry step 1's update: equal up to its summation order
                        assert torch.allclose(outs[t], ref_out[t][lo:hi], rtol=0, atol=1e-5), ("row-wise forward, step 2", t)
                one.backward_sgd(idx, [dx[:, t * D:(t + 1) * D].contiguous() for t in range(len(rows3))], 0.01)
                if legacy:
                    rc.backward_sgd(torch.cat(mine_idx, dim=1).contiguous(), dx[lo:hi].contiguous(), 0.01)
                else:
                    rc.backward_sgd(mine_idx, dx[lo:hi].contiguous(), 0.01)
                torch.cuda.synchronize()
                dist.barrier()
                for t in range(len(rows3)):
                    r0 = rank * rc.rps[t]
                    mine, want = rc.shards[t][:rc.local_rows[t]], refW[t][r0:r0 + rc.local_rows[t]]
                    if not legacy:   # routed slots / rows: bit-exact
                        full = idx[t]
                        own = (full >= r0) & (full < r0 + rc.rps[t])
                        assert torch.equal(rc.slot_views[t][:B], torch.where(own, full - r0, torch.full_like(full, -1))), \
                            ("routed slots differ", t, it)
                        assert bool((rc.slot_views[t][B:] == -1).all()), ("slots past the batch must stay empty", t)
                        bags_here = own.any(dim=1)
                        assert torch.equal(rc.grow_views[t][:B][bags_here], dx[bags_here, t * D:(t + 1) * D]), \
                            ("routed output-grad rows differ", t, it)
                    if it == 0:
                        cnt = torch.bincount(idx[t].reshape(-1), minlength=rows3[t])[r0:r0 + rc.local_rows[t]]
                        assert torch.equal(mine[cnt <= 1], want[cnt <= 1]), ("row-wise collection: single-lookup rows differ", t, legacy)
                    assert torch.allclose(mine, want, rtol=0, atol=2e-6), ("row-wise collection update differs", t, legacy, it)
                # (the checks above read the routed buffers from outside the protocol: nobody may route the
                # next step's lookups into them before every rank is done looking)
                torch.cuda.synchronize()
                dist.barrier()
            dist.barrier()
    if rank == 0:
        print(f"multigpu_check ok: world={world}, fused and nccl exchanges bit-exact vs single GPU")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
