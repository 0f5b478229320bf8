"""world_size-2/4 gloo tests (CPU) of the table-wise exchange: the host-side logic that routes
pooled rows into the Concat layout and output-grads back to the table owners.  Expected values
come from the oracle's restatement of Concat (src/ops/concat.cu:211-251) and the shard maps."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, T, B, D, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pkg = importlib.import_module("dlrm-flexflow_b200")
    own = [t % world for t in range(T)]
    ex = pkg.exchange.TableWiseExchange(T, own, B, D, rank, world, torch.device("cpu"))
    # pooled output of table t, sample b, column d = t*1e6 + b*1e2 + d  (exactly representable)
    def pooled(t):
        b = torch.arange(B, dtype=torch.float32).view(B, 1)
        d = torch.arange(D, dtype=torch.float32).view(1, D)
        return t * 1.0e4 + b * 16.0 + d
    outs = [pooled(t) for t in ex.owned]
    x = ex.forward(outs).clone()
    ex.dx.copy_(x * 2.0)
    grads = [torch.zeros(B, D) for _ in ex.owned]
    ex.backward(grads)
    ok_back = all(torch.equal(g, 2.0 * pooled(t)) for g, t in zip(grads, ex.owned))
    q.put((rank, x.numpy(), ok_back))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,T,B,D", [(2, 5, 7, 4), (4, 26, 33, 8), (2, 1, 8, 16)])
def test_tablewise_exchange_gloo(world, T, B, D, oracle):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, T, B, D, q)) for r in range(world)]
    [p.start() for p in procs]
    res = [q.get(timeout=120) for _ in range(world)]
    [p.join(timeout=60) for p in procs]
    b = np.arange(B, dtype=np.float32).reshape(B, 1)
    d = np.arange(D, dtype=np.float32).reshape(1, D)
    full = oracle.concat_forward([t * 1.0e4 + b * 16.0 + d for t in range(T)])  # [B][T*D]
    for rank, x, ok_back in res:
        lo, hi = oracle.batch_block(B, world, rank)
        assert np.array_equal(x, full[lo:hi]), rank
        assert ok_back, rank


# ---- row-wise routed backward over process boundaries (the slot-position contract of
# ffemb_rowwise_route, include/ffemb.h): rank r routes ITS samples, owner g ends up with slot arrays and
# gradient rows at the GLOBAL bag positions r*ext + sample, and the owners' backward over those arrays
# equals the unsharded reference step.  gloo stands in for the peer stores.
def _rowwise_worker(rank, world, port, R, D, L, B, aggr, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pkg = importlib.import_module("dlrm-flexflow_b200")
    from oracle import oracle as O
    S = pkg.sharding
    rng = np.random.default_rng(77)  # every rank draws the same global problem and keeps its part
    idx = rng.integers(0, R, (B, L), dtype=np.int64)
    idx[: B // 3] = idx[0]
    W = rng.standard_normal((R, D)).astype(np.float32)
    og = rng.standard_normal((B, D)).astype(np.float32)
    lo, hi = S.batch_block(B, world, rank)
    ext = (B + world - 1) // world
    rps = S.rows_per_shard(R, world)
    # route: for every owner, the slots of MY samples and my gradient rows, padded to ext samples
    send_slots = [torch.full((ext, L), -1, dtype=torch.int64) for _ in range(world)]
    send_rows = [torch.zeros((ext, D), dtype=torch.float32) for _ in range(world)]
    for g in range(world):
        mine = S.route_slots(idx[lo:hi], R, world, g)
        send_slots[g][: hi - lo] = torch.from_numpy(mine)
        touched = (mine >= 0).any(axis=1)  # a gradient row goes only to owners of at least one lookup
        rows = np.where(touched[:, None], og[lo:hi], 0.0).astype(np.float32)
        send_rows[g][: hi - lo] = torch.from_numpy(rows)
    # (gloo has no all-to-all: gather every rank's whole send list, keep the entries addressed to me)
    all_slots = [torch.empty((world, ext, L), dtype=torch.int64) for _ in range(world)]
    all_rows = [torch.empty((world, ext, D), dtype=torch.float32) for _ in range(world)]
    dist.all_gather(all_slots, torch.stack(send_slots))
    dist.all_gather(all_rows, torch.stack(send_rows))
    recv_slots = [all_slots[src][rank] for src in range(world)]
    recv_rows = [all_rows[src][rank] for src in range(world)]
    slots = torch.cat(recv_slots).numpy()[:B]  # position = global bag id: nothing compacted
    grow = torch.cat(recv_rows).numpy()[:B]
    n = S.shard_rows(R, world, rank)
    ok_slots = np.array_equal(slots, S.route_slots(idx, R, world, rank))
    shard = W[rank * rps:rank * rps + n]
    if n > 0:
        filled = np.where(slots >= 0, slots, n)  # an empty slot touches a throw-away row
        grad = O.backward_dense(filled, grow, n + 1, aggr=aggr)[:n]
        shard = O.sgd_update(shard, grad, 0.05)
    q.put((rank, ok_slots, shard))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,R,D,L,B,aggr", [(2, 101, 8, 3, 37, 21), (4, 64, 4, 5, 50, 22), (2, 7, 16, 1, 16, 21)])
def test_rowwise_route_contract_gloo(world, R, D, L, B, aggr, ffemb, oracle):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() % 2000) + world
    procs = [ctx.Process(target=_rowwise_worker, args=(r, world, port, R, D, L, B, aggr, q)) for r in range(world)]
    [p.start() for p in procs]
    res = sorted((q.get(timeout=180) for _ in range(world)), key=lambda t: t[0])
    [p.join(timeout=60) for p in procs]
    rng = np.random.default_rng(77)
    idx = rng.integers(0, R, (B, L), dtype=np.int64)
    idx[: B // 3] = idx[0]
    W = rng.standard_normal((R, D)).astype(np.float32)
    og = rng.standard_normal((B, D)).astype(np.float32)
    want = oracle.reference_step(idx, og, W, aggr=aggr, lr=0.05)
    assert all(ok for _, ok, _ in res)
    got = np.concatenate([s for _, _, s in res])
    assert got.shape == want.shape
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
