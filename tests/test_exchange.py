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
