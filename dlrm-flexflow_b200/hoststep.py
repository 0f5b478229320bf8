"""One embedding step driven from HOST buffers — the reference's contract for this operator: its
pooled-output and output-grad regions live in zero-copy host memory (MAP_TO_ZC_MEMORY,
/root/reference/src/ops/embedding.cu:295-298,377-380) and its index batches are copied in from
pinned host memory every step (/root/reference/examples/cpp/DLRM/dlrm.cu:19-58).

HostStep keeps that contract but takes the copies off the kernels' critical path:
  copy-in stream : indices H2D, then output-grads H2D (the big one)
  compute stream : forward as soon as the indices landed; backward as soon as the grads landed;
                   the index sort of the backward runs right after the forward (indices only)
  copy-out stream: pooled outputs D2H as soon as the forward finished — in the opposite PCIe
                   direction of the output-grad upload, so the two overlap (full duplex)
The step returns when all three streams have drained (the caller owns the host buffers again).
torch is plumbing here: pinned buffers, streams, events; every kernel is libffemb's.
"""
import torch


class HostStep:
    def __init__(self, coll, batch, exchange=None):
        """coll: EmbeddingCollection of this GPU's tables (or None on a rank that owns none);
        exchange: FusedTableWiseExchange for N > 1 (then the host-side result is the rank's
        interaction input x and the host-side gradient is its dx)."""
        self.coll, self.fx, self.B = coll, exchange, int(batch)
        dev = coll.device if coll is not None else exchange.device
        self.dev = dev
        self.s_main = torch.cuda.current_stream(dev)
        self.s_in, self.s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        self.ev_idx, self.ev_g, self.ev_fwd, self.ev_free = (torch.cuda.Event() for _ in range(4))
        if coll is not None:
            self.d_idx = [torch.empty((self.B, l), dtype=torch.int64, device=dev) for l in coll.bags]
        else:
            self.d_idx = []
        if exchange is None:
            f32 = dict(dtype=torch.float32, device=dev)
            self.d_g = [torch.empty((self.B, coll.dim), **f32) for _ in coll.bags]
            self.d_out = [torch.empty((self.B, coll.dim), **f32) for _ in coll.bags]

    def step(self, h_idx, h_grad, h_out, lr):
        """h_idx: pinned int64 [B][L] per table; h_grad: pinned output-grads ([B][D] per table, or
        this rank's dx with an exchange); h_out: pinned destination(s) of the pooled outputs (or of
        x with an exchange).  Blocks until the host buffers may be reused."""
        coll, fx, main = self.coll, self.fx, self.s_main
        self.ev_free.record(main)              # device buffers of the previous step are free
        with torch.cuda.stream(self.s_in):
            self.s_in.wait_event(self.ev_free)
            for d, h in zip(self.d_idx, h_idx):
                d.copy_(h, non_blocking=True)
            self.ev_idx.record(self.s_in)
            if fx is None:
                for d, h in zip(self.d_g, h_grad):
                    d.copy_(h, non_blocking=True)
            else:
                fx.dx.copy_(h_grad, non_blocking=True)
            self.ev_g.record(self.s_in)
        main.wait_event(self.ev_idx)
        if fx is None:
            coll.forward(self.d_idx, self.d_out)
        else:
            fx.forward(coll, self.d_idx)
        self.ev_fwd.record(main)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self.ev_fwd)
            if fx is None:
                for h, d in zip(h_out, self.d_out):
                    h.copy_(d, non_blocking=True)
            else:
                h_out.copy_(fx.x, non_blocking=True)
        if coll is not None:
            coll.backward_prepare(self.d_idx)  # sort: indices only, runs while the grads upload
        main.wait_event(self.ev_g)
        if fx is not None:
            fx.backward()
            grads = fx.outgrads
        else:
            grads = self.d_g
        if coll is not None:
            coll.backward_sgd(self.d_idx, grads, lr, prepared=True)
        self.s_out.synchronize()
        main.synchronize()
