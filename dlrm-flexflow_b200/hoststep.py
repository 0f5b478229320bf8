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
step() returns when all three streams have drained (the caller owns the host buffers again).
submit()/wait() is the same step without the blocking tail: the device-side buffers are double
buffered, so step k+1's index / gradient upload queues right behind step k's on the copy-in
stream while step k still computes and downloads — a training loop that prefetches its next batch
keeps the H2D engine (the bottleneck: 1.03 GB up per cfg3 step) busy back to back.
torch is plumbing here: pinned buffers, streams, events; every kernel is libffemb's.
"""
import torch


class HostStep:
    def __init__(self, coll, batch, exchange=None, depth=2):
        """coll: EmbeddingCollection of this GPU's tables (or None on a rank that owns none);
        exchange: FusedTableWiseExchange for N > 1 (then the host-side result is the rank's
        interaction input x and the host-side gradient is its dx; those live in one symmetric
        buffer each, so steps are not overlapped there: depth 1)."""
        self.coll, self.fx, self.B = coll, exchange, int(batch)
        dev = coll.device if coll is not None else exchange.device
        self.dev = dev
        self.depth = 1 if exchange is not None else max(1, int(depth))
        self.s_main = torch.cuda.current_stream(dev)
        self.s_in, self.s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        ev = lambda: [torch.cuda.Event() for _ in range(self.depth)]
        # per buffer set: indices landed / grads landed / forward done / D2H done / backward done
        self.ev_idx, self.ev_g, self.ev_fwd, self.ev_out, self.ev_done = ev(), ev(), ev(), ev(), ev()
        self.d_idx, self.d_g, self.d_out = [], [], []
        f32 = dict(dtype=torch.float32, device=dev)
        for _ in range(self.depth):
            self.d_idx.append([torch.empty((self.B, l), dtype=torch.int64, device=dev) for l in coll.bags]
                              if coll is not None else [])
            if exchange is None:
                self.d_g.append([torch.empty((self.B, coll.dim), **f32) for _ in coll.bags])
                self.d_out.append([torch.empty((self.B, coll.dim), **f32) for _ in coll.bags])
        self.ticket = 0

    def submit(self, h_idx, h_grad, h_out, lr):
        """Enqueue one step (see step()); returns a ticket for wait().  Does not block: the host
        buffers belong to the step until wait(ticket) returns."""
        coll, fx, main = self.coll, self.fx, self.s_main
        k = self.ticket
        s = k % self.depth
        d_idx = self.d_idx[s]
        with torch.cuda.stream(self.s_in):
            self.s_in.wait_event(self.ev_done[s])  # the step that last used this buffer set is done
            for d, h in zip(d_idx, h_idx):
                d.copy_(h, non_blocking=True)
            self.ev_idx[s].record(self.s_in)
            if fx is None:
                for d, h in zip(self.d_g[s], h_grad):
                    d.copy_(h, non_blocking=True)
            else:
                fx.dx.copy_(h_grad, non_blocking=True)
            self.ev_g[s].record(self.s_in)
        main.wait_event(self.ev_idx[s])
        main.wait_event(self.ev_out[s])            # the previous download out of d_out[s] has drained
        if fx is None:
            coll.forward(d_idx, self.d_out[s])
        else:
            fx.forward(coll, d_idx)
        self.ev_fwd[s].record(main)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self.ev_fwd[s])
            if fx is None:
                for h, d in zip(h_out, self.d_out[s]):
                    h.copy_(d, non_blocking=True)
            else:
                h_out.copy_(fx.x, non_blocking=True)
            self.ev_out[s].record(self.s_out)
        if coll is not None:
            coll.backward_prepare(d_idx)       # sort: indices only, runs while the grads upload
        main.wait_event(self.ev_g[s])
        if fx is not None:
            fx.backward()
            grads = fx.outgrads
        else:
            grads = self.d_g[s]
        if coll is not None:
            coll.backward_sgd(d_idx, grads, lr, prepared=True)
        self.ev_done[s].record(main)
        self.ticket = k + 1
        return k

    def wait(self, ticket=None):
        """Block until step `ticket` (default: the last submitted) has finished: its pooled outputs
        are in h_out, its host inputs may be reused, its row updates are in the tables."""
        k = self.ticket - 1 if ticket is None else int(ticket)
        if k < 0:
            return
        # (if the buffer set of step k has been reused since, its events now stand for the later
        # step, which was ordered after k: waiting for it is still correct)
        s = k % self.depth
        self.ev_out[s].synchronize()
        self.ev_done[s].synchronize()

    def join(self, stream=None):
        """Make `stream` (default: the compute stream) wait for every outstanding copy — for
        CUDA-event timing of a run of submit() calls."""
        st = self.s_main if stream is None else stream
        for s in range(self.depth):
            st.wait_event(self.ev_out[s])
            st.wait_event(self.ev_done[s])

    def step(self, h_idx, h_grad, h_out, lr):
        """h_idx: pinned int64 [B][L] per table; h_grad: pinned output-grads ([B][D] per table, or
        this rank's dx with an exchange); h_out: pinned destination(s) of the pooled outputs (or of
        x with an exchange).  Blocks until the host buffers may be reused."""
        self.wait(self.submit(h_idx, h_grad, h_out, lr))
