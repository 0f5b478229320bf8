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
    GRAPH_MAX_LOOKUPS = 2_000_000

    def __init__(self, coll, batch, exchange=None, depth=2, graph="auto"):
        """coll: EmbeddingCollection of this GPU's tables (or None on a rank that owns none); with an
        exchange: the list of per-group collections of this rank's parts (exchange.ShardedStep order,
        None for an empty group).
        exchange: FusedTableWiseExchange for N > 1 (then the host-side result is the rank's
        interaction input x and the host-side gradient is its dx; dx lives in one buffer, so
        steps are not overlapped there: depth 1).
        graph: replay the kernels of a step from two CUDA graphs per buffer set (forward + index sort;
        segmented sum + SGD) instead of launching them one by one — "auto" = single-GPU steps of at
        most GRAPH_MAX_LOOKUPS lookups, which are bound by the host's launch rate, not by the GPU or
        PCIe.  The copies stay eager (three per step), so steps still overlap across buffer sets."""
        self.fx, self.B = exchange, int(batch)
        self.colls = None
        if exchange is not None:
            from .exchange import ShardedStep
            self.colls = list(coll) if isinstance(coll, (list, tuple)) else [coll]
            live = [c for c in self.colls if c is not None]
            coll = live[0] if live else None
            self.sharded = ShardedStep(exchange, self.colls, exchange.device)
            self.group_bags = [list(c.bags) if c is not None else [] for c in self.colls]
        self.coll = coll
        small = coll is not None and exchange is None and int(batch) * sum(coll.bags) <= self.GRAPH_MAX_LOOKUPS
        self.use_graph = exchange is None and coll is not None and (graph is True or graph == "on" or
                                                                    (graph == "auto" and small))
        self._graphs, self._last_lr, self.graph_replays = {}, {}, 0
        dev = coll.device if coll is not None else exchange.device
        self.dev = dev
        self.depth = 1 if exchange is not None else max(1, int(depth))
        self.s_main = torch.cuda.current_stream(dev)
        self.s_in, self.s_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        ev = lambda: [torch.cuda.Event() for _ in range(self.depth)]
        # per buffer set: indices landed / grads landed / forward done / D2H done / backward done
        self.ev_idx, self.ev_g, self.ev_fwd, self.ev_out, self.ev_done = ev(), ev(), ev(), ev(), ev()
        # device side: one slab per kind and buffer set, the per-table tensors are views of it — when the
        # caller's host tensors are views of one slab too (slab_like), a step moves each kind with ONE
        # copy instead of one per table (configs 1/2 are copy-launch-bound: 78 copies -> 3)
        self.d_idx, self.d_g, self.d_out = [], [], []
        self.slab_idx, self.slab_g, self.slab_out = [], [], []
        f32 = dict(dtype=torch.float32, device=dev)
        for _ in range(self.depth):
            bags = [l for g in self.group_bags for l in g] if exchange is not None else (coll.bags if coll is not None else [])
            if bags:
                slab, views = self._slab([(self.B, l) for l in bags], torch.int64, dev)
            else:
                slab, views = None, []
            self.slab_idx.append(slab)
            self.d_idx.append(views)
            if exchange is None:
                for dst_s, dst_v in ((self.slab_g, self.d_g), (self.slab_out, self.d_out)):
                    slab, views = self._slab([(self.B, coll.dim)] * len(coll.bags), torch.float32, dev)
                    dst_s.append(slab)
                    dst_v.append(views)
        self.ticket = 0

    @staticmethod
    def _slab(shapes, dtype, device, pin=False):
        """One flat buffer + per-table views (each view 16-byte aligned: sizes are padded to 16 bytes)."""
        esz = torch.empty((), dtype=dtype).element_size()
        pad = max(1, 16 // esz)
        offs, n = [], 0
        for shp in shapes:
            offs.append(n)
            k = 1
            for d in shp:
                k *= int(d)
            n += (k + pad - 1) // pad * pad
        slab = torch.empty(max(n, 1), dtype=dtype, device=device)
        if pin:
            slab = slab.pin_memory()
        views = []
        for shp, o in zip(shapes, offs):
            k = 1
            for d in shp:
                k *= int(d)
            views.append(slab[o:o + k].view(*shp))
        return slab, views

    def host_buffers(self):
        """Pinned host tensors laid out like the device slabs: (h_idx, h_grad, h_out) lists of per-table
        views; a step on them moves each kind with a single copy."""
        coll = self.coll
        _, hi = self._slab([(self.B, l) for l in coll.bags], torch.int64, "cpu", pin=True)
        _, hg = self._slab([(self.B, coll.dim)] * len(coll.bags), torch.float32, "cpu", pin=True)
        _, ho = self._slab([(self.B, coll.dim)] * len(coll.bags), torch.float32, "cpu", pin=True)
        return hi, hg, ho

    @staticmethod
    def _span(views, slab):
        """(first pointer, elements) when `views` tile one contiguous region laid out like `slab`'s
        views, else None."""
        if slab is None or not views or not isinstance(views, (list, tuple)):
            return None
        esz = views[0].element_size()
        pad = max(1, 16 // esz)
        p = views[0].data_ptr()
        for v in views:
            if not v.is_contiguous() or v.data_ptr() != p:
                return None
            p += (v.numel() + pad - 1) // pad * pad * esz
        return views[0].data_ptr(), (p - views[0].data_ptr()) // esz

    @staticmethod
    def _copy(dst_views, dst_slab, src_views, src_slab_like, to_host=False):
        """Per-table copies, or one copy when the host side is laid out like the device slab."""
        host_views, dev_slab = (dst_views, src_slab_like) if to_host else (src_views, dst_slab)
        span = HostStep._span(host_views, dev_slab)
        if span is not None and span[1] <= dev_slab.numel():
            n = span[1]
            try:
                flat = torch.as_strided(host_views[0], (n,), (1,))   # the contiguous host region the views tile
            except RuntimeError:
                flat = None                                           # not one storage: per-table copies
            if flat is not None:
                if to_host:
                    flat.copy_(dev_slab[:n], non_blocking=True)
                else:
                    dev_slab[:n].copy_(flat, non_blocking=True)
                return 1
        for d, h in zip(dst_views, src_views):
            d.copy_(h, non_blocking=True)
        return len(dst_views)

    def submit(self, h_idx, h_grad, h_out, lr):
        """Enqueue one step (see step()); returns a ticket for wait().  Does not block: the host
        buffers belong to the step until wait(ticket) returns."""
        coll, fx, main = self.coll, self.fx, self.s_main
        k = self.ticket
        s = k % self.depth
        d_idx = self.d_idx[s]
        with torch.cuda.stream(self.s_in):
            self.s_in.wait_event(self.ev_done[s])  # the step that last used this buffer set is done
            self._copy(d_idx, self.slab_idx[s], h_idx, None)
            self.ev_idx[s].record(self.s_in)
            if fx is None:
                self._copy(self.d_g[s], self.slab_g[s], h_grad, None)
            else:
                fx.dx.copy_(h_grad, non_blocking=True)
            self.ev_g[s].record(self.s_in)
        gr = self._step_graphs(s, float(lr)) if self.use_graph else None
        main.wait_event(self.ev_idx[s])
        main.wait_event(self.ev_out[s])            # the previous download out of d_out[s] has drained
        if fx is not None:
            # sharded step: forward + exchange, push (once this step's dx has landed), backward, pipelined
            # over the exchange's groups; the download of x starts the moment x is complete on this rank
            groups, o = [], 0
            for g in self.group_bags:
                groups.append(d_idx[o:o + len(g)])
                o += len(g)
            x = self.sharded.run(groups, lr, dx_ready=self.ev_g[s], stream=main)
            with torch.cuda.stream(self.s_out):
                self.s_out.wait_event(self.sharded.ev_x)
                h_out.copy_(x, non_blocking=True)
                self.ev_out[s].record(self.s_out)
            self.ev_done[s].record(main)
            self.ticket = k + 1
            return k
        if gr is not None:
            with torch.cuda.stream(main):
                gr[0].replay()                     # forward + keygen / radix sort
        else:
            coll.forward(d_idx, self.d_out[s])
        self.ev_fwd[s].record(main)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self.ev_fwd[s])
            self._copy(h_out, None, self.d_out[s], self.slab_out[s], to_host=True)
            self.ev_out[s].record(self.s_out)
        if gr is None:
            coll.backward_prepare(d_idx)       # sort: indices only, runs while the grads upload
        main.wait_event(self.ev_g[s])
        if gr is not None:
            with torch.cuda.stream(main):
                gr[1].replay()                     # segmented sum + SGD row updates
            self.graph_replays += 2
            self.ev_done[s].record(main)
            self.ticket = k + 1
            return k
        coll.backward_sgd(d_idx, self.d_g[s], lr, prepared=True)
        self.ev_done[s].record(main)
        self.ticket = k + 1
        return k

    def _step_graphs(self, s, lr):
        """The two graphs of buffer set s for this learning rate, captured the second time the set is
        used with it (the first use runs eagerly: it is the warm-up — workspace, side stream, kernel
        attributes — and the device buffers hold real indices from then on).  A step with another
        learning rate runs eagerly; the graphs are re-captured when the new rate repeats."""
        have = self._graphs.get(s)
        if have is not None and have[0] == lr:
            return have[1]
        seen, self._last_lr[s] = self._last_lr.get(s), lr
        if seen != lr:
            return None
        coll, d_idx = self.coll, self.d_idx[s]
        ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.graph(ga, capture_error_mode="thread_local"):
            coll.forward(d_idx, self.d_out[s])
            coll.backward_prepare(d_idx)
        with torch.cuda.graph(gb, capture_error_mode="thread_local"):
            coll.backward_sgd(d_idx, self.d_g[s], lr, prepared=True)
        self._graphs[s] = (lr, (ga, gb))
        return self._graphs[s][1]

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
