"""Exchange of pooled embeddings between table owners and the data-parallel part of the model.

The reference has no collective here: each table's pooled output lives in zero-copy host memory
and Legion copies batch slices to the consumer GPUs when the data-parallel Concat partitions it
(/root/reference/src/ops/embedding.cu:295-298, src/ops/concat.cu:88-97).  That
implicit all-to-all is made explicit here:

  forward : owner g holds pooled [B][D] for each owned table; sample block h of every owned
            table goes to rank h and lands in column block t of rank h's interaction input
            x_h [B_h][T*D]  (the Concat layout, src/ops/concat.cu:211-251).
  backward: the transpose — column block t of dx_h goes back to owner(t) rows [lo_h, hi_h).

TableWiseExchange is the collective form (torch.distributed all_to_all_single: NCCL on GPUs,
gloo in the CPU tests).  The batch split is Legion's blocked split (sharding.batch_block).
"""
import torch
import torch.distributed as dist

from .sharding import batch_block


class TableWiseExchange:
    def __init__(self, num_tables, owner, batch, dim, rank, world, device, group=None):
        self.T, self.owner, self.B, self.D = int(num_tables), list(owner), int(batch), int(dim)
        self.rank, self.world, self.device, self.group = rank, world, device, group
        self.owned_by = [[t for t in range(self.T) if self.owner[t] == g] for g in range(world)]
        self.owned = self.owned_by[rank]
        self.blocks = [batch_block(self.B, world, h) for h in range(world)]
        lo, hi = self.blocks[rank]
        self.local_batch = hi - lo
        n = len(self.owned)
        # forward: send to h the block of h of each owned table; receive from g its tables' rows of my block
        self.fwd_send = [n * (b[1] - b[0]) * self.D for b in self.blocks]
        self.fwd_recv = [len(self.owned_by[g]) * self.local_batch * self.D for g in range(world)]
        f32 = dict(dtype=torch.float32, device=device)
        self.sendbuf = torch.empty(max(1, sum(self.fwd_send)), **f32)
        self.recvbuf = torch.empty(max(1, sum(self.fwd_recv)), **f32)
        self.x = torch.zeros((self.local_batch, self.T * self.D), **f32)    # interaction input (embedding part)
        self.dx = torch.zeros((self.local_batch, self.T * self.D), **f32)   # its gradient

    def _offsets(self, counts):
        off, o = [], 0
        for c in counts:
            off.append(o)
            o += c
        return off

    def forward(self, outs):
        """outs: pooled [B][D] per owned table (owner side) -> self.x on every rank."""
        n, D = len(self.owned), self.D
        off = self._offsets(self.fwd_send)
        for h, (lo, hi) in enumerate(self.blocks):
            if n and hi > lo:
                dst = self.sendbuf[off[h]:off[h] + self.fwd_send[h]].view(n, hi - lo, D)
                for i in range(n):
                    dst[i].copy_(outs[i][lo:hi])
        dist.all_to_all_single(self.recvbuf[:sum(self.fwd_recv)], self.sendbuf[:sum(self.fwd_send)],
                               self.fwd_recv, self.fwd_send, group=self.group)
        roff = self._offsets(self.fwd_recv)
        xv = self.x.view(self.local_batch, self.T, D)
        for g in range(self.world):
            ts = self.owned_by[g]
            if ts and self.local_batch:
                src = self.recvbuf[roff[g]:roff[g] + self.fwd_recv[g]].view(len(ts), self.local_batch, D)
                for i, t in enumerate(ts):
                    xv[:, t, :].copy_(src[i])
        return self.x

    def backward(self, outgrads):
        """self.dx on every rank -> outgrads: [B][D] per owned table on its owner."""
        n, D = len(self.owned), self.D
        roff = self._offsets(self.fwd_recv)
        dxv = self.dx.view(self.local_batch, self.T, D)
        for g in range(self.world):
            ts = self.owned_by[g]
            if ts and self.local_batch:
                dst = self.recvbuf[roff[g]:roff[g] + self.fwd_recv[g]].view(len(ts), self.local_batch, D)
                for i, t in enumerate(ts):
                    dst[i].copy_(dxv[:, t, :])
        dist.all_to_all_single(self.sendbuf[:sum(self.fwd_send)], self.recvbuf[:sum(self.fwd_recv)],
                               self.fwd_send, self.fwd_recv, group=self.group)
        off = self._offsets(self.fwd_send)
        for h, (lo, hi) in enumerate(self.blocks):
            if n and hi > lo:
                src = self.sendbuf[off[h]:off[h] + self.fwd_send[h]].view(n, hi - lo, D)
                for i in range(n):
                    outgrads[i][lo:hi].copy_(src[i])
        return outgrads


class FusedTableWiseExchange:
    """The exchange fused into the kernels, over NVLink peer memory (no NCCL on the data path), cut into
    pipeline groups so that it overlaps the compute around it.

    Units are PARTS (sharding.Part): a whole table on its owner, or a column range of a heavy table.
    Every rank deals its parts into K groups (sharding.plan_shards); group k of the step is

      forward_group(k): ffemb_forward_exchange — the gather-pool kernel of the part owner stores every
              pooled row straight into the interaction input x_h [B_h][T*D] of the rank h that owns the
              sample (the Concat layout, /root/reference/src/ops/concat.cu:211-251), tile by tile while
              it computes; then ffemb_sync_signal(X_k).
      push_group(k)   : ffemb_sync_wait(X_k) — group k's pooled rows of EVERY rank have landed here — then
              ffemb_exchange_backward_parts pushes column block (t, cols) of this rank's dx into rows
              [lo_h, hi_h) of the owner's output-grad buffer for every part of group k (bulk copies over
              NVLink), then ffemb_sync_signal(G_k).  Runs on its own stream: under forward_group(k + 1)
              and under the backward of group k - 1.
      wait_grads(k)   : ffemb_sync_wait(G_k) — every rank's block of group k's output-grads has landed;
              the owner's fused backward + SGD of group k follows on the same stream.

    The flags are epoch counters in a peer-mapped pad (include/ffemb.h, ffemb_sync_*): no host
    synchronisation, no collective call, capturable in a CUDA graph.  x is double-buffered by step parity:
    a rank may run ahead into the next step's forward while a peer still reads the previous x (the
    write-after-read hazard of a single buffer); a consumer of x(step) only has to be ordered before the
    same rank's forward of step + 1.  Buffers are torch symmetric memory (peer-mapped allocations:
    plumbing only).  K = 1 with whole tables is the un-pipelined form of round 1.
    """

    X_SLOT, G_SLOT = 0, 8   # flag slots: X_k = k, G_k = 8 + k  (FFEMB_SYNC_SLOTS = 16)

    def __init__(self, num_tables, owner, batch, dim, rank, world, device, group=None, parts=None, groups=None):
        import torch.distributed._symmetric_memory as symm
        from .binding import ffemb_sync_t, lib
        from .sharding import Part, num_groups
        self.T, self.B, self.D = int(num_tables), int(batch), int(dim)
        self.rank, self.world, self.device = rank, world, device
        self.group = group if group is not None else dist.group.WORLD
        if parts is None:
            parts = [Part(t, 0, self.D, int(owner[t]), 0) for t in range(self.T)]
        self.parts = sorted(parts, key=lambda p: (p.table, p.col0))
        self.K = int(groups) if groups is not None else num_groups(self.parts)
        if not 1 <= self.K <= 8:
            raise ValueError("1..8 pipeline groups")
        self.owner = list(owner) if owner is not None else None
        self.mine = [[p for p in self.parts if p.rank == rank and p.group == k] for k in range(self.K)]
        self.owned_parts = [p for k in range(self.K) for p in self.mine[k]]
        self.owned = [p.table for p in self.owned_parts]
        self.ext = (self.B + world - 1) // world
        self.lo, self.hi = batch_block(self.B, world, rank)
        self.local_batch = self.hi - self.lo
        self.stride = self.T * self.D
        # output-grad buffers of a rank: its parts back to back ([B][w] each), group by group
        self.g_off = {}
        sizes = [0] * world
        for g in range(world):
            for k in range(self.K):
                for p in self.parts:
                    if p.rank == g and p.group == k:
                        self.g_off[(p.table, p.col0)] = sizes[g]
                        sizes[g] += self.B * p.width
        f32 = dict(dtype=torch.float32, device=device)
        self.xbuf = symm.empty((2, self.ext, self.stride), **f32)        # interaction input, by step parity
        self.gbuf = symm.empty((max(4, max(sizes)),), **f32)              # output-grads of the owned parts
        self.pad = symm.empty((lib().ffemb_sync_pad_bytes() // 8,), dtype=torch.int64, device=device)
        self.hx = symm.rendezvous(self.xbuf, self.group)
        self.hg = symm.rendezvous(self.gbuf, self.group)
        self.hp = symm.rendezvous(self.pad, self.group)
        self.xbuf.zero_()
        self.gbuf.zero_()
        self.pad.zero_()
        self.dx = torch.zeros((self.local_batch, self.stride), **f32)    # gradient of x, produced locally
        self.outgrads_by_group = [
            [self.gbuf[self.g_off[(p.table, p.col0)]:self.g_off[(p.table, p.col0)] + self.B * p.width].view(self.B, p.width)
             for p in self.mine[k]] for k in range(self.K)]
        self.outgrads = [g for k in range(self.K) for g in self.outgrads_by_group[k]]
        self.parity = 1          # begin_step() flips it: the first step writes x[0]
        self._ex = [None, None]
        self._push = [None] * self.K
        sy = ffemb_sync_t()
        for h in range(world):
            sy.peer_pad[h] = int(self.hp.buffer_ptrs[h])
        sy.num_peers, sy.rank = world, rank
        self._sync = sy
        torch.cuda.synchronize(device)
        dist.barrier(group=self.group)

    # ---- the interaction input of the current step ----------------------------------------------
    @property
    def x(self):
        return self.xbuf[self.parity]

    def begin_step(self):
        self.parity ^= 1

    def _exchange_desc(self):
        if self._ex[self.parity] is None:
            from .binding import ffemb_exchange_t
            ex = ffemb_exchange_t()
            off = self.parity * self.ext * self.stride * 4
            for h in range(self.world):
                ex.peer_x[h] = int(self.hx.buffer_ptrs[h]) + off
            ex.num_peers, ex.batch_per_peer, ex.row_stride = self.world, self.ext, self.stride
            self._ex[self.parity] = ex
        return self._ex[self.parity]

    def signal(self, slot, stream=None):
        import ctypes as C
        from .binding import check, lib
        from .embedding import _stream_ptr
        check(lib().ffemb_sync_signal(C.byref(self._sync), slot, _stream_ptr(stream)))

    def wait(self, slot, stream=None):
        import ctypes as C
        from .binding import check, lib
        from .embedding import _stream_ptr
        check(lib().ffemb_sync_wait(C.byref(self._sync), slot, _stream_ptr(stream)))

    # ---- group-wise pieces of a step ---------------------------------------------------------------
    def forward_group(self, k, coll, indices, stream=None):
        """coll: EmbeddingCollection of this rank's parts of group k (in self.mine[k] order; None when
        the rank has none); fills columns of x on ALL ranks, then signals X_k."""
        import ctypes as C
        from .binding import check, ffemb_table_t, lib
        from .embedding import _stream_ptr
        mine = self.mine[k]
        if mine:
            n, w = len(mine), mine[0].width
            arr = (ffemb_table_t * n)()
            for i, p in enumerate(mine):
                arr[i].idx, arr[i].out, arr[i].table = indices[i].data_ptr(), None, coll.weights[i].data_ptr()
                arr[i].R, arr[i].L, arr[i].reserved = int(coll.weights[i].shape[0]), coll.bags[i], p.table * self.D + p.col0
            check(lib().ffemb_forward_exchange(arr, n, w, self.B, coll.aggr, C.byref(self._exchange_desc()),
                                               _stream_ptr(stream)))
        self.signal(self.X_SLOT + k, stream)

    def wait_x(self, k, stream=None):
        self.wait(self.X_SLOT + k, stream)

    def push_group(self, k, stream=None):
        """Column blocks of dx of every part of group k (of all ranks) -> their owners; signals G_k."""
        from .binding import check, ffemb_table_t, lib
        from .embedding import _stream_ptr
        if self._push[k] is None:
            ps = [p for p in self.parts if p.group == k]
            arr = (ffemb_table_t * max(1, len(ps)))()
            for i, p in enumerate(ps):
                arr[i].out = int(self.hg.buffer_ptrs[p.rank]) + self.g_off[(p.table, p.col0)] * 4
                arr[i].reserved, arr[i].L = p.table * self.D + p.col0, p.width
            # resident warps per SM (hint): the push of a pipelined step runs under the neighbouring group's
            # compute kernels — few warps (measured at N = 2: 1.754 ms per step with 2, 1.841 with the default
            # 4).  At N = 8 the push is bound by the fabric, not by its warps (0.201 ms with 16, 0.204 with 4)
            arr[0].R = 2 if (self.K > 1 and self.world <= 4) else 0
            self._push[k] = (arr, len(ps))
        arr, n = self._push[k]
        if self.local_batch and n:
            check(lib().ffemb_exchange_backward_parts(arr, n, self.dx.data_ptr(), self.local_batch, self.lo,
                                                      self.stride, _stream_ptr(stream)))
        self.signal(self.G_SLOT + k, stream)

    def wait_grads(self, k, stream=None):
        self.wait(self.G_SLOT + k, stream)

    # ---- the un-pipelined calls (K == 1, or every group back to back on one stream) -----------------
    def forward(self, coll, indices, stream=None):
        """coll / indices: one EmbeddingCollection (all owned parts, K == 1) or per-group lists.
        Fills x on all ranks and returns it once every rank's rows have landed."""
        self.begin_step()
        colls = coll if isinstance(coll, (list, tuple)) else [coll]
        idxs = indices if isinstance(coll, (list, tuple)) else [indices]
        for k in range(self.K):
            self.forward_group(k, colls[k], idxs[k], stream)
        for k in range(self.K):
            self.wait_x(k, stream)
        return self.x

    def backward(self, stream=None):
        """dx on every rank -> self.outgrads ([B][w] per owned part) on the owners."""
        for k in range(self.K):
            self.push_group(k, stream)
        for k in range(self.K):
            self.wait_grads(k, stream)
        return self.outgrads


class ShardedStep:
    """One training step of this rank's share of the embedding tables, pipelined over the exchange's
    groups on three streams (the schedule a per-GPU Legion leaf task would enqueue;
    csrc/flexflow/embedding_ffemb.cu has the C++ form):

        main : forward(0) sig  forward(1) sig ...            wait G_0  backward(0)  wait G_1  backward(1) ...
        push :          wait X_0  push(0) sig   wait X_1  push(1) sig ...
        side : keygen + radix sort of group 0, 1, ...   (needs the indices only)

    `dx_ready`: optional event after which this rank's dx holds the step's gradient (with a model in
    between it is recorded after the dense layers' backward; the host-buffer step records it after the
    upload).  The data dependence kept per group: the gradient block of a part is pushed only once that
    part's pooled rows have landed on every rank."""

    def __init__(self, fx, colls, device):
        self.fx, self.colls, self.dev = fx, colls, device
        # the push and the index sort are short, latency-critical pieces that must get SM slots while a
        # long gather-pool / segmented-sum grid is draining: high-priority streams
        import os
        self.s_push = torch.cuda.Stream(device=device, priority=int(os.environ.get("FFEMB_PUSH_PRIO", "-1")))
        self.s_side = torch.cuda.Stream(device=device, priority=int(os.environ.get("FFEMB_SIDE_PRIO", "-1")))
        self.ev_sorted = [torch.cuda.Event() for _ in range(fx.K)]
        self.ev_x = torch.cuda.Event()
        self.trace = None   # set to [] to collect (label, start event, end event) of every piece of a step

    def _span(self, label, stream, fn):
        if self.trace is None:
            return fn()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        r = fn()
        b.record(stream)
        self.trace.append((label, a, b))
        return r

    def trace_ms(self):
        """[(label, start ms, end ms)] relative to the first recorded event (call after a synchronize)."""
        if not self.trace:
            return []
        t0 = self.trace[0][1]
        return [(l, round(t0.elapsed_time(a), 4), round(t0.elapsed_time(b), 4)) for l, a, b in self.trace]

    def run(self, idx_groups, lr, dx_ready=None, overlap=True, stream=None):
        fx, K = self.fx, self.fx.K
        main = stream if stream is not None else torch.cuda.current_stream(self.dev)
        push, side = self.s_push, self.s_side
        fx.begin_step()
        self._span("step_start", main, lambda: None)
        push.wait_stream(main)
        if overlap:
            side.wait_stream(main)   # the previous step's reductions are done with the sort buffers
            for k in range(K):
                if self.colls[k] is not None:
                    self._span(f"sort{k}", side, lambda: self.colls[k].backward_prepare(idx_groups[k], stream=side))
                self.ev_sorted[k].record(side)
        for k in range(K):
            self._span(f"fwd{k}", main, lambda: fx.forward_group(k, self.colls[k], idx_groups[k], main))
        with torch.cuda.stream(push):
            for k in range(K):
                self._span(f"waitx{k}", push, lambda: fx.wait_x(k, push))
                if k == K - 1:
                    self.ev_x.record(push)   # x is complete on this rank
                if dx_ready is not None and k == 0:
                    push.wait_event(dx_ready)
                self._span(f"push{k}", push, lambda: fx.push_group(k, push))
        for k in range(K):
            self._span(f"waitg{k}", main, lambda: fx.wait_grads(k, main))
            if self.colls[k] is not None:
                if overlap:
                    main.wait_event(self.ev_sorted[k])
                self._span(f"bwd{k}", main, lambda: self.colls[k].backward_sgd(
                    idx_groups[k], fx.outgrads_by_group[k], lr, stream=main, prepared=overlap))
        main.wait_stream(push)
        if overlap:
            main.wait_stream(side)   # (a group without parts of this rank leaves its sort event unwaited)
        return fx.x

    def capture(self, idx_groups, lr, parity):
        """The whole pipelined step — three streams, the cross-rank flag kernels included — as ONE CUDA
        graph (ShardedStepGraph).  At 8 GPUs a rank's share of a step is ~50 launches of 5-100 us: enqueued
        one by one from the host they take longer than they run."""
        return ShardedStepGraph(self, idx_groups, lr, parity)


class ShardedStepGraph:
    """ShardedStep.run captured once for one set of index buffers and one parity of the double-buffered
    interaction input; replay() = one graph launch.  Capturable because no entry point synchronises or
    allocates, the flags keep their epochs in the pad (include/ffemb.h, ffemb_sync_*), and the streams
    fork from / join the capture stream through events.  Every rank must replay the same sequence of
    steps; consecutive steps must alternate parity (x(step) may still be read while step + 1 is written).
    Same kernels in the same order as the eager step: bit-identical results."""

    def __init__(self, sstep, idx_groups, lr, parity):
        fx = sstep.fx
        self.sstep, self.parity = sstep, int(parity)
        self.graph = torch.cuda.CUDAGraph()
        keep = fx.parity
        fx.parity = self.parity ^ 1      # begin_step() flips it
        trace, sstep.trace = sstep.trace, None
        with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
            sstep.run(idx_groups, lr, overlap=True)
        fx.parity, sstep.trace = keep, trace

    def replay(self):
        self.sstep.fx.parity = self.parity
        self.graph.replay()
        return self.sstep.fx.x


class RowWiseTable:
    """One table of `rows` rows sharded ROW-WISE over the ranks (for tables beyond one GPU's HBM;
    not in the reference, see include/ffemb.h).  Shard map: sharding.rows_per_shard / row_owner.

    forward     : this rank's samples only; rows are read from the owner's shard over NVLink peer
                  memory inside the gather-pool kernel (bit-identical to the unsharded table).
    backward_sgd: indices and output-grads of the whole batch are all-gathered (the one real
                  exchange of this path: torch.distributed all_gather, NCCL on GPUs), localised
                  (foreign rows -> the scratch row appended to the shard) and the ordinary
                  deterministic backward + SGD runs on the shard.
    """

    def __init__(self, rows, dim, bag, batch, rank, world, device, aggr=21, group=None):
        import torch.distributed._symmetric_memory as symm
        from .embedding import EmbeddingCollection
        from .sharding import rows_per_shard, shard_rows
        self.R, self.D, self.L, self.B = int(rows), int(dim), int(bag), int(batch)
        self.rank, self.world, self.device, self.aggr = rank, world, device, int(aggr)
        self.group = group if group is not None else dist.group.WORLD
        self.rps = rows_per_shard(self.R, world)
        self.local_rows = shard_rows(self.R, world, rank)
        self.ext = (self.B + world - 1) // world
        self.lo, self.hi = batch_block(self.B, world, rank)
        self.shard = symm.empty((self.rps + 1, self.D), dtype=torch.float32, device=device)  # + scratch row
        self.h = symm.rendezvous(self.shard, self.group)
        self.shard.zero_()
        self.idx_full = torch.zeros((world * self.ext, self.L), dtype=torch.int64, device=device)
        self.idx_local = torch.empty((self.B, self.L), dtype=torch.int64, device=device)
        self.g_full = torch.zeros((world * self.ext, self.D), dtype=torch.float32, device=device)
        self._pad_i = torch.zeros((self.ext, self.L), dtype=torch.int64, device=device)
        self._pad_g = torch.zeros((self.ext, self.D), dtype=torch.float32, device=device)
        self.coll = EmbeddingCollection([self.shard], [self.L], self.aggr)
        torch.cuda.synchronize(device)
        dist.barrier(group=self.group)

    def load_rows(self, full_table):
        """Test/bench helper: take this rank's rows out of an unsharded [R][D] tensor."""
        lo = self.rank * self.rps
        self.shard[:self.local_rows].copy_(full_table[lo:lo + self.local_rows])
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def forward(self, idx, out=None, stream=None):
        """idx: int64 [B_h][L] global row ids of this rank's samples -> pooled [B_h][D]."""
        import ctypes as C
        from .binding import check, lib
        from .embedding import _stream_ptr
        n = int(idx.shape[0])
        if out is None:
            out = torch.empty((n, self.D), dtype=torch.float32, device=self.device)
        ptrs = (C.c_void_p * self.world)(*[int(p) for p in self.h.buffer_ptrs])
        check(lib().ffemb_forward_rowwise(idx.data_ptr(), out.data_ptr(), ptrs, self.world, self.rps,
                                          self.L, self.D, n, self.aggr, _stream_ptr(stream)))
        return out

    def backward_sgd(self, idx, outgrad, lr, stream=None):
        """idx [B_h][L], outgrad [B_h][D] of this rank's samples; updates this rank's shard."""
        from .binding import check, lib
        from .embedding import _stream_ptr
        n = self.hi - self.lo
        self._pad_i[:n].copy_(idx)
        self._pad_g[:n].copy_(outgrad)
        dist.all_gather_into_tensor(self.idx_full, self._pad_i, group=self.group)  # also orders every
        dist.all_gather_into_tensor(self.g_full, self._pad_g, group=self.group)    # rank's forward first
        check(lib().ffemb_rowwise_localize(self.idx_full.data_ptr(), self.idx_local.data_ptr(),
                                           self.B * self.L, self.rps, self.rank,
                                           self.shard[self.rps].data_ptr(), self.D, _stream_ptr(stream)))
        self.coll.backward_sgd([self.idx_local], [self.g_full[:self.B]], lr, stream=stream)
        self.h.barrier(channel=0)  # shards are consistent before anyone's next forward reads them


class RowWiseCollection:
    """Several row-wise sharded tables driven together (BASELINE config 4: 8 x 100M x 128), SURVEY 8e.

      forward     : one remote-gather launch per table — rows are read from the owners' shards over NVLink
                    inside the gather-pool kernel; adds in j order on the sample's rank (bit-identical to
                    the unsharded table).
      backward_sgd: ROUTE (ffemb_rowwise_route, one launch for all tables): every lookup of this rank's
                    samples is bucketised by owner = row / ceil(R/G) and its local row number stored into
                    the owner's slot array, the bag's output-grad row into the owner's routed-gradient
                    buffer — remote stores over NVLink, the all-to-all-v of indices + gradient rows done
                    by the kernel itself; slot = position of the lookup in the GLOBAL batch, so nothing
                    is counted and no counts are exchanged (foreign slots hold -1);
                    flag (ffemb_sync_signal / _wait): all ranks' routes have landed;
                    REDUCE (ffemb_backward_sgd_padded, one call for all tables): the ordinary sorted
                    backward + SGD over the owner's slot arrays; empty slots sort to the end and are
                    never loaded — an owner gathers and updates its own lookups only;
                    flag: every shard is updated before anyone's next forward reads it.
    No collective call, no host synchronisation; per rank and step the NVLink carries one output-grad
    row per (bag, owner) pair + 8 bytes per lookup and owner instead of the whole batch's gradient.
    `legacy=True` keeps the round-1 form (two all-gathers, every owner walks the whole batch) as the
    measured comparison."""

    R_SLOT, D_SLOT = 0, 1   # flags: routes landed / shards updated

    def __init__(self, rows, dim, bags, batch, rank, world, device, aggr=21, group=None, legacy=False):
        import torch.distributed._symmetric_memory as symm
        from .binding import ffemb_route_t, ffemb_sync_t, lib
        from .embedding import EmbeddingCollection
        from .sharding import rows_per_shard, shard_rows
        self.rows, self.D, self.bags, self.B = [int(r) for r in rows], int(dim), [int(l) for l in bags], int(batch)
        self.T = len(self.rows)
        self.rank, self.world, self.device, self.aggr = rank, world, device, int(aggr)
        self.group = group if group is not None else dist.group.WORLD
        self.legacy = bool(legacy)
        self.rps = [rows_per_shard(r, world) for r in self.rows]
        self.local_rows = [shard_rows(r, world, rank) for r in self.rows]
        self.ext = (self.B + world - 1) // world
        self.lo, self.hi = batch_block(self.B, world, rank)
        self.shards, self.handles = [], []
        for t in range(self.T):
            s = symm.empty((self.rps[t] + 1, self.D), dtype=torch.float32, device=device)  # (+ legacy scratch row)
            self.handles.append(symm.rendezvous(s, self.group))
            s.zero_()
            self.shards.append(s)
        self.sumL = sum(self.bags)
        i64, f32 = dict(dtype=torch.int64, device=device), dict(dtype=torch.float32, device=device)
        Bp = world * self.ext   # the batch padded to whole blocks: the owners' slot arrays cover it
        self.Bp = Bp
        if self.legacy:
            self._pad_i = torch.zeros((self.ext, self.sumL), **i64)
            self._pad_g = torch.zeros((self.ext, self.T * self.D), **f32)
            self.idx_full = torch.zeros((Bp, self.sumL), **i64)
            self.g_full = torch.zeros((Bp, self.T * self.D), **f32)
            self.idx_glob = [torch.empty((self.B, l), **i64) for l in self.bags]
            self.idx_local = [torch.empty((self.B, l), **i64) for l in self.bags]
            self.coll = EmbeddingCollection(self.shards, self.bags, self.aggr)
        else:
            # slot arrays [T][Bp][L_t] (int64, -1 = empty) and routed gradient rows [T][Bp][D], peer-mapped
            self.slots = symm.empty((max(1, Bp * self.sumL),), **i64)
            self.grows = symm.empty((max(1, self.T * Bp), self.D), **f32)
            self.pad = symm.empty((lib().ffemb_sync_pad_bytes() // 8,), **i64)
            self.hs = symm.rendezvous(self.slots, self.group)
            self.hr = symm.rendezvous(self.grows, self.group)
            self.hp = symm.rendezvous(self.pad, self.group)
            self.slots.fill_(-1)   # (a ragged last block leaves its tail slots untouched: empty)
            self.grows.zero_()
            self.pad.zero_()
            rt = ffemb_route_t()
            sy = ffemb_sync_t()
            for h in range(world):
                rt.peer_slots[h] = int(self.hs.buffer_ptrs[h])
                rt.peer_rows[h] = int(self.hr.buffer_ptrs[h])
                sy.peer_pad[h] = int(self.hp.buffer_ptrs[h])
            rt.num_peers, rt.rank, rt.samples_per_peer = world, rank, self.ext
            sy.num_peers, sy.rank = world, rank
            self._route, self._sync = rt, sy
            off, self.slot_views = 0, []
            for l in self.bags:
                self.slot_views.append(self.slots[off:off + Bp * l].view(Bp, l))
                off += Bp * l
            self.grow_views = [self.grows[t * Bp:(t + 1) * Bp] for t in range(self.T)]
            # the owner's reduce: shards as tables of rps rows (the scratch row stays out), batch = Bp
            self.coll = EmbeddingCollection([s[:self.rps[t]] for t, s in enumerate(self.shards)], self.bags, self.aggr)
            self._dirty = False   # a backward has signalled D_SLOT and nobody has waited on it yet
        torch.cuda.synchronize(device)
        dist.barrier(group=self.group)

    def load_rows(self, full_tables):
        for t, full in enumerate(full_tables):
            lo = self.rank * self.rps[t]
            self.shards[t][:self.local_rows[t]].copy_(full[lo:lo + self.local_rows[t]])
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def _flag(self, fn, slot, stream):
        import ctypes as C
        from .binding import check
        from .embedding import _stream_ptr
        check(fn(C.byref(self._sync), slot, _stream_ptr(stream)))

    def forward(self, idx_list, outs=None, stream=None):
        """idx_list[t]: int64 [B_h][L_t] global row ids of this rank's samples -> pooled [B_h][D] per table."""
        import ctypes as C
        from .binding import check, lib
        from .embedding import _stream_ptr
        n = int(idx_list[0].shape[0])
        if outs is None:
            outs = [torch.empty((n, self.D), dtype=torch.float32, device=self.device) for _ in range(self.T)]
        if not self.legacy and self._dirty:   # every rank's shards carry the previous step's update
            self._flag(lib().ffemb_sync_wait, self.D_SLOT, stream)
            self._dirty = False
        for t in range(self.T):
            ptrs = (C.c_void_p * self.world)(*[int(p) for p in self.handles[t].buffer_ptrs])
            check(lib().ffemb_forward_rowwise(idx_list[t].data_ptr(), outs[t].data_ptr(), ptrs, self.world,
                                              self.rps[t], self.bags[t], self.D, n, self.aggr, _stream_ptr(stream)))
        return outs

    def backward_sgd(self, idx, dx, lr, stream=None):
        """dx: [B_h][T*D] — the gradient of this rank's interaction input (table t's block at columns
        [t*D, +D)); idx: the per-table index tensors [B_h][L_t] of its samples (legacy form: ONE
        sample-major int64 [B_h][sum L] tensor).  Updates this rank's shards of every table."""
        from .binding import check, ffemb_table_t, lib
        from .embedding import _stream_ptr, load_sparse_inputs
        n = self.hi - self.lo
        if self.legacy:
            self._pad_i[:n].copy_(idx)
            self._pad_g[:n].copy_(dx)
            dist.all_gather_into_tensor(self.idx_full, self._pad_i, group=self.group)  # also orders every
            dist.all_gather_into_tensor(self.g_full, self._pad_g, group=self.group)    # rank's forward first
            load_sparse_inputs(self.idx_full, self.bags, self.B, outs=self.idx_glob, stream=stream)
            for t in range(self.T):
                check(lib().ffemb_rowwise_localize(self.idx_glob[t].data_ptr(), self.idx_local[t].data_ptr(),
                                                   self.B * self.bags[t], self.rps[t], self.rank,
                                                   self.shards[t][self.rps[t]].data_ptr(), self.D, _stream_ptr(stream)))
            self.coll.backward_sgd_concat(self.idx_local, self.g_full, lr, stream=stream)
            self.handles[0].barrier(channel=0)  # shards are consistent before anyone's next forward reads them
            return
        import ctypes as C
        arr = (ffemb_table_t * self.T)()
        for t in range(self.T):
            arr[t].idx, arr[t].R, arr[t].L, arr[t].reserved = idx[t].data_ptr(), self.rows[t], self.bags[t], t * self.D
        if self._dirty:   # (two backwards in a row without a forward: the owners are done with the last routes)
            self._flag(lib().ffemb_sync_wait, self.D_SLOT, stream)
        check(lib().ffemb_rowwise_route(arr, self.T, self.D, n, dx.data_ptr(), int(dx.shape[1]),
                                        C.byref(self._route), _stream_ptr(stream)))
        self._flag(lib().ffemb_sync_signal, self.R_SLOT, stream)
        self._flag(lib().ffemb_sync_wait, self.R_SLOT, stream)
        self.coll.backward_sgd_padded(self.slot_views, self.grow_views, lr, stream=stream)
        self._flag(lib().ffemb_sync_signal, self.D_SLOT, stream)
        self._dirty = True
