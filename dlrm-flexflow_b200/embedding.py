"""Host-side mirror of the reference Embedding operator for this path.

Names and argument meaning follow /root/reference/include/model.h:1136-1202:
  Embedding.forward_kernel / backward_kernel  — the static raw-pointer wrappers (the seam);
  EmbeddingMeta                                — per-shard state made once by init_task
                                                 (src/ops/embedding.cu:104-113), here it also
                                                 owns the backward workspace;
  Embedding(num_entries, out_dim, aggr)        — the op: init / forward / backward.
EmbeddingCollection is the table-batched form (all tables of one GPU per launch).
Everything computes in libffemb.so; tensors are torch CUDA tensors used as device buffers.
"""
import ctypes as C

import torch

from . import binding as B
from .binding import AGGR_MODE_SUM, check, ffemb_table_t, lib


def _stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def _req(t, dtype, name):
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == dtype and t.is_contiguous()):
        raise TypeError(f"{name}: expected a contiguous CUDA tensor of {dtype}")
    return t


class EmbeddingMeta:
    """OpMeta analogue: aggr + a persistent workspace (no allocation on the hot path)."""

    def __init__(self, aggr, profiling=False):
        self.aggr = int(aggr)
        self.profiling = bool(profiling)
        self.workspace = None

    def ensure_workspace(self, num_lookups, total_rows, dim, device):
        need = lib().ffemb_backward_workspace(int(num_lookups), int(total_rows), int(dim))
        if need == 0 and num_lookups > 0:
            raise B.FFEmbError(B.FFEMB_EINVAL, "invalid workspace query")
        if self.workspace is None or self.workspace.numel() < need or self.workspace.device != device:
            self.workspace = torch.empty(max(need, 256), dtype=torch.uint8, device=device)
        return self.workspace


class Embedding:
    """One table.  Mirrors class Embedding (include/model.h:1136-1196)."""

    def __init__(self, num_entries, out_dim, aggr=AGGR_MODE_SUM, name=None):
        self.num_entries, self.out_channels, self.aggr = int(num_entries), int(out_dim), int(aggr)
        self.name = name
        self.meta = None

    # -- the seam: same argument order as the reference's static wrappers --------------------
    @staticmethod
    def forward_kernel(input_ptr, output_ptr, weight_ptr, in_dim, out_dim, batch_size, aggr,
                       outputSize=None, stream=None):
        """Embedding::forward_kernel (src/ops/embedding.cu:219-231). Pointers are ints."""
        if outputSize is not None and outputSize != batch_size * out_dim:
            raise ValueError("outputSize must equal batch_size * out_dim")
        check(lib().ffemb_forward(input_ptr, output_ptr, weight_ptr, in_dim, out_dim, batch_size,
                                  int(aggr), _stream_ptr(stream)))

    @staticmethod
    def backward_kernel(input_ptr, output_ptr, weight_grad_ptr, in_dim, out_dim, batch_size, aggr,
                        outputSize=None, stream=None):
        """Embedding::backward_kernel (src/ops/embedding.cu:308-320): unordered atomic scatter
        into the dense, pre-zeroed weight-grad — the API-exact comparison variant."""
        if outputSize is not None and outputSize != batch_size * out_dim:
            raise ValueError("outputSize must equal batch_size * out_dim")
        check(lib().ffemb_backward_dense_atomic(input_ptr, output_ptr, weight_grad_ptr, in_dim,
                                                out_dim, batch_size, int(aggr), _stream_ptr(stream)))

    # -- op-level API --------------------------------------------------------------------------
    def init(self):
        """Embedding::init -> init_task: make the per-shard meta."""
        self.meta = EmbeddingMeta(self.aggr)
        return self.meta

    def _shapes(self, input, weight):
        _req(input, torch.int64, "input")
        _req(weight, torch.float32, "weight")
        if input.dim() != 2 or weight.dim() != 2:
            raise ValueError("input must be [B][L], weight [R][D]")
        if weight.shape[0] != self.num_entries or weight.shape[1] != self.out_channels:
            raise ValueError("weight shape does not match (num_entries, out_dim)")
        return input.shape[0], input.shape[1]

    def forward(self, input, weight, output=None, stream=None):
        Bsz, L = self._shapes(input, weight)
        if output is None:
            output = torch.empty((Bsz, self.out_channels), dtype=torch.float32, device=weight.device)
        _req(output, torch.float32, "output")
        if tuple(output.shape) != (Bsz, self.out_channels):
            raise ValueError("output shape must be [B][D]")  # the task body's asserts
        self.forward_kernel(input.data_ptr(), output.data_ptr(), weight.data_ptr(), L,
                            self.out_channels, Bsz, self.aggr, stream=stream)
        return output

    def backward(self, input, output_grad, weight_grad, deterministic=True, stream=None):
        """Unfused, API-exact: accumulate into the dense weight-grad."""
        Bsz, L = self._shapes(input, weight_grad)
        _req(output_grad, torch.float32, "output_grad")
        if not deterministic:
            self.backward_kernel(input.data_ptr(), output_grad.data_ptr(), weight_grad.data_ptr(),
                                 L, self.out_channels, Bsz, self.aggr, stream=stream)
            return weight_grad
        if self.meta is None:
            self.init()
        ws = self.meta.ensure_workspace(Bsz * L, self.num_entries, self.out_channels, weight_grad.device)
        check(lib().ffemb_backward_dense(input.data_ptr(), output_grad.data_ptr(),
                                         weight_grad.data_ptr(), L, self.out_channels, Bsz,
                                         self.num_entries, self.aggr, ws.data_ptr(), ws.numel(),
                                         _stream_ptr(stream)))
        return weight_grad

    def backward_sgd(self, input, output_grad, weight, lr, stream=None):
        """Fused backward + sparse SGD row update (momentum = weight_decay = 0)."""
        Bsz, L = self._shapes(input, weight)
        _req(output_grad, torch.float32, "output_grad")
        if self.meta is None:
            self.init()
        ws = self.meta.ensure_workspace(Bsz * L, self.num_entries, self.out_channels, weight.device)
        check(lib().ffemb_backward_sgd(input.data_ptr(), output_grad.data_ptr(), weight.data_ptr(),
                                       L, self.out_channels, Bsz, self.num_entries, self.aggr,
                                       float(lr), ws.data_ptr(), ws.numel(), _stream_ptr(stream)))
        return weight

    def backward_sgd_static(self, input, output_grad, weight, lr, stream=None):
        """backward_sgd with the scratch in the library's file-scope device array (ffemb_backward_sgd_static,
        no counterpart in the reference): nothing is allocated — no workspace tensor either — so it can be
        called where allocation is forbidden.  Shapes beyond FFEMB_STATIC_SCRATCH_BYTES raise FFEmbError
        (FFEMB_EWORKSPACE); calls that overlap on one device race."""
        Bsz, L = self._shapes(input, weight)
        _req(output_grad, torch.float32, "output_grad")
        check(lib().ffemb_backward_sgd_static(input.data_ptr(), output_grad.data_ptr(), weight.data_ptr(),
                                              L, self.out_channels, Bsz, self.num_entries, self.aggr,
                                              float(lr), _stream_ptr(stream)))
        return weight


class EmbeddingCollection:
    """All tables that live on one GPU, driven through the table-batched entry points."""

    def __init__(self, weights, bags, aggr=AGGR_MODE_SUM):
        if len(weights) == 0 or len(weights) != len(bags):
            raise ValueError("need one bag width per table")
        self.weights = [_req(w, torch.float32, "weight") for w in weights]
        self.dim = int(weights[0].shape[1])
        if any(w.dim() != 2 or w.shape[1] != self.dim for w in weights):
            raise ValueError("all tables of a collection share the embedding dim")
        self.bags = [int(l) for l in bags]
        self.aggr = int(aggr)
        self.device = weights[0].device
        self.meta = EmbeddingMeta(aggr)
        self.total_rows = sum(int(w.shape[0]) for w in weights)

    def _descs(self, indices, outs, tables, batch):
        T = len(self.weights)
        arr = (ffemb_table_t * T)()
        for t in range(T):
            _req(indices[t], torch.int64, "indices")
            _req(outs[t], torch.float32, "out")
            if tuple(indices[t].shape) != (batch, self.bags[t]) or tuple(outs[t].shape) != (batch, self.dim):
                raise ValueError(f"table {t}: expected indices [{batch}][{self.bags[t]}] and out [{batch}][{self.dim}]")
            arr[t].idx, arr[t].out, arr[t].table = indices[t].data_ptr(), outs[t].data_ptr(), tables[t].data_ptr()
            arr[t].R, arr[t].L = int(tables[t].shape[0]), self.bags[t]
        return arr

    def _chunks(self):
        T = len(self.weights)
        return [(s, min(T, s + B.FFEMB_MAX_TABLES)) for s in range(0, T, B.FFEMB_MAX_TABLES)]

    def forward(self, indices, outs=None, stream=None):
        batch = int(indices[0].shape[0])
        if outs is None:
            outs = [torch.empty((batch, self.dim), dtype=torch.float32, device=self.device)
                    for _ in self.weights]
        arr = self._descs(indices, outs, self.weights, batch)
        for s, e in self._chunks():
            sub = C.cast(C.byref(arr, s * C.sizeof(ffemb_table_t)), C.POINTER(ffemb_table_t))
            check(lib().ffemb_forward_batched(sub, e - s, self.dim, batch, self.aggr, _stream_ptr(stream)))
        return outs

    def _workspaces(self, batch):
        """One persistent workspace per chunk of <= 64 tables (prepare/apply pairs share it)."""
        if not hasattr(self, "_ws") or self._ws_batch != batch:
            self._ws, self._ws_batch = [], batch
            for s, e in self._chunks():
                n = batch * sum(self.bags[s:e])
                rows = sum(int(w.shape[0]) for w in self.weights[s:e])
                if self.aggr != AGGR_MODE_SUM:   # room for the once-per-bag quotients of the AVG backward
                    need = lib().ffemb_backward_workspace_avg(int(n), int(rows), self.dim, int(batch), e - s)
                else:
                    need = lib().ffemb_backward_workspace(int(n), int(rows), self.dim)
                self._ws.append(torch.empty(max(need, 256), dtype=torch.uint8, device=self.device))
        return self._ws

    def backward_prepare(self, indices, stream=None):
        """Keygen + radix sort of the lookups (needs the indices only): enqueue it early, e.g. on a
        side stream under the forward pass / exchange; then call backward_sgd(..., prepared=True)."""
        batch = int(indices[0].shape[0])
        T = len(self.weights)
        arr = (ffemb_table_t * T)()
        for t in range(T):
            _req(indices[t], torch.int64, "indices")
            arr[t].idx, arr[t].out, arr[t].table = indices[t].data_ptr(), None, self.weights[t].data_ptr()
            arr[t].R, arr[t].L = int(self.weights[t].shape[0]), self.bags[t]
        for k, (s, e) in enumerate(self._chunks()):
            ws = self._workspaces(batch)[k]
            sub = C.cast(C.byref(arr, s * C.sizeof(ffemb_table_t)), C.POINTER(ffemb_table_t))
            check(lib().ffemb_backward_prepare(sub, e - s, self.dim, batch, self.aggr, ws.data_ptr(),
                                               ws.numel(), _stream_ptr(stream)))

    def backward_sgd(self, indices, out_grads, lr, stream=None, prepared=False):
        batch = int(indices[0].shape[0])
        arr = self._descs(indices, out_grads, self.weights, batch)
        fn = lib().ffemb_backward_sgd_prepared if prepared else lib().ffemb_backward_sgd_batched
        for k, (s, e) in enumerate(self._chunks()):
            ws = self._workspaces(batch)[k]
            sub = C.cast(C.byref(arr, s * C.sizeof(ffemb_table_t)), C.POINTER(ffemb_table_t))
            check(fn(sub, e - s, self.dim, batch, self.aggr, float(lr), ws.data_ptr(), ws.numel(),
                     _stream_ptr(stream)))

    def backward_sgd_padded(self, indices, out_grads, lr, stream=None):
        """backward_sgd where an index < 0 is an EMPTY SLOT (the owner's side of the row-wise backward,
        ffemb_backward_sgd_padded): skipped without loading anything."""
        batch = int(indices[0].shape[0])
        if len(self.weights) > B.FFEMB_MAX_TABLES:
            raise ValueError("at most 64 tables per call")
        arr = self._descs(indices, out_grads, self.weights, batch)
        ws = self._workspaces(batch)[0]
        check(lib().ffemb_backward_sgd_padded(arr, len(self.weights), self.dim, batch, self.aggr, float(lr),
                                              ws.data_ptr(), ws.numel(), _stream_ptr(stream)))

    def backward_sgdx(self, indices, out_grads, lr, weight_decay=0.0, momentum=0.0, nesterov=False,
                      V=None, stream=None):
        """Backward fused with the reference's full sgd_update (weight decay / momentum / nesterov,
        /root/reference/src/runtime/optimizer_kernel.cu:23-41) — every row of every table changes.
        V: momentum buffers ([R][D] per table, zero-initialised by the caller like
        SGDOptimizer::init, src/runtime/optimizer.cc:50-87); required when momentum > 0."""
        from .binding import ffemb_sgd_t
        batch = int(indices[0].shape[0])
        T = len(self.weights)
        if T > B.FFEMB_MAX_TABLES:
            raise ValueError("at most 64 tables per call")
        arr = self._descs(indices, out_grads, self.weights, batch)
        if not hasattr(self, "_touched"):
            self._touched = [torch.zeros(max(1, int(lib().ffemb_sgd_touched_words(int(w.shape[0])))),
                                         dtype=torch.int32, device=self.device) for w in self.weights]
        opt = ffemb_sgd_t(float(lr), float(weight_decay), float(momentum), int(bool(nesterov)))
        tp = (C.c_void_p * T)(*[t.data_ptr() for t in self._touched])
        vp = None
        if momentum > 0:
            if V is None or len(V) != T:
                raise ValueError("momentum > 0 needs one momentum buffer per table")
            vp = (C.c_void_p * T)(*[_req(v, torch.float32, "V").data_ptr() for v in V])
        ws = self._workspaces(batch)[0]
        check(lib().ffemb_backward_sgdx_batched(arr, T, self.dim, batch, self.aggr, C.byref(opt), vp, tp,
                                                ws.data_ptr(), ws.numel(), _stream_ptr(stream)))

    def backward_adam(self, indices, out_grads, alpha_t, M, V, beta1=0.9, beta2=0.999, weight_decay=0.0,
                      epsilon=1e-8, stream=None):
        """Backward fused with the reference's adam_update
        (/root/reference/src/runtime/optimizer_kernel.cu:206-226).  alpha_t: the bias-corrected step
        of this iteration (AdamOptimizer::next); M, V: [R][D] moment buffers per table."""
        from .binding import ffemb_adam_t
        batch = int(indices[0].shape[0])
        T = len(self.weights)
        if T > B.FFEMB_MAX_TABLES or len(M) != T or len(V) != T:
            raise ValueError("one M and one V buffer per table, at most 64 tables per call")
        arr = self._descs(indices, out_grads, self.weights, batch)
        if not hasattr(self, "_touched"):
            self._touched = [torch.zeros(max(1, int(lib().ffemb_sgd_touched_words(int(w.shape[0])))),
                                         dtype=torch.int32, device=self.device) for w in self.weights]
        opt = ffemb_adam_t(float(alpha_t), float(beta1), float(beta2), float(weight_decay), float(epsilon))
        tp = (C.c_void_p * T)(*[t.data_ptr() for t in self._touched])
        mp = (C.c_void_p * T)(*[_req(m, torch.float32, "M").data_ptr() for m in M])
        vp = (C.c_void_p * T)(*[_req(v, torch.float32, "V").data_ptr() for v in V])
        ws = self._workspaces(batch)[0]
        check(lib().ffemb_backward_adam_batched(arr, T, self.dim, batch, self.aggr, C.byref(opt), mp, vp, tp,
                                                ws.data_ptr(), ws.numel(), _stream_ptr(stream)))

    # ---- straight into / out of the Concat tensors (one GPU; ffemb_forward_concat / _backward_sgd_concat)
    def _concat_descs(self, indices, batch, cols):
        T = len(self.weights)
        if T > B.FFEMB_MAX_TABLES:
            raise ValueError("the Concat-direct calls take at most 64 tables")
        arr = (ffemb_table_t * T)()
        for t in range(T):
            _req(indices[t], torch.int64, "indices")
            if tuple(indices[t].shape) != (batch, self.bags[t]):
                raise ValueError(f"table {t}: expected indices [{batch}][{self.bags[t]}]")
            arr[t].idx, arr[t].out, arr[t].table = indices[t].data_ptr(), None, self.weights[t].data_ptr()
            arr[t].R, arr[t].L, arr[t].reserved = int(self.weights[t].shape[0]), self.bags[t], int(cols[t])
        return arr

    def forward_concat(self, indices, x, cols=None, stream=None):
        """Pooled rows stored straight into the Concat output x [B][row_stride]
        (/root/reference/src/ops/concat.cu:211-251): table t at columns [cols[t], cols[t] + D)."""
        batch = int(indices[0].shape[0])
        _req(x, torch.float32, "x")
        cols = [t * self.dim for t in range(len(self.weights))] if cols is None else cols
        arr = self._concat_descs(indices, batch, cols)
        check(lib().ffemb_forward_concat(arr, len(self.weights), self.dim, batch, self.aggr, x.data_ptr(),
                                         int(x.shape[1]), _stream_ptr(stream)))
        return x

    def backward_sgd_concat(self, indices, dx, lr, cols=None, stream=None, prepared=False):
        """Backward + SGD reading every table's output-grad in place from the gradient of the Concat
        output dx [B][row_stride] (replaces the add_with_stride pass of concat.cu:325-366 and the
        zero-fill of the per-table output-grad regions)."""
        batch = int(indices[0].shape[0])
        _req(dx, torch.float32, "dx")
        cols = [t * self.dim for t in range(len(self.weights))] if cols is None else cols
        arr = self._concat_descs(indices, batch, cols)
        ws = self._workspaces(batch)[0]
        check(lib().ffemb_backward_sgd_concat(arr, len(self.weights), self.dim, batch, self.aggr, float(lr),
                                              dx.data_ptr(), int(dx.shape[1]), int(prepared), ws.data_ptr(),
                                              ws.numel(), _stream_ptr(stream)))

    def backward_dense(self, indices, out_grads, weight_grads, stream=None):
        batch = int(indices[0].shape[0])
        arr = self._descs(indices, out_grads, weight_grads, batch)
        for s, e in self._chunks():
            n = batch * sum(self.bags[s:e])
            rows = sum(int(w.shape[0]) for w in weight_grads[s:e])
            ws = self.meta.ensure_workspace(n, rows, self.dim, self.device)
            sub = C.cast(C.byref(arr, s * C.sizeof(ffemb_table_t)), C.POINTER(ffemb_table_t))
            check(lib().ffemb_backward_dense_batched(sub, e - s, self.dim, batch, self.aggr,
                                                     ws.data_ptr(), ws.numel(), _stream_ptr(stream)))


class StepGraph:
    """One forward + backward-with-SGD step of a collection captured into a CUDA graph: for
    launch-bound steps (small batches: a dozen kernels of a few microseconds each) a replay is one
    launch instead of the dozen.  The entry points qualify because they neither synchronise nor
    allocate, and the side-stream fork of the small-table kernels is a fork/join on events
    (include/ffemb.h, conventions) — the captured step is the same kernels in the same order, so
    its results are bit-identical to the eager calls'.

    Buffers are static: write the step's indices into `self.idx[t]` and the output-grads into
    `self.out_grads[t]` (or let the caller's kernels write there), then `replay()`; pooled rows
    are in `self.outs[t]`.  `between` (optional callable) is captured between the two halves —
    the place of the caller's own work on the capture stream (interaction, dense layers).
    (The sgdx / adam entry points are not capturable: their scalar arguments change every step.)"""

    def __init__(self, coll, batch, lr, idx=None, outs=None, out_grads=None, between=None, overlap_sort=True):
        dev = coll.device
        self.coll, self.batch, self.lr = coll, int(batch), float(lr)
        new = lambda shape, dt: torch.zeros(shape, dtype=dt, device=dev)
        self.idx = idx if idx is not None else [new((batch, l), torch.int64) for l in coll.bags]
        self.outs = outs if outs is not None else [new((batch, coll.dim), torch.float32) for _ in coll.bags]
        self.out_grads = out_grads if out_grads is not None else \
            [new((batch, coll.dim), torch.float32) for _ in coll.bags]
        # one eager step first: workspace allocation, the lazily created side stream and the kernel
        # attributes all happen outside the capture (lr = 0: no table value changes)
        coll.forward(self.idx, self.outs)
        coll.backward_sgd(self.idx, self.out_grads, 0.0)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream(device=dev) if overlap_sort else None
        with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
            main = torch.cuda.current_stream(dev)
            if side is not None:
                # the index sort needs the indices only: a second branch of the graph, next to the forward
                # pass (and the caller's work in `between`); the segmented sum joins it (at these sizes the
                # sort is one CTA per table and the forward a few dozen CTAs: they run side by side)
                side.wait_stream(main)
                coll.backward_prepare(self.idx, stream=side)
            coll.forward(self.idx, self.outs)
            if between is not None:
                between(self)
            if side is not None:
                main.wait_stream(side)
            coll.backward_sgd(self.idx, self.out_grads, self.lr, prepared=side is not None)

    def replay(self):
        self.graph.replay()
        return self.outs


def load_sparse_inputs(full, bags, batch, first_sample=0, sample_ids=None, cols=None, outs=None, stream=None):
    """DataLoader::load_sparse_input for ALL tables in one launch
    (/root/reference/examples/cpp/DLRM/dlrm.cu:19-58, dlrm.cc:482-520; ffemb_load_sparse_inputs).

    full: int64 [num_samples][row_len] — a CUDA tensor, or a pinned CPU tensor (the reference keeps
    the dataset in zero-copy host memory; the kernel then reads it over PCIe itself).
    bags[t]: bag width of table t; cols[t]: first column of its block (default: blocks back to back,
    the reference's my_input_idx * in_dim).  Returns the per-table [batch][L] CUDA index tensors."""
    if full.dtype != torch.int64 or full.dim() != 2 or not full.is_contiguous():
        raise ValueError("full must be a contiguous int64 [num_samples][row_len] tensor")
    if not full.is_cuda and not full.is_pinned():
        raise ValueError("a host-side dataset must be pinned (zero-copy) memory")
    dev = torch.device("cuda", torch.cuda.current_device()) if not full.is_cuda else full.device
    T = len(bags)
    if cols is None:
        cols, c = [], 0
        for l in bags:
            cols.append(c)
            c += int(l)
    if outs is None:
        outs = [torch.empty((batch, int(l)), dtype=torch.int64, device=dev) for l in bags]
    arr = (ffemb_table_t * T)()
    for t in range(T):
        arr[t].idx, arr[t].L, arr[t].reserved = outs[t].data_ptr(), int(bags[t]), int(cols[t])
    ids = None
    if sample_ids is not None:
        ids = _req(sample_ids, torch.int32, "sample_ids").data_ptr()
    for s in range(0, T, B.FFEMB_MAX_TABLES):
        e = min(T, s + B.FFEMB_MAX_TABLES)
        sub = C.cast(C.byref(arr, s * C.sizeof(ffemb_table_t)), C.POINTER(ffemb_table_t))
        check(lib().ffemb_load_sparse_inputs(sub, e - s, full.data_ptr(), int(full.shape[1]), ids,
                                             int(first_sample), int(batch), _stream_ptr(stream)))
    return outs
