"""oracle/oracle.py — TEST INFRASTRUCTURE: numpy/ctypes front-end of the parity checkers.

Three libraries, none of them product code:
  libffemb_oracle.so            plain-C restatement (oracle/ffemb_oracle.c), always available
  _ref/libffref_cpu[_ndebug].so the reference's own src/ops/embedding.cc, compiled unmodified
  _ref/libffref_gpu.so          the reference's own src/ops/embedding.cu (+ zero-fill, sgd_update)
                                for sm_100a; device pointers only (GPU box)
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module.  Reference line numbers are cited in the C sources.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
AGGR_NONE, AGGR_SUM, AGGR_AVG, AGGR_MEAN = 20, 21, 22, 122

_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_i64, _int, _f32 = C.c_int64, C.c_int, C.c_float


def build(ref=True):
    """make oracle (+ ref when /root/reference is present). Returns the list of built libs."""
    targets = ["oracle"]
    if ref and os.path.isdir("/root/reference/src/ops"):
        targets.append("ref")
        if os.path.exists(os.path.join(HERE, "..", "dlrm-flexflow_b200", "libffemb.so")):
            targets.append("dropin")
    subprocess.run(["make", "-s", "-C", HERE] + targets, check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
    return [p for p in (os.path.join(HERE, "libffemb_oracle.so"),
                        os.path.join(HERE, "_ref", "libffref_cpu.so"),
                        os.path.join(HERE, "_ref", "libffref_cpu_ndebug.so"),
                        os.path.join(HERE, "_ref", "libffref_gpu.so"),
                        os.path.join(HERE, "_ref", "libffdropin_gpu.so")) if os.path.exists(p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(HERE, "libffemb_oracle.so")
        if not os.path.exists(path):
            build(ref=False)
        L = C.CDLL(path)
        L.oracle_embed_forward.argtypes = [_i64p, _f32p, _f32p, _i64, _i64, _i64, _int, _i64]
        L.oracle_embed_backward.argtypes = [_i64p, _f32p, _f32p, _i64, _i64, _i64, _int, _i64]
        L.oracle_embed_backward_f64.argtypes = [_i64p, _f32p, _f64p, _i64, _i64, _i64, _int, _i64]
        L.oracle_sgd_update.argtypes = [_i64, _f32, _f32, _f32, _int, _f32p, C.c_void_p, _f32p, _int]
        L.oracle_sgd_update.restype = None
        L.oracle_adam_update.argtypes = [_i64, _f32, _f32, _f32, _f32, _f32, _f32p, _f32p, _f32p, _f32p]
        L.oracle_adam_update.restype = None
        L.oracle_reference_step.argtypes = [_i64p, _f32p, _f32p, _f32p, _i64, _i64, _i64, _i64,
                                            _int, _f32, _int]
        L.oracle_table_owner.argtypes = [_i64, _i64]
        L.oracle_batch_block.argtypes = [_i64, _i64, _i64, C.POINTER(_i64), C.POINTER(_i64)]
        L.oracle_batch_block.restype = None
        L.oracle_rows_per_shard.argtypes = [_i64, _i64]
        L.oracle_rows_per_shard.restype = _i64
        L.oracle_row_owner.argtypes = [_i64, _i64, _i64]
        L.oracle_row_local.argtypes = [_i64, _i64, _i64]
        L.oracle_row_local.restype = _i64
        L.oracle_route_rows.argtypes = [_i64p, _i64, _i64, _i64, _i64p, _i64p, _i64p]
        L.oracle_count_unique.argtypes = [_i64p, _i64]
        L.oracle_count_unique.restype = _i64
        L.oracle_load_sparse_input.argtypes = [_i64p, _i64, _i64, _i64, C.c_void_p, _i64, _i64, _i64p]
        L.oracle_load_sparse_input.restype = None
        L.oracle_concat_forward.argtypes = [_f32p, C.POINTER(C.c_void_p), _i64p, _i64, _i64]
        L.oracle_concat_forward.restype = None
        L.oracle_concat_backward.argtypes = [_f32p, C.POINTER(C.c_void_p), _i64p, _i64, _i64]
        L.oracle_concat_backward.restype = None
        _lib = L
    return _lib


def _chk(rc):
    if rc == 1:
        raise ValueError("oracle: invalid aggr mode (reference asserts aggr == AGGR_MODE_AVG)")
    if rc == 2:
        raise IndexError("oracle: index out of range")


def _shape(idx, W):
    idx = np.ascontiguousarray(idx, np.int64)
    assert idx.ndim == 2 and W.ndim == 2
    return idx, idx.shape[0], idx.shape[1], W.shape[0], W.shape[1]


def forward(idx, W, aggr=AGGR_SUM, check_range=True):
    """embed_forward (src/ops/embedding.cu:166-190). idx [B][L] int64, W [R][D] f32 -> [B][D]."""
    W = np.ascontiguousarray(W, np.float32)
    idx, B, L, R, D = _shape(idx, W)
    out = np.empty((B, D), np.float32)
    _chk(lib().oracle_embed_forward(idx, out, W, L, D, B, aggr, R if check_range else 0))
    return out


def backward_dense(idx, outgrad, R, aggr=AGGR_SUM, Wgrad=None):
    """embed_backward (src/ops/embedding.cu:192-217) into a dense [R][D] gradient."""
    idx = np.ascontiguousarray(idx, np.int64)
    outgrad = np.ascontiguousarray(outgrad, np.float32)
    B, L = idx.shape
    D = outgrad.shape[1]
    if Wgrad is None:
        Wgrad = np.zeros((R, D), np.float32)
    _chk(lib().oracle_embed_backward(idx, outgrad, Wgrad, L, D, B, aggr, R))
    return Wgrad


def backward_dense_f64(idx, outgrad, R, aggr=AGGR_SUM):
    idx = np.ascontiguousarray(idx, np.int64)
    outgrad = np.ascontiguousarray(outgrad, np.float32)
    B, L = idx.shape
    D = outgrad.shape[1]
    Wgrad = np.zeros((R, D), np.float64)
    _chk(lib().oracle_embed_backward_f64(idx, outgrad, Wgrad, L, D, B, aggr, R))
    return Wgrad


def sgd_update(W, Wgrad, lr, weight_decay=0.0, use_fma=True):
    """sgd_update (src/runtime/optimizer_kernel.cu:23-41), momentum 0. Returns the new table."""
    W = np.array(W, np.float32, order="C", copy=True)
    Wgrad = np.ascontiguousarray(Wgrad, np.float32)
    lib().oracle_sgd_update(W.size, lr, weight_decay, 0.0, 0, Wgrad.reshape(-1), None,
                            W.reshape(-1), int(use_fma))
    return W


def sgd_update_full(W, Wgrad, V, lr, weight_decay=0.0, momentum=0.0, nesterov=False, use_fma=True):
    """sgd_update with momentum (src/runtime/optimizer_kernel.cu:23-41). Returns (new W, new V)."""
    W = np.array(W, np.float32, order="C", copy=True)
    V = np.array(V, np.float32, order="C", copy=True)
    Wgrad = np.ascontiguousarray(Wgrad, np.float32)
    lib().oracle_sgd_update(W.size, lr, weight_decay, momentum, int(nesterov), Wgrad.reshape(-1),
                            V.ctypes.data, W.reshape(-1), int(use_fma))
    return W, V


def adam_update(W, Wgrad, M, V, alpha_t, beta1=0.9, beta2=0.999, weight_decay=0.0, epsilon=1e-8):
    """adam_update (src/runtime/optimizer_kernel.cu:206-226). Returns (new W, new M, new V)."""
    W, M, V = (np.array(x, np.float32, order="C", copy=True) for x in (W, M, V))
    Wgrad = np.ascontiguousarray(Wgrad, np.float32)
    lib().oracle_adam_update(W.size, alpha_t, beta1, beta2, weight_decay, epsilon, Wgrad.reshape(-1),
                             M.reshape(-1), V.reshape(-1), W.reshape(-1))
    return W, M, V


def adam_alpha_t(alpha, beta1, beta2, step):
    """AdamOptimizer::next (src/runtime/optimizer.cc:248-254) applied `step` times, in double like there."""
    b1t = b2t = 1.0
    a = alpha
    for _ in range(step):
        b1t *= beta1
        b2t *= beta2
        a = alpha * np.sqrt(1 - b2t) / (1 - b1t)
    return float(a)


def reference_step(idx, outgrad, W, aggr=AGGR_SUM, lr=0.01, use_fma=True):
    """zero-fill -> embed_backward -> sgd_update: the table after one reference step."""
    W = np.array(W, np.float32, order="C", copy=True)
    idx, B, L, R, D = _shape(idx, W)
    outgrad = np.ascontiguousarray(outgrad, np.float32)
    scratch = np.empty((R, D), np.float32)
    _chk(lib().oracle_reference_step(idx, outgrad, W, scratch, L, D, B, R, aggr, lr, int(use_fma)))
    return W


def table_owner(t, G):
    return lib().oracle_table_owner(t, G)


def batch_block(B, G, g):
    lo, hi = _i64(), _i64()
    lib().oracle_batch_block(B, G, g, C.byref(lo), C.byref(hi))
    return lo.value, hi.value


def rows_per_shard(R, G):
    return lib().oracle_rows_per_shard(R, G)


def route_rows(idx_flat, R, G):
    idx_flat = np.ascontiguousarray(idx_flat, np.int64).reshape(-1)
    n = idx_flat.size
    offsets = np.zeros(G + 1, np.int64)
    perm = np.empty(n, np.int64)
    local = np.empty(n, np.int64)
    _chk(lib().oracle_route_rows(idx_flat, n, R, G, offsets, perm, local))
    return offsets, perm, local


def load_sparse_input(full, col, width, batch, first_sample=0, sample_ids=None):
    """DataLoader::load_sparse_input (examples/cpp/DLRM/dlrm.cu:19-58): the [batch][width] index
    tensor of one table cut out of the sample-major dataset full [num_samples][row_len]."""
    full = np.ascontiguousarray(full, np.int64)
    out = np.empty((batch, width), np.int64)
    ids = None
    if sample_ids is not None:
        ids = np.ascontiguousarray(sample_ids, np.int32)
    lib().oracle_load_sparse_input(full, full.shape[1], col, width,
                                   ids.ctypes.data if ids is not None else None, first_sample, batch, out)
    return out


def count_unique(idx):
    idx = np.ascontiguousarray(idx, np.int64).reshape(-1)
    return lib().oracle_count_unique(idx, idx.size)


def concat_forward(inputs):
    """Concat axis=1 (src/ops/concat.cu:211-251): [B][w_i] inputs -> [B][sum w_i]."""
    inputs = [np.ascontiguousarray(x, np.float32) for x in inputs]
    B = inputs[0].shape[0]
    widths = np.array([x.shape[1] for x in inputs], np.int64)
    out = np.empty((B, int(widths.sum())), np.float32)
    ptrs = (C.c_void_p * len(inputs))(*[x.ctypes.data for x in inputs])
    lib().oracle_concat_forward(out, ptrs, widths, len(inputs), B)
    return out


def concat_backward(out_grad, widths):
    out_grad = np.ascontiguousarray(out_grad, np.float32)
    B = out_grad.shape[0]
    widths = np.array(widths, np.int64)
    grads = [np.zeros((B, int(w)), np.float32) for w in widths]
    ptrs = (C.c_void_p * len(grads))(*[g.ctypes.data for g in grads])
    lib().oracle_concat_backward(out_grad, ptrs, widths, len(grads), B)
    return grads


# ------------------------------------------------------------------------------------------
# The reference's own code (oracle/_ref). CPU: numpy arrays. GPU: raw device pointers (ints).
# ------------------------------------------------------------------------------------------
_ref_cpu = {}
_ref_gpu = None
_vp = C.c_void_p


def ref_cpu_available(ndebug=False):
    return os.path.exists(os.path.join(HERE, "_ref",
                                       "libffref_cpu_ndebug.so" if ndebug else "libffref_cpu.so"))


def ref_cpu(ndebug=False):
    """libffref_cpu.so (asserts on: idx < 1,000,000, bag 1) or the -DNDEBUG build."""
    if ndebug not in _ref_cpu:
        name = "libffref_cpu_ndebug.so" if ndebug else "libffref_cpu.so"
        L = C.CDLL(os.path.join(HERE, "_ref", name))
        L.ffref_cpu_forward_task.argtypes = [_i64p, _f32p, _f32p, _i64, _i64, _i64, _i64]
        L.ffref_cpu_forward_task.restype = None
        L.ffref_cpu_backward_task.argtypes = [_i64p, _f32p, _f32p, _i64, _i64, _i64, _i64]
        L.ffref_cpu_backward_task.restype = None
        L.ffref_cpu_lookup.argtypes = [_i64p, _f32p, _f32p, _i64, _i64, _i64, _i64, _int]
        L.ffref_cpu_lookup.restype = None
        L.ffref_cpu_run_tables.argtypes = [_int, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp),
                                           C.POINTER(_vp), C.POINTER(_vp), _i64p, _i64, _i64,
                                           _i64, _int, _int, _int, _int]
        L.ffref_cpu_run_tables.restype = C.c_double
        _ref_cpu[ndebug] = L
    return _ref_cpu[ndebug]


def ref_strategy_load_op(path, name):
    """The reference's load_strategies_from_file + lookup of op `name` (src/runtime/strategy.cc:
    95-146) -> (device_type, dims, device_ids)."""
    L = ref_cpu()
    out = (C.c_int * 1100)()
    n = L.ffref_strategy_load_op(path.encode(), name.encode(), out, 1100)
    if n < 0:
        raise KeyError(f"{name}: reference loader returned {n}")
    nd = out[1]
    return out[0], list(out[2:2 + nd]), list(out[2 + nd:n])


def ref_strategy_save_tablewise(path, num_tables, G):
    """The DLRM generator's table-wise strategy written by the reference's save_strategies_to_file."""
    rc = ref_cpu().ffref_strategy_save_tablewise(path.encode(), num_tables, G)
    assert rc == 0


def ref_cpu_forward_task(idx, W, ndebug=False):
    """Embedding::forward_task_cpu (src/ops/embedding.cc:377-415); bag size must be 1."""
    W = np.ascontiguousarray(W, np.float32)
    idx, B, L, R, D = _shape(idx, W)
    out = np.full((B, D), np.nan, np.float32)
    ref_cpu(ndebug).ffref_cpu_forward_task(idx, out, W, L, D, B, R)
    return out


def ref_cpu_backward_task(idx, outgrad, Wgrad, ndebug=False):
    """Embedding::backward_task_cpu (src/ops/embedding.cc:417-453); accumulates into Wgrad."""
    idx = np.ascontiguousarray(idx, np.int64)
    outgrad = np.ascontiguousarray(outgrad, np.float32)
    B, L = idx.shape
    R, D = Wgrad.shape
    ref_cpu(ndebug).ffref_cpu_backward_task(idx, outgrad, Wgrad, L, D, B, R)
    return Wgrad


def ref_cpu_lookup(idx, W, normalize=False, ndebug=True):
    """EmbeddingLookup_int64_t_float_float__avx2_fma (src/ops/embedding.cc:23-319) directly."""
    W = np.ascontiguousarray(W, np.float32)
    idx, B, L, R, D = _shape(idx, W)
    out = np.full((B, D), np.nan, np.float32)
    ref_cpu(ndebug).ffref_cpu_lookup(idx, out, W, L, D, B, R, int(normalize))
    return out


def ref_cpu_run_tables(idx_batches, outs, Ws, outgrads, Wgrads, P, steps, fwd_only=False,
                       ndebug=False):
    """Time the reference CPU tasks: table t on worker thread t % P (see ref_glue_cpu.cc).
    idx_batches[t]: int64 [nbatches][B][L]."""
    T = len(Ws)
    nb, B, L = idx_batches[0].shape
    D = Ws[0].shape[1]
    arr = lambda xs: (_vp * T)(*[x.ctypes.data for x in xs])
    R = np.array([w.shape[0] for w in Ws], np.int64)
    return ref_cpu(ndebug).ffref_cpu_run_tables(T, arr(idx_batches), arr(outs), arr(Ws),
                                                arr(outgrads), arr(Wgrads), R, L, D, B, P, steps,
                                                nb, int(fwd_only))


def ref_gpu_available():
    return os.path.exists(os.path.join(HERE, "_ref", "libffref_gpu.so"))


def _bind_gpu(L):
    if True:
        L.ffref_gpu_forward_task.argtypes = [_vp, _vp, _vp, _i64, _i64, _i64, _i64, _int]
        L.ffref_gpu_forward_task.restype = None
        L.ffref_gpu_backward_task.argtypes = [_vp, _vp, _vp, _i64, _i64, _i64, _i64, _int]
        L.ffref_gpu_backward_task.restype = None
        L.ffref_gpu_forward_kernel.argtypes = [_vp, _vp, _vp, _int, _int, _int, _int, _vp]
        L.ffref_gpu_forward_kernel.restype = None
        L.ffref_gpu_backward_kernel.argtypes = [_vp, _vp, _vp, _int, _int, _int, _int, _vp]
        L.ffref_gpu_backward_kernel.restype = None
        L.ffref_gpu_zero_task.argtypes = [_vp, _i64, _i64]
        L.ffref_gpu_zero_task.restype = None
        L.ffref_gpu_sgd_task.argtypes = [_vp, _vp, _i64, _i64, C.c_double, C.c_double]
        L.ffref_gpu_sgd_task.restype = None
        L.ffref_gpu_set_workspace.argtypes = [_vp, C.c_size_t]
        L.ffref_gpu_set_workspace.restype = None
        if hasattr(L, "ffref_gpu_sgd_task_full"):
            L.ffref_gpu_sgd_task_full.argtypes = [_vp, _vp, _vp, _i64, _i64, C.c_double, C.c_double,
                                                  C.c_double, _int]
            L.ffref_gpu_sgd_task_full.restype = None
        if hasattr(L, "ffref_gpu_adam_task"):
            L.ffref_gpu_adam_task.argtypes = [_vp, _vp, _vp, _vp, _i64, _i64, C.c_double, C.c_double,
                                              C.c_double, C.c_double, C.c_double]
            L.ffref_gpu_adam_task.restype = None
        if hasattr(L, "ffref_gpu_load_sparse_input"):  # not in the drop-in library
            L.ffref_gpu_load_sparse_input.argtypes = [_vp, _i64, _int, _int, _int, _int, _int, _vp]
    return L


def ref_gpu():
    """libffref_gpu.so: all arguments are device pointers passed as ints (tensor.data_ptr())."""
    global _ref_gpu
    if _ref_gpu is None:
        _ref_gpu = _bind_gpu(C.CDLL(os.path.join(HERE, "_ref", "libffref_gpu.so")))
    return _ref_gpu


_dropin_gpu = None


def dropin_gpu_available():
    return os.path.exists(os.path.join(HERE, "_ref", "libffdropin_gpu.so"))


def dropin_gpu():
    """libffdropin_gpu.so: the SAME task frames and the reference's own accessor objects, with
    the reference's embedding.cu swapped for dlrm-flexflow_b200/csrc/flexflow/embedding_ffemb.cu
    (which calls libffemb.so).  Same entry points as ref_gpu() + the fused leaf task."""
    global _dropin_gpu
    if _dropin_gpu is None:
        L = _bind_gpu(C.CDLL(os.path.join(HERE, "_ref", "libffdropin_gpu.so")))
        L.ffdropin_gpu_backward_sgd_task.argtypes = [_vp, _vp, _vp, _i64, _i64, _i64, _i64, _int, _f32]
        L.ffdropin_gpu_backward_sgd_task.restype = None
        pp = C.POINTER(C.c_void_p)
        L.ffdropin_gpu_group_task.argtypes = [_int, _int, pp, pp, pp, C.POINTER(C.c_int64), _i64, _i64,
                                              C.POINTER(C.c_int64), _int, _f32, _int]
        L.ffdropin_gpu_group_task.restype = None
        _dropin_gpu = L
    return _dropin_gpu


def launch_record(which, what, idx_ptrs, out_ptrs, W_ptrs, bags, D, B, rows, Wgrad_ptrs=None, aggr=AGGR_SUM,
                  lr=0.01, prepared=False):
    """The launcher side run against a Legion stand-in that writes the index launch down (ref_glue_gpu.cu,
    ffref_launch_record).  which = "fused": the per-GPU group launchers of
    csrc/flexflow/embedding_group_ffemb.cu (what 0 forward / 1 prepare / 2 backward+SGD, any number of
    tables); which = "reference": the reference's own Embedding::forward / backward from its object code
    (what 0 / 2, one table).  Pointers are only carried: no device is touched.
    -> dict(task_id, tag, mapper_id, must, points, args=..., local_arg_is_lead_meta, consts=...,
            reqs=[dict(privilege, coherence, tag, fields, base, extent)])."""
    n = len(bags)
    L = ref_gpu() if which == "reference" else dropin_gpu()
    pp = C.POINTER(C.c_void_p)
    L.ffref_launch_record.argtypes = [_int, _int, pp, pp, pp, pp, C.POINTER(C.c_int64), _i64, _i64,
                                      C.POINTER(C.c_int64), _int, _f32, _int, C.POINTER(C.c_int64), _int]
    L.ffref_launch_record.restype = _int
    arr = lambda xs: (C.c_void_p * n)(*[int(x) for x in xs]) if xs is not None else None
    i64 = lambda xs: (C.c_int64 * n)(*[int(x) for x in xs])
    H = 32
    rec = (C.c_int64 * (H + 8 * 3 * n))()
    k = L.ffref_launch_record(int(what), n, arr(idx_ptrs), arr(out_ptrs), arr(W_ptrs), arr(Wgrad_ptrs), i64(bags),
                              int(D), int(B), i64(rows), int(aggr), float(lr), int(prepared), rec, len(rec))
    if k < 0:
        raise RuntimeError(f"ffref_launch_record: {k}")
    r = list(rec[:k])
    lr_back = np.array([r[8]], np.int32).view(np.float32)[0]
    reqs = [dict(privilege=e[0], coherence=e[1], tag=e[2], fields=[e[4]] * e[3], base=e[5], extent=(e[6], e[7]))
            for e in (r[H + 8 * i:H + 8 + 8 * i] for i in range(r[1]))]
    return dict(task_id=r[0], tag=r[2], mapper_id=r[3], must=bool(r[4]), points=r[5],
                args=dict(len=r[6], num_tables=r[7], lr=float(lr_back), prepared=r[9], parity=r[10]),
                point_args=r[11], local_arg_is_lead_meta=bool(r[12]), local_arglen=r[13],
                consts=dict(READ_ONLY=r[14], WRITE_ONLY=r[15], READ_WRITE=r[16], EXCLUSIVE=r[17], FID_DATA=r[18],
                            MAP_TO_ZC_MEMORY=r[19], TASK_FORWARD=r[20], TASK_PREPARE=r[21], TASK_BACKWARD_SGD=r[22],
                            EMBED_FWD_TASK_ID=r[23], EMBED_BWD_TASK_ID=r[24]),
                reqs=reqs)


def create_weights_record(which, name="embedding3", rows=1000, dim=64):
    """Weight creation written down (ref_glue_gpu.cu, ffref_create_weights_record): which = "reference" runs
    the reference's own Embedding::create_weights (src/ops/embedding.cu:58-74, libffref_gpu.so), "fused" the
    drop-in's ffemb_fused_create_weights (libffdropin_gpu.so).  No device is touched."""
    L = ref_gpu() if which == "reference" else dropin_gpu()
    L.ffref_create_weights_record.argtypes = [C.c_char_p, _int, _int, C.POINTER(C.c_int64), C.c_char_p]
    L.ffref_create_weights_record.restype = _int
    rec = (C.c_int64 * 16)()
    buf = C.create_string_buffer(64)
    assert L.ffref_create_weights_record(name.encode(), int(rows), int(dim), rec, buf) == 16
    r = list(rec)
    return dict(task_is_calls=r[0], task_is_dims=r[1], weight_calls=r[2], dims=(r[3], r[4]), data_type=r[5],
                create_grad=bool(r[6]), sync_type=r[7], op_is_this=bool(r[8]), initializer_is_ops=bool(r[9]),
                consts=dict(DT_FLOAT=r[10], NONE=r[11], PS=r[12], NCCL=r[13]), task_points=r[14],
                weight_num_dim=r[15], task_is_name=buf.value.decode())
