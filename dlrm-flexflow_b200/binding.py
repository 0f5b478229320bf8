"""ctypes binding of include/ffemb.h — the same stub a reference maintainer would write for a
Python-side caller (INTEGRATION.md shows the C++ one used by the Legion task bodies)."""
import ctypes as C
import os
import re
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)

# include/ffconst.h:12-16
AGGR_MODE_NONE, AGGR_MODE_SUM, AGGR_MODE_AVG = 20, 21, 22
AGGR_MODE_MEAN = 122  # opt-in true mean (not a reference enum value)

FFEMB_OK, FFEMB_EINVAL, FFEMB_ECUDA, FFEMB_EWORKSPACE, FFEMB_ERANGE = 0, 1, 2, 3, 4
FFEMB_MAX_TABLES = 64
FFEMB_MAX_PEERS = 8


class FFEmbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"ffemb error {code}: {msg}")
        self.code = code


class ffemb_table_t(C.Structure):
    _fields_ = [("idx", C.c_void_p), ("out", C.c_void_p), ("table", C.c_void_p),
                ("R", C.c_int64), ("L", C.c_int32), ("reserved", C.c_int32)]


class ffemb_sgd_t(C.Structure):
    _fields_ = [("lr", C.c_float), ("weight_decay", C.c_float), ("momentum", C.c_float),
                ("nesterov", C.c_int32)]


class ffemb_adam_t(C.Structure):
    _fields_ = [("alpha_t", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float),
                ("weight_decay", C.c_float), ("epsilon", C.c_float)]


class ffemb_exchange_t(C.Structure):
    _fields_ = [("peer_x", C.c_void_p * FFEMB_MAX_PEERS), ("num_peers", C.c_int32),
                ("reserved", C.c_int32), ("batch_per_peer", C.c_int64), ("row_stride", C.c_int64)]


class ffemb_sync_t(C.Structure):
    _fields_ = [("peer_pad", C.c_void_p * FFEMB_MAX_PEERS), ("num_peers", C.c_int32), ("rank", C.c_int32)]


class ffemb_route_t(C.Structure):
    _fields_ = [("peer_slots", C.c_void_p * FFEMB_MAX_PEERS), ("peer_rows", C.c_void_p * FFEMB_MAX_PEERS),
                ("num_peers", C.c_int32), ("rank", C.c_int32), ("samples_per_peer", C.c_int64)]


def library_path():
    return os.path.join(HERE, "libffemb.so")


def build_library(verbose=False):
    """nvcc -gencode arch=compute_100a,code=sm_100a ... -> dlrm-flexflow_b200/libffemb.so"""
    r = subprocess.run(["make", "-s", "-j4", "-C", os.path.join(HERE, "csrc")],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or r.returncode:
        print(r.stdout)
    if r.returncode:
        raise RuntimeError("building libffemb.so failed")
    return library_path()


def declared_symbols():
    """Function names declared in include/ffemb.h."""
    text = open(os.path.join(ROOT, "include", "ffemb.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ffemb_[a-z0-9_]+)\s*\(", text)))


def exported_symbols():
    out = subprocess.run(["nm", "-D", "--defined-only", library_path()], stdout=subprocess.PIPE,
                         text=True, check=True).stdout
    return sorted(l.split()[-1] for l in out.splitlines() if " T " in l)


_lib = None


def lib():
    """Load libffemb.so. Fails loudly when it has not been built: there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        raise FFEmbError(-1, f"{path} is missing: run __graft_entry__.build() "
                             "(make -C dlrm-flexflow_b200/csrc); no CPU fallback exists")
    L = C.CDLL(path)
    vp, i64, i32, f32, sz = C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_size_t
    tp = C.POINTER(ffemb_table_t)
    L.ffemb_last_error.restype = C.c_char_p
    L.ffemb_version.restype = i32
    L.ffemb_forward.argtypes = [vp, vp, vp, i64, i64, i64, i32, vp]
    L.ffemb_forward_batched.argtypes = [tp, i32, i64, i64, i32, vp]
    L.ffemb_forward_exchange.argtypes = [tp, i32, i64, i64, i32, C.POINTER(ffemb_exchange_t), vp]
    L.ffemb_exchange_backward.argtypes = [tp, i32, i64, vp, i64, i64, i64, vp]
    L.ffemb_exchange_backward_parts.argtypes = [tp, i32, vp, i64, i64, i64, vp]
    L.ffemb_sync_pad_bytes.restype = sz
    L.ffemb_sync_signal.argtypes = [C.POINTER(ffemb_sync_t), i32, vp]
    L.ffemb_sync_wait.argtypes = [C.POINTER(ffemb_sync_t), i32, vp]
    L.ffemb_forward_rowwise.argtypes = [vp, vp, C.POINTER(vp), i32, i64, i64, i64, i64, i32, vp]
    L.ffemb_rowwise_localize.argtypes = [vp, vp, i64, i64, i32, vp, i64, vp]
    L.ffemb_rowwise_route.argtypes = [tp, i32, i64, i64, vp, i64, C.POINTER(ffemb_route_t), vp]
    L.ffemb_backward_sgd_padded.argtypes = [tp, i32, i64, i64, i32, f32, vp, sz, vp]
    L.ffemb_backward_dense_atomic.argtypes = [vp, vp, vp, i64, i64, i64, i32, vp]
    L.ffemb_backward_dense.argtypes = [vp, vp, vp, i64, i64, i64, i64, i32, vp, sz, vp]
    L.ffemb_backward_sgd.argtypes = [vp, vp, vp, i64, i64, i64, i64, i32, f32, vp, sz, vp]
    L.ffemb_backward_sgd_static.argtypes = [vp, vp, vp, i64, i64, i64, i64, i32, f32, vp]
    L.ffemb_static_scratch_bytes.restype = sz
    L.ffemb_backward_sgd_batched.argtypes = [tp, i32, i64, i64, i32, f32, vp, sz, vp]
    L.ffemb_backward_prepare.argtypes = [tp, i32, i64, i64, i32, vp, sz, vp]
    L.ffemb_backward_sgd_prepared.argtypes = [tp, i32, i64, i64, i32, f32, vp, sz, vp]
    L.ffemb_sgd_touched_words.argtypes = [i64]
    L.ffemb_sgd_touched_words.restype = i64
    L.ffemb_backward_sgdx_batched.argtypes = [tp, i32, i64, i64, i32, C.POINTER(ffemb_sgd_t), C.POINTER(vp),
                                              C.POINTER(vp), vp, sz, vp]
    L.ffemb_backward_adam_batched.argtypes = [tp, i32, i64, i64, i32, C.POINTER(ffemb_adam_t), C.POINTER(vp),
                                              C.POINTER(vp), C.POINTER(vp), vp, sz, vp]
    L.ffemb_forward_concat.argtypes = [tp, i32, i64, i64, i32, vp, i64, vp]
    L.ffemb_backward_sgd_concat.argtypes = [tp, i32, i64, i64, i32, f32, vp, i64, i32, vp, sz, vp]
    L.ffemb_backward_dense_batched.argtypes = [tp, i32, i64, i64, i32, vp, sz, vp]
    L.ffemb_backward_workspace.argtypes = [i64, i64, i64]
    L.ffemb_backward_workspace.restype = sz
    L.ffemb_backward_workspace_avg.argtypes = [i64, i64, i64, i64, i32]
    L.ffemb_backward_workspace_avg.restype = sz
    L.ffemb_backward_path.argtypes = [i64, i64, i64, i64]
    L.ffemb_load_sparse_inputs.argtypes = [tp, i32, vp, i64, vp, i64, i64, vp]
    L.ffemb_check_indices.argtypes = [vp, i64, i64, vp, vp]
    L.ffemb_profile_enable.argtypes = [i32]
    L.ffemb_profile_collect.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_int64)]
    L.ffemb_launch_count.restype = i64
    for name in declared_symbols():
        if not hasattr(L, name):
            raise FFEmbError(-1, f"libffemb.so does not export {name}")
    _lib = L
    return L


PHASES = ["forward", "keygen", "sort", "reduce1", "reduce_up", "atomic", "small", "exchange", "input", "dense",
          "sync", "prescale", "hot"]


def profile_enable(on=True):
    check(lib().ffemb_profile_enable(int(on)))


def profile_collect():
    """-> ({phase: ms}, {phase: launches}) accumulated since the last collect (synchronises)."""
    ms = (C.c_float * len(PHASES))()
    n = (C.c_int64 * len(PHASES))()
    check(lib().ffemb_profile_collect(ms, n))
    return dict(zip(PHASES, list(ms))), dict(zip(PHASES, list(n)))


def launch_count():
    return int(lib().ffemb_launch_count())


def check(rc):
    if rc != FFEMB_OK:
        raise FFEmbError(rc, lib().ffemb_last_error().decode())
