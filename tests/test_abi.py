"""CPU tests of the drop-in boundary: libffemb.so loads, exports every symbol include/ffemb.h
declares, and rejects bad arguments before touching a device (no compute without a GPU)."""
import ctypes as C
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_built_and_exports_every_declared_symbol(ffemb):
    assert os.path.exists(ffemb.library_path()), "run __graft_entry__.build() first"
    declared = ffemb.declared_symbols()
    assert "ffemb_forward" in declared and "ffemb_backward_sgd_batched" in declared
    exported = set(ffemb.exported_symbols())
    missing = [s for s in declared if s not in exported]
    assert not missing, f"declared in ffemb.h but not exported: {missing}"
    # nothing but the C ABI leaks out of the library
    assert all(s.startswith("ffemb_") for s in exported), sorted(exported)
    L = ffemb.lib()
    assert L.ffemb_version() == 100


def test_argument_validation_returns_codes_without_a_device(ffemb):
    L = ffemb.lib()
    # AGGR_MODE_NONE (20) hits assert(aggr == AGGR_MODE_AVG) in the reference kernels
    assert L.ffemb_forward(8, 8, 8, 1, 16, 4, ffemb.AGGR_MODE_NONE, None) == 1
    assert b"aggr" in L.ffemb_last_error()
    assert L.ffemb_forward(8, 8, 8, 1, 0, 4, ffemb.AGGR_MODE_SUM, None) == 1      # D = 0
    assert L.ffemb_forward(8, 8, 8, 1, 16, -1, ffemb.AGGR_MODE_SUM, None) == 1    # B < 0
    assert L.ffemb_forward(8, None, 8, 1, 16, 4, ffemb.AGGR_MODE_SUM, None) == 1  # NULL out
    assert L.ffemb_forward(8, 8, 8, -1, 16, 4, ffemb.AGGR_MODE_SUM, None) == 1    # L < 0
    assert L.ffemb_forward_batched(None, 1, 16, 4, ffemb.AGGR_MODE_SUM, None) == 1
    # empty work is a no-op, not an error (and launches nothing)
    assert L.ffemb_forward(None, None, None, 1, 16, 0, ffemb.AGGR_MODE_SUM, None) == 0


def test_exchange_entry_points_validate_before_touching_a_device(ffemb):
    import ctypes as C
    L = ffemb.lib()
    from importlib import import_module
    b = import_module("dlrm-flexflow_b200.binding")
    assert L.ffemb_sync_pad_bytes() >= 16 * 8 * 8 and L.ffemb_sync_pad_bytes() % 8 == 0
    sy = b.ffemb_sync_t()
    sy.num_peers, sy.rank = 2, 0
    assert L.ffemb_sync_signal(C.byref(sy), 0, None) == 1 and b"peer_pad" in L.ffemb_last_error()  # NULL pads
    sy.peer_pad[0], sy.peer_pad[1] = 64, 128
    assert L.ffemb_sync_wait(C.byref(sy), 16, None) == 1      # slot out of range
    sy.rank = 2
    assert L.ffemb_sync_signal(C.byref(sy), 0, None) == 1     # rank outside the peer set
    assert L.ffemb_sync_signal(None, 0, None) == 1
    parts = (b.ffemb_table_t * 1)()
    parts[0].out, parts[0].L, parts[0].reserved = 256, 64, 100
    assert L.ffemb_exchange_backward_parts(parts, 1, 256, 8, 0, 128, None) == 1   # columns 100..163 of 128
    parts[0].L = 0
    assert L.ffemb_exchange_backward_parts(parts, 1, 256, 8, 0, 128, None) == 1   # zero width
    assert L.ffemb_exchange_backward_parts(parts, 0, None, 0, 0, 128, None) == 0  # nothing to do


def test_workspace_query(ffemb):
    L = ffemb.lib()
    assert L.ffemb_backward_workspace(0, 10, 16) == 0
    assert L.ffemb_backward_workspace(10, 10, 0) == 0
    small = L.ffemb_backward_workspace(1000, 1000, 16)
    big = L.ffemb_backward_workspace(100000, 1000, 16)
    assert 0 < small < big
    # keys widen to 64 bits once the global row id no longer fits 32
    assert L.ffemb_backward_workspace(100000, 1 << 33, 16) > big
    # missing workspace is reported, not dereferenced
    rc = L.ffemb_backward_sgd(16, 16, 16, 1, 16, 4, 100, ffemb.AGGR_MODE_SUM, 0.01, None, 0, None)
    assert rc == 3 and b"workspace" in L.ffemb_last_error()


def test_compute_fails_loudly_without_library(ffemb, monkeypatch):
    from importlib import import_module
    b = import_module("dlrm-flexflow_b200.binding")
    monkeypatch.setattr(b, "_lib", None)
    monkeypatch.setattr(b, "library_path", lambda: "/nonexistent/libffemb.so")
    with pytest.raises(ffemb.FFEmbError):
        b.lib()


def test_dropin_binding_defines_the_reference_symbols():
    """csrc/flexflow/embedding_ffemb.cu, compiled against the reference's own include/model.h,
    defines exactly the members of class Embedding it replaces (include/model.h:1154-1186)."""
    so = os.path.join(ROOT, "oracle", "_ref", "libffdropin_gpu.so")
    if not os.path.exists(so):
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    out = subprocess.run(["nm", "-DC", "--defined-only", so], stdout=subprocess.PIPE, text=True, check=True).stdout
    for sym in ("Embedding::forward_kernel(", "Embedding::backward_kernel(", "Embedding::forward_task(",
                "Embedding::backward_task(", "ffemb_embedding_backward_sgd_task("):
        assert sym in out, sym
    und = subprocess.run(["nm", "-D", "--undefined-only", so], stdout=subprocess.PIPE, text=True, check=True).stdout
    assert "ffemb_forward" in und and "ffemb_backward_sgd" in und  # the numeric work is libffemb's
    assert "embed_forward" not in out and "embed_backward" not in out  # none of the reference kernels


def test_header_is_plain_c_and_struct_layouts_match_the_ctypes_mirror(ffemb, tmp_path):
    """include/ffemb.h must compile as C99 (the boundary is a C ABI: no C++ types in it) and as C++11
    (the reference's dialect, CMakeLists.txt), and the ctypes mirror in binding.py must lay the
    descriptor structs out exactly as the C compiler does."""
    import ctypes as C
    import subprocess
    B = ffemb.binding
    structs = ["ffemb_table_t", "ffemb_sgd_t", "ffemb_adam_t", "ffemb_exchange_t", "ffemb_sync_t", "ffemb_route_t"]
    src = tmp_path / "layout.c"
    body = "".join(f'  printf("{s} %zu\\n", sizeof({s}));\n' for s in structs)
    body += '  printf("R %zu\\n", offsetof(ffemb_table_t, R));\n  printf("L %zu\\n", offsetof(ffemb_table_t, L));\n'
    body += '  printf("num_peers %zu\\n", offsetof(ffemb_exchange_t, num_peers));\n'
    body += '  printf("row_stride %zu\\n", offsetof(ffemb_exchange_t, row_stride));\n'
    body += '  printf("samples_per_peer %zu\\n", offsetof(ffemb_route_t, samples_per_peer));\n'
    body += '  printf("cap %llu\\n", (unsigned long long)FFEMB_STATIC_SCRATCH_BYTES);\n'
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "ffemb.h"\nint main(void) {\n' + body + "  return 0;\n}\n")
    inc = os.path.join(ROOT, "include")
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", f"-I{inc}", str(src), "-o", str(exe)],
                   check=True)
    subprocess.run(["g++", "-std=c++11", "-Wall", "-Wextra", "-Werror", f"-I{inc}", "-x", "c++", "-fsyntax-only", str(src)],
                   check=True)
    out = dict(l.split() for l in subprocess.run([str(exe)], stdout=subprocess.PIPE, text=True, check=True).stdout.splitlines())
    for s in structs:
        assert int(out[s]) == C.sizeof(getattr(B, s)), s
    assert int(out["R"]) == B.ffemb_table_t.R.offset and int(out["L"]) == B.ffemb_table_t.L.offset
    assert int(out["num_peers"]) == B.ffemb_exchange_t.num_peers.offset
    assert int(out["row_stride"]) == B.ffemb_exchange_t.row_stride.offset
    assert int(out["samples_per_peer"]) == B.ffemb_route_t.samples_per_peer.offset
    assert int(out["cap"]) == int(B.lib().ffemb_static_scratch_bytes())


def test_rowwise_route_and_padded_backward_validate_before_touching_a_device(ffemb):
    import ctypes as C
    L, b = ffemb.lib(), ffemb.binding
    tabs = (b.ffemb_table_t * 1)()
    tabs[0].idx, tabs[0].R, tabs[0].L, tabs[0].reserved = 64, 1000, 2, 0
    rt = b.ffemb_route_t()
    rt.num_peers, rt.rank, rt.samples_per_peer = 2, 0, 8
    # NULL peer buffers
    assert L.ffemb_rowwise_route(tabs, 1, 16, 8, 256, 16, C.byref(rt), None) == 1 and b"peer_slots" in L.ffemb_last_error()
    for h in range(2):
        rt.peer_slots[h], rt.peer_rows[h] = 1024, 4096
    assert L.ffemb_rowwise_route(tabs, 1, 16, 9, 256, 16, C.byref(rt), None) == 1      # more samples than a peer's block
    assert L.ffemb_rowwise_route(tabs, 1, 18, 8, 256, 20, C.byref(rt), None) == 1      # D not a multiple of 4
    assert L.ffemb_rowwise_route(tabs, 1, 16, 8, 260, 16, C.byref(rt), None) == 1      # dx misaligned
    assert L.ffemb_rowwise_route(tabs, 1, 16, 8, 256, 8, C.byref(rt), None) == 1       # row_stride < D
    rt.rank = 2
    assert L.ffemb_rowwise_route(tabs, 1, 16, 8, 256, 16, C.byref(rt), None) == 1      # rank outside the peer set
    rt.rank = 1
    tabs[0].reserved = 4
    assert L.ffemb_rowwise_route(tabs, 1, 16, 8, 256, 16, C.byref(rt), None) == 1      # block past the row
    tabs[0].reserved, tabs[0].R = 0, 0
    assert L.ffemb_rowwise_route(tabs, 1, 16, 8, 256, 16, C.byref(rt), None) == 1      # rows of the whole table unknown
    tabs[0].R = 1000
    assert L.ffemb_rowwise_route(tabs, 1, 16, 0, None, 16, C.byref(rt), None) == 0     # no samples: nothing launched
    assert L.ffemb_rowwise_route(tabs, 1, 16, 8, 256, 16, None, None) == 1
    # the owner's side: same workspace contract as every batched backward
    tabs[0].out, tabs[0].table = 256, 512
    assert L.ffemb_backward_sgd_padded(tabs, 1, 16, 8, ffemb.AGGR_MODE_SUM, 0.01, None, 0, None) == 3
    assert b"workspace" in L.ffemb_last_error()
    assert L.ffemb_backward_sgd_padded(tabs, 1, 16, 8, ffemb.AGGR_MODE_NONE, 0.01, None, 0, None) == 1
