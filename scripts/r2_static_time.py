"""Cost of the static-scratch entry point against ffemb_backward_sgd with a caller's workspace, and the
device's free memory across the calls (nothing may be allocated).  Output: gpurun_out/static_time.json"""
import importlib, json, os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("dlrm-flexflow_b200")
L_ = pkg.binding.lib()
torch.cuda.set_device(0)
out = []
for (R, D, L, B) in [(5000, 128, 3, 1000), (200_000, 64, 8, 8192), (1_000_000, 128, 4, 65536)]:
    rng = np.random.default_rng(1)
    idx = torch.from_numpy(rng.integers(0, R, (B, L), dtype=np.int64)).cuda()
    g = torch.randn(B, D, device="cuda")
    W = torch.randn(R, D, device="cuda")
    need = int(L_.ffemb_backward_workspace(B * L, R, D))
    ws = torch.empty(need + 256, dtype=torch.uint8, device="cuda")
    wp = (ws.data_ptr() + 255) & ~255
    s = torch.cuda.current_stream().cuda_stream
    def static():
        pkg.binding.check(L_.ffemb_backward_sgd_static(idx.data_ptr(), g.data_ptr(), W.data_ptr(), L, D, B, R, 21, 1e-3, s))
    def plain():
        pkg.binding.check(L_.ffemb_backward_sgd(idx.data_ptr(), g.data_ptr(), W.data_ptr(), L, D, B, R, 21, 1e-3, wp, need, s))
    res = {"R": R, "D": D, "L": L, "B": B, "scratch_bytes": need}
    for name, fn in (("static", static), ("workspace", plain)):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(100):
            fn()
        b.record()
        torch.cuda.synchronize()
        res[name + "_ms"] = a.elapsed_time(b) / 100
        res[name + "_free_bytes_delta"] = torch.cuda.mem_get_info()[0] - free0
    out.append(res)
    print(res)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/static_time.json", "w"), indent=1)
