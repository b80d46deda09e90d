"""Times droid_chol_solve_f64 alone at the size of the 1000-keyframe reduced camera system (n = 5994)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "droid-slam_b200"))
import torch
import droid_backends as B
n = int(sys.argv[1]) if len(sys.argv) > 1 else 5994
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
torch.manual_seed(0)
M = torch.randn(n, n // 4, device="cuda", dtype=torch.float64)
A = M @ M.T + n * torch.eye(n, device="cuda", dtype=torch.float64)
b = torch.randn(n, device="cuda", dtype=torch.float64)
x, info = B.chol_solve(A, b)
torch.cuda.synchronize()
print("info", int(info.item()), "resid", float((A @ x - b).abs().max() / b.abs().max()))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
buf = torch.empty(n * n + n, device="cuda", dtype=torch.float64)
import ctypes
inf = torch.zeros(1, dtype=torch.int32, device="cuda")
wsb = int(B._lib.droid_chol_workspace_bytes(n)); ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
ts = []
B.prof_enable(True); B.prof_read()
for _ in range(reps):
    buf[: n * n] = A.reshape(-1); buf[n * n:] = b
    torch.cuda.synchronize()
    e0.record()
    B._lib.droid_chol_solve_f64(ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(buf.data_ptr() + 8 * n * n), n,
                                ctypes.c_void_p(inf.data_ptr()), ctypes.c_void_p(ws.data_ptr()), ctypes.c_size_t(wsb),
                                ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
print({k: v for k, v in B.prof_read().items() if v[1]})
print("n", n, "ms", ts, "TFLOP/s", (n ** 3 / 3) / (min(ts) * 1e-3) / 1e12)
