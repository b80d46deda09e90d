"""Envelope Cholesky: correctness against numpy on a banded matrix with a few wide rows, and timing at n = 5994."""
import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "droid-slam_b200"))
import numpy as np, torch
import droid_backends as B


def banded(n, bw, seed=0, wide=()):
    rng = np.random.default_rng(seed)
    A = np.zeros((n, n))
    first = np.maximum(np.arange(n) - bw, 0)
    for r in wide:
        first[r] = rng.integers(0, max(1, r // 4))
    for r in range(n):
        A[r, first[r]:r] = rng.standard_normal(r - first[r]) * 0.1
    A = A + A.T + np.diag(np.full(n, 0.1 * 2 * bw + 5.0 + 0.1 * n * (len(wide) > 0)))
    return A, first.astype(np.int32)


for n, bw, wide in ((400, 100, ()), (1200, 200, (700, 1100)), (2500, 384, (2000,)), (2431, 130, ())):
    A, first = banded(n, bw, seed=n, wide=wide)
    b = np.random.default_rng(1).standard_normal(n)
    x, info = B.chol_solve(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda(), torch.from_numpy(first).cuda())
    ref = np.linalg.solve(A, b)
    print(n, bw, "info", int(info.item()), "rel err", float(np.abs(x.cpu().numpy() - ref).max() / np.abs(ref).max()))

n = 5994
A, first = banded(n, 384, seed=3)
At = torch.from_numpy(A).cuda(); bt = torch.randn(n, device="cuda", dtype=torch.float64); ft = torch.from_numpy(first).cuda()
buf = torch.empty(n * n + n, device="cuda", dtype=torch.float64)
inf = torch.zeros(1, dtype=torch.int32, device="cuda")
wsb = int(B._lib.droid_chol_workspace_bytes(n)); ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
B.prof_enable(True); B.prof_read()
for mode in ("env", "dense"):
    ts = []
    for _ in range(4):
        buf[: n * n] = At.reshape(-1); buf[n * n:] = bt
        torch.cuda.synchronize()
        e0.record()
        B._lib.droid_chol_solve_env_f64(ctypes.c_void_p(buf.data_ptr()), ctypes.c_void_p(buf.data_ptr() + 8 * n * n), n,
                                        ctypes.c_void_p(ft.data_ptr() if mode == "env" else 0), ctypes.c_void_p(inf.data_ptr()),
                                        ctypes.c_void_p(ws.data_ptr()), ctypes.c_size_t(wsb),
                                        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        e1.record(); torch.cuda.synchronize()
        ts.append(round(e0.elapsed_time(e1), 3))
    print(mode, "ms", ts, {k: (round(v[0], 3), v[1]) for k, v in B.prof_read().items() if v[1]})
    x = buf[n * n:].clone()
    print("  resid", float((At @ x - bt).abs().max()))
