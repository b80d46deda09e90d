import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "droid-slam_b200"))
import numpy as np, torch
import droid_backends as B
n = 2560; bw = 384
rng = np.random.default_rng(0)
A = np.zeros((n, n)); first = np.maximum(np.arange(n) - bw, 0).astype(np.int32)
for r in range(n): A[r, first[r]:r] = rng.standard_normal(r - first[r]) * 0.1
A = A + A.T + np.diag(np.full(n, 0.2 * bw + 5.0))
b = rng.standard_normal(n)
x, info = B.chol_solve(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda(), torch.from_numpy(first).cuda())
torch.cuda.synchronize()
print("ok", int(info.item()))
