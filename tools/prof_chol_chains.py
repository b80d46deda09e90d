"""Timing of droid_chol_solve_chains_f64 on a BA-like structure: n = 5994, band 348, k chains + (k-1) separator groups
of 192 rows (group m reaches into the tail of chain m and the head -- the tail for the last chain -- of chain m+1)."""
import ctypes, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "droid-slam_b200"))
import droid_backends as B
from droid_backends import _lib, _p, _stream

n, bw, sg = 5994, 348, 192
NONE = np.iinfo(np.int32).max
rng = np.random.default_rng(0)
for k in (0, 1, 2, 3, 4, 6, 8):
    nsep = sg * max(k - 1, 0) if k > 1 else (sg if k == 1 else 0)
    body = n - nsep
    ends = [int(round(body * (c + 1) / k / 2)) * 2 for c in range(k)] if k else []
    b0 = [0] + ends + [n]
    firstc = np.full((k + 1, n), NONE, np.int64)
    A = torch.zeros((n, n), dtype=torch.float64, device="cuda")
    r = np.arange(n)
    for c in range(k + 1):
        lo, hi = b0[c], b0[c + 1]
        if c < k or k == 0:
            firstc[c, lo:hi] = np.maximum(lo, r[lo:hi] - bw)
        else:
            firstc[c, lo:hi] = lo
    if k >= 1:
        ngroups = max(k - 1, 1)
        for m in range(ngroups):
            rows = slice(b0[k] + m * sg, b0[k] + (m + 1) * sg)
            firstc[m, rows] = b0[m + 1] - bw                      # tail of chain m
            if m + 1 < k:
                firstc[m + 1, rows] = b0[m + 2] - bw if m + 1 == k - 1 else b0[m + 1]
    fc = torch.from_numpy(firstc.astype(np.int32)).cuda()
    # values: fill the envelope with small noise, strong diagonal
    M = torch.zeros((n, n), dtype=torch.float64, device="cuda")
    cols = torch.arange(n, device="cuda")[None, :]
    rowsT = torch.arange(n, device="cuda")[:, None]
    mask = torch.zeros((n, n), dtype=torch.bool, device="cuda")
    for c in range(k + 1):
        f = fc[c].long()[:, None]
        ok = (f <= rowsT) & (f >= b0[c]) & (f < b0[c + 1])
        mask |= ok & (cols >= f) & (cols < min(b0[c + 1], n)) & (cols < rowsT)
    M = torch.randn((n, n), dtype=torch.float64, device="cuda") * 0.01 * mask
    M = M + M.T + torch.eye(n, dtype=torch.float64, device="cuda") * 20.0
    bvec = torch.randn(n, dtype=torch.float64, device="cuda")
    want = torch.linalg.solve(M, bvec)
    buf = torch.empty((n * n + n,), dtype=torch.float64, device="cuda")
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    wsb = int(_lib.droid_chol_chains_workspace_bytes(n, k, n - (ends[-1] if k else 0)))
    ws = torch.empty(wsb, dtype=torch.uint8, device="cuda")
    cends = (ctypes.c_int * max(k, 1))(*ends) if k else None
    def run():
        buf[: n * n] = M.reshape(-1); buf[n * n:] = bvec
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = _lib.droid_chol_solve_chains_f64(_p(buf), ctypes.c_void_p(buf.data_ptr() + 8 * n * n), n, _p(fc), k, cends,
                                             _p(info), _p(ws), ctypes.c_size_t(wsb), _stream(M))
        e1.record(); torch.cuda.synchronize()
        assert rc == 0, rc
        return e0.elapsed_time(e1)
    for _ in range(3): run()
    ts = sorted(run() for _ in range(10))
    err = float((buf[n * n:] - want).abs().max() / want.abs().max())
    print("k=%d ends=%s sep=%d : %.3f ms (min %.3f)  rel err %.1e info %d" % (k, ends, n - (ends[-1] if k else 0), ts[5], ts[0], err, int(info.item())))
