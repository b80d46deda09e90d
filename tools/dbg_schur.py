"""Captures the reduced camera system (H - S | v - v_S) of the first BA iteration through the reduce hook and
compares the tensor-core Schur kernel (default) against the CUDA-core kernel (DROID_SCHUR_TC=0, second process)."""
import os, sys, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "droid-slam_b200"))
import numpy as np, torch

def capture(name, kw):
    import synth, droid_backends as B
    p = synth.make_ba_problem(name, **kw)
    K = ("poses","disps","intrinsics","disps_sens","targets","weights","eta","ii","jj")
    d = {k: torch.from_numpy(np.ascontiguousarray(p[k])).cuda() for k in K}
    got = []
    def _reduce(user, buf, n, stream):
        if not got:
            got.append(B._as_tensor(buf, n, d["poses"].device).clone())
        return 0
    B.ba(d["poses"], d["disps"], d["intrinsics"], d["disps_sens"], d["targets"], d["weights"], d["eta"], d["ii"], d["jj"],
         p["t0"], p["t1"], 1, 1e-4, 0.1, False, _shard=(0, 1 << 30), _reduce=_reduce)
    torch.cuda.synchronize()
    return got[0].cpu().numpy(), p["t1"] - p["t0"]

if __name__ == "__main__":
    name = sys.argv[1] if len(sys.argv) > 1 else "C1"
    kw = eval(sys.argv[2]) if len(sys.argv) > 2 else {}
    if os.environ.get("DBG_CHILD"):
        a, P = capture(name, kw)
        np.save(os.environ["DBG_CHILD"], a)
        sys.exit(0)
    a, P = capture(name, kw)
    env = dict(os.environ, DROID_SCHUR_TC="0", DBG_CHILD="/tmp/_dbg_ref.npy")
    subprocess.check_call([sys.executable, __file__] + sys.argv[1:], env=env)
    r = np.load("/tmp/_dbg_ref.npy")
    n6 = 6 * P
    A, b = a[: n6 * n6].reshape(n6, n6), a[n6 * n6:]
    Ar, br = r[: n6 * n6].reshape(n6, n6), r[n6 * n6:]
    print("n6", n6, "max|A|", np.abs(Ar).max(), "max|A-Ar|", np.abs(A - Ar).max(), "max|b-br|", np.abs(b - br).max(), "max|b|", np.abs(br).max())
    D = np.abs(A - Ar)
    blk = D.reshape(P, 6, P, 6).max(axis=(1, 3))
    np.set_printoptions(precision=2, linewidth=200, suppress=False)
    print("block max abs diff (rows = pose):\n", blk[:8, :8])
    i, j = np.unravel_index(np.argmax(D), D.shape)
    print("worst at", i, j, "got", A[i, j], "ref", Ar[i, j])
    bi, bj = i // 6, j // 6
    print("got block\n", A[6*bi:6*bi+6, 6*bj:6*bj+6], "\nref block\n", Ar[6*bi:6*bi+6, 6*bj:6*bj+6])
