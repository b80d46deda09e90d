"""Differential debug: permuted edge order / chains, current library vs another build (DROID_B200_LIB)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "droid-slam_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import droid_backends as B, synth
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
def run(p, iters=1):
    po, di = cu(p["poses"]), cu(p["disps"])
    dx, dz = B.ba(po, di, cu(p["intrinsics"]), cu(p["disps_sens"]), cu(p["targets"]), cu(p["weights"]), cu(p["eta"]),
                  cu(p["ii"]), cu(p["jj"]), p["t0"], p["t1"], iters, 1e-5, 1e-2, False)
    torch.cuda.synchronize()
    return dx.cpu().numpy(), dz.cpu().numpy(), B.ba_status()
rel = lambda a, b: float(np.abs(a - b).max() / np.abs(b).max())
p = synth.make_ba_problem("C4", n_frames=60)
a = run(p)
q = dict(p); perm = np.roll(np.arange(len(p["ii"])), 7)
for k in ("ii", "jj", "targets", "weights"): q[k] = np.ascontiguousarray(p[k][perm])
c = run(q)
print(B.__version__, "status", a[2], c[2], "dx rel", rel(c[0], a[0]), "dz rel", rel(c[1], a[1]))
bad = [m for m in range(a[1].shape[0]) if rel(c[1][m], a[1][m]) > 1e-3]
print("frames with different dz:", bad, "edges of frame 59 in q:", np.nonzero(q["ii"] == 59)[0])
for rollby in (1, 3, 64, 500):
    q = dict(p); perm = np.roll(np.arange(len(p["ii"])), rollby)
    for k in ("ii", "jj", "targets", "weights"): q[k] = np.ascontiguousarray(p[k][perm])
    c = run(q)
    print("roll", rollby, "dx rel", rel(c[0], a[0]), "dz rel", rel(c[1], a[1]), [m for m in range(60) if rel(c[1][m], a[1][m]) > 1e-3])
