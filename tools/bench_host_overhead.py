"""Host-side cost of the ctypes wrappers: wall time per call of small operators issued back to back (no sync inside the
loop), this module vs the reference's pybind extension."""
import sys, time
sys.path.insert(0, "/root/repo/droid-slam_b200"); sys.path.insert(0, "/root/repo/oracle/_ref")
import numpy as np, torch
import droid_backends as B, synth
try:
    import droid_backends_ref as R
except Exception:
    R = None
dev = "cuda"
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
p = synth.make_ba_problem("C3")
P, D, K = cu(p["gt_poses"]), cu(p["gt_disps"]), cu(p["intrinsics"])
ii, jj = cu(p["ii"][:2]), cu(p["jj"][:2])
vol = torch.randn((8, 48, 64, 48, 64), device=dev).half()
vols = [vol] + [torch.randn((8, 48, 64, 48 >> l, 64 >> l), device=dev).half() for l in (1, 2, 3)]
c = torch.rand((8, 2, 48, 64), device=dev) * 40
def t(fn, n=2000):
    for _ in range(50): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    dt = time.perf_counter() - t0; torch.cuda.synchronize()
    return dt / n * 1e6
for name, f, g in (("frame_distance(2 pairs)", lambda: B.frame_distance(P, D, K, ii, jj, 0.3), (lambda: R.frame_distance(P, D, K, ii, jj, 0.3)) if R else None),
                   ("corr_index_forward", lambda: B.corr_index_forward(vol, c, 3), (lambda: R.corr_index_forward(vol, c, 3)) if R else None),
                   ("corr_index_pyramid", lambda: B.corr_index_pyramid(vols, c, 3), None),
                   ("projmap(2 edges)", lambda: B.projmap(P, D, K, ii, jj), (lambda: R.projmap(P, D, K, ii, jj)) if R else None)):
    print("%-26s this module %6.1f us/call   reference ext %s" % (name, t(f), ("%6.1f us/call" % t(g)) if g else "-"))
