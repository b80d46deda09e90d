"""frame_distance on 10^6 pairs: this repo (workspace path, plain path) vs the reference extension; bit-exactness."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("droid-slam_b200", "oracle", os.path.join("oracle", "_ref")): sys.path.insert(0, os.path.join(ROOT, p))
import droid_backends as B, synth
try:
    import droid_backends_ref as R
except Exception:
    R = None
pk = synth.make_ba_problem("C4", n_frames=1000, knn=2)
dev = "cuda"
P, D, K = (torch.from_numpy(pk[k]).to(dev) for k in ("gt_poses", "gt_disps", "intrinsics"))
a, b = torch.meshgrid(torch.arange(1000, device=dev), torch.arange(1000, device=dev), indexing="ij")
a, b = a.reshape(-1).contiguous(), b.reshape(-1).contiguous()
def t(fn, n=5):
    fn(); torch.cuda.synchronize(); ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]
print("ws path   %.3f ms" % t(lambda: B.frame_distance(P, D, K, a, b, 0.3)))
print("swapped   %.3f ms" % t(lambda: B.frame_distance(P, D, K, b, a, 0.3)))
print("plain     %.3f ms" % t(lambda: B.frame_distance(P, D, K, a, b, 0.3, _plain=True)))
if R is not None:
    print("reference %.3f ms" % t(lambda: R.frame_distance(P, D, K, a, b, 0.3), 3))
    for x, y in ((a, b), (b, a)):
        r = R.frame_distance(P, D, K, x, y, 0.3)
        print("bit-exact ws", bool(torch.equal(B.frame_distance(P, D, K, x, y, 0.3), r)), "plain", bool(torch.equal(B.frame_distance(P, D, K, x, y, 0.3, _plain=True), r)),
              "sentinels", int((r == 1000.0).sum()))
