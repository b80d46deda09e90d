"""One global-BA call at the bench workload (C4, 1000 keyframes, knn 16, 2 iterations) for ncu launch lists."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "droid-slam_b200"))
import numpy as np, torch, synth, droid_backends as B
p = synth.make_ba_problem("C4", n_frames=1000, knn=16)
K = ("poses","disps","intrinsics","disps_sens","targets","weights","eta","ii","jj")
d = {k: torch.from_numpy(np.ascontiguousarray(p[k])).cuda() for k in K}
B.ba(d["poses"], d["disps"], d["intrinsics"], d["disps_sens"], d["targets"], d["weights"], d["eta"], d["ii"], d["jj"], p["t0"], p["t1"], 2, 1e-5, 1e-2, False)
torch.cuda.synchronize()
print("ok", B.ba_status())
