"""One frontend-window BA call (C3: 12 keyframes, 60 edges, 48x64, iterations=2) for an ncu launch list."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "droid-slam_b200"))
import numpy as np, torch, synth, droid_backends as B
p = synth.make_ba_problem("C3")
K = ("poses","disps","intrinsics","disps_sens","targets","weights","eta","ii","jj")
d = {k: torch.from_numpy(np.ascontiguousarray(p[k])).cuda() for k in K}
for _ in range(3):
    B.ba(d["poses"].clone(), d["disps"].clone(), d["intrinsics"], d["disps_sens"], d["targets"], d["weights"], d["eta"], d["ii"], d["jj"], p["t0"], p["t1"], 2, 1e-4, 0.1, False)
torch.cuda.synchronize()
print("ok")
