"""DROID_CHOL_TRACE=1 python tools/trace_chol.py : stage times of the solve inside one headline ba() iteration."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "droid-slam_b200"))
import droid_backends as B, synth
p = synth.make_ba_problem("C4", n_frames=int(os.environ.get("FRAMES", "1000")), knn=16)
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
d = {k: cu(p[k]) for k in ("poses", "disps", "intrinsics", "disps_sens", "targets", "weights", "eta", "ii", "jj")}
for it in range(3):
    print("---- call", it, file=sys.stderr)
    B.ba(d["poses"].clone(), d["disps"].clone(), d["intrinsics"], d["disps_sens"], d["targets"], d["weights"], d["eta"], d["ii"], d["jj"],
         p["t0"], p["t1"], 1, 1e-5, 1e-2, False)
torch.cuda.synchronize()
print("tree:", B.ba_tree())
