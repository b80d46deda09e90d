"""Per-call fixed cost of ba() at the headline size (graph bookkeeping, pose order, host read-back):
fixed = 2 T(iterations=1) - T(iterations=2)."""
import sys
sys.path.insert(0, "/root/repo/droid-slam_b200")
import numpy as np, torch
import droid_backends as B, synth
dev = "cuda"
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
p = synth.make_ba_problem("C4", n_frames=1000, knn=16)
d = {k: cu(p[k]) for k in ("poses", "disps", "intrinsics", "disps_sens", "targets", "weights", "eta", "ii", "jj")}
def run(it, n=12):
    ts = []
    for _ in range(n):
        po, di = d["poses"].clone(), d["disps"].clone()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        B.ba(po, di, d["intrinsics"], d["disps_sens"], d["targets"], d["weights"], d["eta"], d["ii"], d["jj"], p["t0"], p["t1"], it, 1e-5, 1e-2, False)
        b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    ts.sort(); return ts[len(ts) // 2]
t1, t2, t4 = run(1), run(2), run(4)
print("T1 %.3f ms  T2 %.3f ms  T4 %.3f ms  per-iteration %.3f  fixed per call %.3f ms" % (t1, t2, t4, (t4 - t2) / 2, t2 - 2 * (t4 - t2) / 2))
