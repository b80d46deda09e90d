"""Per-phase timing of the headline BA (C4, 1000 keyframes) + parity vs the reference extension (one call of 2 iterations)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("droid-slam_b200", "oracle", os.path.join("oracle", "_ref")): sys.path.insert(0, os.path.join(ROOT, p))
import droid_backends as B, synth
p = synth.make_ba_problem("C4", n_frames=int(os.environ.get("FRAMES", "1000")), knn=16)
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
d = {k: cu(p[k]) for k in ("poses", "disps", "intrinsics", "disps_sens", "targets", "weights", "eta", "ii", "jj")}
p0, d0 = d["poses"].clone(), d["disps"].clone()
def step(mod=B):
    d["poses"].copy_(p0); d["disps"].copy_(d0)
    return mod.ba(d["poses"], d["disps"], d["intrinsics"], d["disps_sens"], d["targets"], d["weights"], d["eta"], d["ii"], d["jj"],
                  p["t0"], p["t1"], 2, 1e-5, 1e-2, False)
for _ in range(3): step()
torch.cuda.synchronize()
B.prof_enable(True); B.prof_read()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): step()
e1.record(); torch.cuda.synchronize()
ph = B.prof_read(); B.prof_enable(False)
ms = e0.elapsed_time(e1) / 10
print("ms/step %.3f  iters/s %.1f  status %d" % (ms, 2000 / ms, B.ba_status()))
print({k: round(v[0] / max(v[1], 1), 4) for k, v in ph.items() if v[1] > 0})
if os.environ.get("PARITY", "1") == "1":
    import droid_backends_ref as R
    dx, dz = step(); mp, md = d["poses"].clone(), d["disps"].clone()
    rx, rz = step(R); torch.cuda.synchronize()
    rel = lambda a, b: float((a.double() - b.double()).abs().max() / b.double().abs().max())
    print("vs reference ext: dx %.2e dz %.2e poses %.2e disps %.2e" % (rel(dx, rx), rel(dz.reshape(-1), rz.reshape(-1)), rel(mp, d["poses"]), rel(md, d["disps"])))
