import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "droid-slam_b200"))
import numpy as np, torch
import droid_backends as B, synth
Bn, H, W = 2, 48, 64
p = synth.make_corr_problem(Bn, H, W, 128, seed=1)
f1 = torch.from_numpy(p["fmaps"][p["ii"]]).cuda(); f2 = torch.from_numpy(p["fmaps"][p["jj"]]).cuda()
lv = B.corr_build(f1, f2); torch.cuda.synchronize()
a = (f1 / 4.0).reshape(Bn, 128, H * W); b = (f2 / 4.0).reshape(Bn, 128, H * W)
ref = torch.matmul(a.transpose(1, 2), b).float()
got = lv[0].reshape(Bn, H * W, H * W).float()
d = (got - ref).abs()
tol = ref.abs() * 2.0 ** -9 + 1e-4
bad = d > tol
print("bad frac", float(bad.float().mean()), "max", float(d.max()), "ref max", float(ref.abs().max()))
bm = bad[0].float()
print("by m tile (128):", [round(float(bm[i*128:(i+1)*128].mean()), 3) for i in range(H*W//128)])
print("by n row:", [round(float(bm[:, r*W:(r+1)*W].mean()), 3) for r in range(H)])
print("by n col (first row):", [round(float(bm[:, c].mean()), 2) for c in range(W)])
print("by m in tile0:", [round(float(bm[i].mean()), 2) for i in range(0, 128, 8)])
i = int(torch.nonzero(bad[0])[0][0]) if bad[0].any() else 0
print("first bad m", i, "row vals got", got[0, i, :8].tolist(), "ref", ref[0, i, :8].tolist())
# does got match some permutation? check got[0,m,n] vs ref with m or n shuffled inside 8-blocks
