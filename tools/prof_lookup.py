"""corr_index_pyramid at BASELINE configs[1] (24 edges, 48x64, 4 levels, radius 3, fp16) for ncu."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "droid-slam_b200"))
import numpy as np, torch
import droid_backends as B, synth
Bn, H, W = 24, 48, 64
q = synth.make_corr_problem(Bn, H, W, 128, seed=1234)
f1 = torch.from_numpy(q["fmaps"][q["ii"]]).cuda(); f2 = torch.from_numpy(q["fmaps"][q["jj"]]).cuda()
coords = torch.from_numpy(q["coords"]).cuda()
vols = B.corr_build(f1, f2)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    flush.zero_()
    out = B.corr_index_pyramid(vols, coords, 3)
torch.cuda.synchronize()
print("ok", out.shape)
