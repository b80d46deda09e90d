"""ncu target: the fused 4-level tensor-core alt-correlation on one update_lowmem chunk (smooth coordinates)."""
import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/droid-slam_b200")
import numpy as np, torch
import droid_backends as B, synth
dev = torch.device("cuda:0")
pr = synth.make_ba_problem("C4", n_frames=250, knn=16)
H, W, C = pr["ht"], pr["wd"], 128
sel = np.nonzero((pr["ii"] >= 120) & (pr["ii"] < 128))[0]
ii = torch.from_numpy(pr["ii"][sel]).to(dev); jj = torch.from_numpy(pr["jj"][sel]).to(dev)
g = torch.Generator(device="cpu").manual_seed(5)
f = (torch.randn((250, C, H, W), generator=g) * 0.5).to(dev)
pyr = []
for l in range(4):
    pyr.append(f.permute(0, 2, 3, 1).contiguous().half()); f = torch.nn.functional.avg_pool2d(f, 2, stride=2)
coords = torch.from_numpy(np.ascontiguousarray(pr["targets"][sel].transpose(0, 2, 3, 1))).to(dev)[:, None].contiguous()
mode = sys.argv[1] if len(sys.argv) > 1 else "fused"
for _ in range(3):
    if mode == "fused":
        B.altcorr_pyramid_indexed(pyr, ii, jj, coords, 3)
    else:
        f1 = pyr[0][ii].float().contiguous(); f2 = pyr[0][jj].float().contiguous()
        B.altcorr_forward(f1, f2, coords, 3)
torch.cuda.synchronize()
