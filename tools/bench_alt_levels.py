"""Per-level timing of the tensor-core alt-correlation (indexed fp16 path) on one update_lowmem chunk."""
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
rnd = torch.rand_like(coords) * torch.tensor([W, H], device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def t(fn, n=10):
    for _ in range(3): fn()
    ts = []
    for _ in range(n):
        flush.zero_(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b) * 1e3)
    return sorted(ts)[n // 2]
for name, c0 in (("smooth", coords), ("random", rnd)):
    for l in range(4):
        c = (c0 / 2 ** l).contiguous()
        print(name, "level", l, "%.1f us" % t(lambda: B.altcorr_forward_indexed(pyr[0], pyr[l], ii, jj, c, 3)),
              "| fp32 drop-in %.1f us" % t(lambda: B.altcorr_forward(pyr[0][ii].float(), pyr[l][jj].float(), c, 3)) if l == 0 else "")
    print(name, "fused 4 levels %.1f us" % t(lambda: B.altcorr_pyramid_indexed(pyr, ii, jj, c0, 3)))
