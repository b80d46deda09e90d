"""corr_build (tcgen05) vs torch.matmul + 3x avg_pool2d at BASELINE configs[1]: 24 edges, 48x64x128 fp16."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "droid-slam_b200"))
import numpy as np, torch, torch.nn.functional as F
import droid_backends as B, synth
Bn, H, W = 24, 48, 64
p = synth.make_corr_problem(Bn, H, W, 128, seed=1)
f1 = torch.from_numpy(p["fmaps"][p["ii"]]).cuda(); f2 = torch.from_numpy(p["fmaps"][p["jj"]]).cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def ours(): return B.corr_build(f1, f2)
def ref():
    a = (f1 / 4.0).reshape(Bn, 128, H * W); b = (f2 / 4.0).reshape(Bn, 128, H * W)
    c = torch.matmul(a.transpose(1, 2), b).view(Bn * H * W, 1, H, W)
    out = [c]
    for _ in range(3):
        c = F.avg_pool2d(c, 2, stride=2); out.append(c)
    return out
def timeit(fn, n=10):
    for _ in range(3): fn()
    ts = []
    for _ in range(n):
        flush.zero_(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))
HW = H * W
flops = 2.0 * HW * HW * 128 * Bn
obytes = Bn * HW * (HW + HW // 4 + HW // 16 + HW // 64) * 2 + 2 * Bn * 128 * HW * 2
t1, t2 = timeit(ours), timeit(ref)
print("corr_build tcgen05: %.1f us  %.1f TFLOP/s  %.0f GB/s written+read (algorithmic %.1f MB)" % (t1 * 1e3, flops / t1 / 1e9, obytes / t1 / 1e6, obytes / 1e6))
print("torch matmul+3 pools: %.1f us  %.1f TFLOP/s  %.0f GB/s" % (t2 * 1e3, flops / t2 / 1e9, obytes / t2 / 1e6))
