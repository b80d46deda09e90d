import sys, json
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/droid-slam_b200")
import torch, bench
import droid_backends as B
dev = torch.device("cuda:0")
def timeit(fn, n=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b) * 1e-3)
    ts.sort(); return ts[len(ts) // 2]
out = bench.altcorr_and_geometry_metrics(dev, 6539.5, timeit)
print(json.dumps({k: v for k, v in out.items() if "altcorr" in k}, indent=1))
