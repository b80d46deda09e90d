"""Write-only HBM bandwidth on this GPU (cudaMemset-style fill of 2 GiB), for the roofline of write-bound kernels."""
import torch
x = torch.empty(2 << 30, dtype=torch.uint8, device="cuda")
for _ in range(3): x.zero_()
ts = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); x.zero_(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
ts.sort()
print("fill 2 GiB: median %.3f ms = %.0f GB/s written" % (ts[5], (2 << 30) / ts[5] / 1e6))
y = torch.empty_like(x)
for _ in range(3): y.copy_(x)
ts = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); y.copy_(x); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
ts.sort()
print("copy 2 GiB: median %.3f ms = %.0f GB/s read+written" % (ts[5], 2 * (2 << 30) / ts[5] / 1e6))
