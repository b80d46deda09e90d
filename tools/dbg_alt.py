import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/droid-slam_b200"); sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo/oracle")
import numpy as np, torch
import droid_backends as B
import test_gpu_parity as T
p, pyr, coords = T._tc_case(5, 40, 64, 47, True)
ii, jj = T.cu(p["ii"]), T.cu(p["jj"])
for l in range(4):
    c = (coords / 2 ** l).contiguous()
    f1, f2 = pyr[0][ii].float().contiguous(), pyr[l][jj].float().contiguous()
    torch.cuda.synchronize(); print("level", l, flush=True)
    plain, = B.altcorr_forward(f1, f2, c, 3, _plain=True); torch.cuda.synchronize(); print(" plain ok", flush=True)
    ix = B.altcorr_forward_indexed(pyr[0], pyr[l], ii, jj, c, 3); torch.cuda.synchronize(); print(" indexed ok", flush=True)
    tc, = B.altcorr_forward(f1, f2, c, 3); torch.cuda.synchronize(); print(" ws ok", flush=True)
    fin = torch.isfinite(plain)
    print(" max diff", float((tc[fin] - plain[fin]).abs().max()), float(plain[fin].abs().max()), bool(torch.equal(tc[fin], ix[fin])))
