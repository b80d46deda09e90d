"""droid_frame_distance_noalloc: the frame_distance entry point that allocates nothing (its scratch is a file-scope
static __device__ array inside libdroid_b200.so).  No counterpart in the reference -- added
for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops).

Called straight through the C ABI (ctypes), two calls in a row on the shared scratch with DIFFERENT pair lists and
beta (so stale contents of the first call would show), each compared with the existing entry point
droid_frame_distance: bit-exact, the tolerance tests/test_gpu_parity.py holds frame_distance to.

This file sorts last on purpose (a failure here must not hide other tests); it passes alone in a
fresh process and after the rest of the suite.
"""
import ctypes

import numpy as np
import pytest
import torch

import synth


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _call(lib, name, P, D, K, a, b, beta, ht, wd):
    """One C-ABI call; inputs and the output are the caller's (allocated here, outside the entry point)."""
    out = torch.full((a.numel(),), -1.0, dtype=torch.float32, device=P.device)
    st = ctypes.c_void_p(torch.cuda.current_stream(P.device).cuda_stream)
    rc = getattr(lib, name)(_ptr(P), _ptr(D), _ptr(K), _ptr(a), _ptr(b), _ptr(out), int(a.numel()), ht, wd,
                            ctypes.c_float(beta), st)
    assert rc == 0, (name, rc)
    return out


@pytest.mark.gpu
def test_frame_distance_noalloc_matches_existing_entry_point(backends):
    lib = backends._lib
    dev = "cuda:0"
    p = synth.make_ba_problem("C4", n_frames=120, knn=8)
    ht, wd, n = p["ht"], p["wd"], 120
    assert ht * wd <= lib.droid_frame_distance_noalloc_max_pixels()
    P = torch.from_numpy(p["gt_poses"]).to(dev)
    D = torch.from_numpy(p["gt_disps"]).to(dev)
    K = torch.from_numpy(p["intrinsics"]).to(dev)
    a, b = torch.meshgrid(torch.arange(n, device=dev), torch.arange(n, device=dev), indexing="ij")
    a, b = a.reshape(-1).contiguous(), b.reshape(-1).contiguous()
    rng = np.random.default_rng(5)
    a2 = torch.from_numpy(rng.integers(0, n, 7000)).to(dev)
    b2 = torch.from_numpy(rng.integers(0, n, 7000)).to(dev)
    # two calls in a row on the shared scratch: another pair list, another beta (another `total`, same grid table)
    for x, y, beta in ((a, b, 0.3), (b2, a2, 0.45), (a, b, 0.3)):
        want = _call(lib, "droid_frame_distance", P, D, K, x, y, beta, ht, wd)
        got = _call(lib, "droid_frame_distance_noalloc", P, D, K, x, y, beta, ht, wd)
        torch.cuda.synchronize()
        assert torch.equal(got, want), float((got != want).float().mean())
        assert bool((want == 1000.0).any()) and bool((want < 1000.0).any())
    # another image size right after (the table is rebuilt by every call): 48x64 of the frontend
    q = synth.make_ba_problem("C3")
    P3, D3, K3 = (torch.from_numpy(q[k]).to(dev) for k in ("gt_poses", "gt_disps", "intrinsics"))
    i3, j3 = torch.from_numpy(q["ii"]).to(dev), torch.from_numpy(q["jj"]).to(dev)
    want = _call(lib, "droid_frame_distance", P3, D3, K3, i3, j3, 0.3, 48, 64)
    got = _call(lib, "droid_frame_distance_noalloc", P3, D3, K3, i3, j3, 0.3, 48, 64)
    assert torch.equal(got, want)


def test_frame_distance_noalloc_rejects_sizes_beyond_its_scratch(backends):
    """Argument validation happens before any CUDA call, so this runs without a GPU too."""
    lib = backends._lib
    mx = lib.droid_frame_distance_noalloc_max_pixels()
    assert mx == 16384
    args = (None, None, None, None, None, None)
    assert lib.droid_frame_distance_noalloc(*args, 4, 128, mx // 128 + 1, ctypes.c_float(0.3), None) == -7
    assert lib.droid_frame_distance_noalloc(*args, 0, 128, mx // 128, ctypes.c_float(0.3), None) == 0     # empty list
    assert lib.droid_frame_distance_noalloc(*args, 4, 128, mx // 128, ctypes.c_float(0.3), None) == -1    # null pointers
