"""Seeded input cases shared by tests/golden/make_golden_gpu.py (which runs the UNMODIFIED reference
extension on the GPU box and stores its outputs) and the tests that check the oracle / the CUDA path
against those stored outputs."""
import hashlib

import numpy as np

import synth


def digest(arrs):
    h = hashlib.sha256()
    for a in arrs:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()[:16]


def corr_case(dtype, seed=11):
    rng = np.random.default_rng(seed)
    B, H1, W1, H2, W2 = 2, 8, 12, 16, 24
    vol = rng.normal(size=(B, H1, W1, H2, W2)).astype(dtype)
    c = synth.make_corr_problem(B, H1, W1, 8, seed=seed, oob_frac=0.1)["coords"]
    c = (c * np.float32(W2 / W1)).astype(np.float32)
    g = rng.normal(size=(B, 7, 7, H1, W1)).astype(dtype)
    return vol, c, g


def altcorr_case(seed=12):
    p = synth.make_corr_problem(2, 8, 16, 64, seed=seed, oob_frac=0.05, dtype=np.float32)
    f = (p["fmaps"] / 4.0).astype(np.float32).transpose(0, 2, 3, 1)
    f1 = np.ascontiguousarray(f[p["ii"]]); f2 = np.ascontiguousarray(f[p["jj"]])
    coords = np.ascontiguousarray(p["coords"].transpose(0, 2, 3, 1)[:, None])
    g = np.random.default_rng(seed).normal(size=(2, 1, 49, 8, 16)).astype(np.float32)
    return f1, f2, coords, g


def geom_case():
    p = synth.make_ba_problem("C1", seed=21)
    poses = p["poses"].copy(); poses[7, :3] += [0, 0, 30.0]
    n = poses.shape[0]
    a, b = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    return dict(poses=poses, disps=p["disps"], intr=p["intrinsics"], ii=p["ii"][:6].copy(), jj=p["jj"][:6].copy(),
                fa=a.reshape(-1).astype(np.int64), fb=b.reshape(-1).astype(np.int64),
                ix=np.arange(n, dtype=np.int64), thresh=np.full(n, 0.05, np.float32),
                gt_poses=p["gt_poses"], gt_disps=p["gt_disps"])


BA_CASES = {
    "ba_C1": dict(name="C1", kw={}, iters=2, lm=1e-4, ep=0.1, motion_only=False),
    "ba_C1_1it": dict(name="C1", kw={}, iters=1, lm=1e-4, ep=0.1, motion_only=False),
    "ba_C3": dict(name="C3", kw={}, iters=2, lm=1e-4, ep=0.1, motion_only=False),
    "ba_C1_stereo_sensor": dict(name="C1", kw=dict(stereo=True, sensor=True), iters=2, lm=1e-4, ep=0.1, motion_only=False),
    "ba_C3_motion_only": dict(name="C3", kw={}, iters=2, lm=1e-4, ep=0.1, motion_only=True),
    "ba_C4_24_backend": dict(name="C4", kw=dict(n_frames=24), iters=2, lm=1e-5, ep=1e-2, motion_only=False, t0=5),
}


def ba_case(key):
    c = BA_CASES[key]
    p = synth.make_ba_problem(c["name"], **c["kw"])
    if "t0" in c:
        p["t0"] = c["t0"]
    return p, c
