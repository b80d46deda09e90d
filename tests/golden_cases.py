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


def reproject_case():
    """geom_case with intrinsics that differ per frame, a stereo edge, edges into / out of the far-away frame 7 and a
    noisy target field (motion features)."""
    q = geom_case()
    n = q["poses"].shape[0]
    K = np.tile(q["intr"][None], (n, 1)).astype(np.float32)
    K *= (1.0 + 0.01 * np.arange(n, dtype=np.float32))[:, None]
    ii = np.concatenate([q["ii"], [3, 7, 6]]).astype(np.int64)
    jj = np.concatenate([q["jj"], [3, 6, 7]]).astype(np.int64)
    ht, wd = q["disps"].shape[1:]
    rng = np.random.default_rng(31)
    y, x = np.meshgrid(np.arange(ht, dtype=np.float32), np.arange(wd, dtype=np.float32), indexing="ij")
    target = (np.stack([x, y], -1)[None] + rng.normal(0, 30.0, size=(len(ii), ht, wd, 2))).astype(np.float32)
    return dict(poses=q["poses"], disps=q["disps"], intr=K, ii=ii, jj=jj, target=target)


FILTER_CASES = {"frontend": (60, 48, 30, 1), "backend": (4000, 9000, 400, 2), "empty_graph": (40, 0, 12, 3),
                "nothing_new": (0, 50, 20, 4), "all_known": (25, 25, 10, 5)}


def filter_case(key):
    """(candidates, existing edges) for FactorGraph.__filter_repeated_edges: random pairs over t frames, about half of
    the candidates already in the graph, duplicates inside the candidate list."""
    n, ne, t, seed = FILTER_CASES[key]
    rng = np.random.default_rng(seed)
    ie = rng.integers(0, t, size=ne).astype(np.int64)
    je = np.clip(ie + rng.integers(-5, 6, size=ne), 0, t - 1).astype(np.int64)
    ii = rng.integers(0, t, size=n).astype(np.int64)
    jj = np.clip(ii + rng.integers(-5, 6, size=n), 0, t - 1).astype(np.int64)
    if ne > 0 and n > 0:
        k = rng.integers(0, ne, size=n)
        m = rng.uniform(size=n) < 0.5
        ii[m], jj[m] = ie[k][m], je[k][m]
    if key == "all_known":
        ii, jj = ie.copy(), je.copy()
    if n > 4:
        ii[3], jj[3] = ii[1], jj[1]                                    # a duplicate inside the candidate list
    return dict(ii=ii, jj=jj, ii_exist=ie, jj_exist=je)


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
