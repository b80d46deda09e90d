"""Seeded cases for the proximity-edge selection (FactorGraph.add_proximity_factors), shared by
tests/golden/make_golden_proximity.py (which runs the unmodified reference method on them) and the tests."""
import hashlib

import numpy as np

# how the reference calls the method: frontend droid_frontend.py:68-70 / 106-107 (window 25, radius 2, nms 1,
# thresh 16, max_factors 48), backend droid_backend.py:33-38 (radius 2, nms 3, thresh 22, max_factors 16 t)
CASES = {
    "frontend": dict(t=30, t0=25, t1=5, rad=2, nms=1, thresh=16.0, max_factors=48, stereo=False, n_exist=30, seed=1),
    "frontend_early": dict(t=9, t0=4, t1=0, rad=2, nms=1, thresh=16.0, max_factors=48, stereo=False, n_exist=6, seed=2),
    "backend": dict(t=120, t0=0, t1=0, rad=2, nms=3, thresh=22.0, max_factors=16 * 120, stereo=False, n_exist=0, seed=3),
    "backend_capped": dict(t=90, t0=0, t1=0, rad=2, nms=2, thresh=22.0, max_factors=400, stereo=False, n_exist=40, seed=4),
    "stereo": dict(t=24, t0=0, t1=0, rad=1, nms=2, thresh=18.0, max_factors=200, stereo=True, n_exist=10, seed=5),
    "nothing_close": dict(t=16, t0=0, t1=0, rad=2, nms=2, thresh=0.5, max_factors=100, stereo=False, n_exist=0, seed=6),
}


def make_case(key):
    c = dict(CASES[key])
    rng = np.random.default_rng(c["seed"])
    t, t0, t1 = c["t"], c["t0"], c["t1"]
    # a looping trajectory: frames that are far apart in time come close in space, like the backend sees them
    s = np.arange(t) * (2 * np.pi / max(t // 2, 8))
    pos = np.stack([10 * np.cos(s), 10 * np.sin(s), 0.05 * np.arange(t)], -1) + rng.normal(0, 0.3, size=(t, 3))
    ii, jj = np.meshgrid(np.arange(t0, t), np.arange(t1, t), indexing="ij")
    ii, jj = ii.reshape(-1), jj.reshape(-1)
    d = np.linalg.norm(pos[ii] - pos[jj], axis=-1) * 2.5 + rng.uniform(0, 1e-3, size=len(ii))
    d[rng.uniform(size=len(d)) < 0.05] = 1000.0            # frame_distance's "too few valid pixels" value
    c["d"] = d.astype(np.float32)
    n = c["n_exist"]
    c["ii_exist"] = rng.integers(0, t, size=n).astype(np.int64)
    c["jj_exist"] = np.clip(c["ii_exist"] + rng.integers(-6, 7, size=n), 0, t - 1).astype(np.int64)
    return c


def digest(case):
    h = hashlib.sha256()
    for k in ("d", "ii_exist", "jj_exist"):
        h.update(np.ascontiguousarray(case[k]).tobytes())
    h.update(repr([case[k] for k in ("t", "t0", "t1", "rad", "nms", "thresh", "max_factors", "stereo")]).encode())
    return h.hexdigest()[:16]
