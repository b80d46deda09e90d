"""Seeded cases for cvx_upsample (droid_slam/droid_net.py:21-35), shared by tests/golden/make_golden_upsample.py and the
tests.  mask: [batch, 9*8*8, ht, wd] logits, data: [batch, ht, wd, dim]."""
import hashlib

import numpy as np

CASES = {
    "disp_small": dict(batch=2, ht=6, wd=10, dim=1, seed=1, scale=1.0),
    "disp_sharp": dict(batch=1, ht=5, wd=33, dim=1, seed=2, scale=12.0),     # near one-hot softmax, ragged width
    "flow2": dict(batch=2, ht=4, wd=7, dim=2, seed=3, scale=2.0),
}


def make_case(key):
    c = CASES[key]
    rng = np.random.default_rng(c["seed"])
    data = rng.uniform(0.1, 2.0, size=(c["batch"], c["ht"], c["wd"], c["dim"])).astype(np.float32)
    mask = (rng.normal(size=(c["batch"], 9 * 64, c["ht"], c["wd"])) * c["scale"]).astype(np.float32)
    return data, mask


def digest(data, mask):
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(data).tobytes()); h.update(np.ascontiguousarray(mask).tobytes())
    return h.hexdigest()[:16]
