"""Generator of tests/golden/truth64_ba_C4_<frames>.npz -- run in the build container (CPU, minutes).

The oracle's BA algorithm (oracle/droid_oracle.py, a restatement of src/droid_kernels.cu:176-424,:1314-1434) evaluated
in FLOAT64 THROUGHOUT on the headline problem (synth.make_ba_problem("C4", n_frames=1000, knn=16): 1000 keyframes,
20 862 edges, 40x64; lm=1e-5, ep=1e-2), ONE Gauss-Newton iteration.  It is the rounding-error yardstick of
tests/test_gpu_parity.py::test_ba_headline_parity_vs_reference_extension: the unmodified reference extension and this
repo's CUDA path are two float32 evaluations of the same algorithm; on this ill-conditioned system they can differ
from each other by more than either differs from the exact answer, so each is also held against this file.

Stored: dx [P,6] float64, dz of the frames in FRAMES [len,HW] float64, poses after the iteration.
Usage: python tests/golden/make_truth64_ba.py [frames]
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (os.path.join(ROOT, "droid-slam_b200"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)

import droid_oracle as O  # noqa: E402
import synth  # noqa: E402

FRAMES = (0, 1, 2, 17, 249, 250, 251, 499, 500, 777, 998, 999)


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    p = synth.make_ba_problem("C4", n_frames=n, knn=16)
    O.F32, O.MIN_DEPTH = np.float64, np.float64(0.25)
    q = {k: (v.astype(np.float64) if isinstance(v, np.ndarray) and v.dtype == np.float32 else v) for k, v in p.items()}
    t = time.time()
    dx, dz = O.ba(q["poses"], q["disps"], q["intrinsics"], q["disps_sens"], q["targets"], q["weights"], q["eta"],
                  q["ii"], q["jj"], q["t0"], q["t1"], 1, 1e-5, 1e-2, False)
    fr = np.array([f for f in FRAMES if f < n], np.int64)
    out = os.path.join(HERE, "truth64_ba_C4_%d.npz" % n)
    np.savez_compressed(out, dx=dx, dz=dz[fr].astype(np.float32), frames=fr, poses=q["poses"],
                        edges=np.int64(len(p["ii"])), seconds=time.time() - t)
    print("wrote", out, "in %.0f s" % (time.time() - t), "dx max", np.abs(dx).max())


if __name__ == "__main__":
    main()
