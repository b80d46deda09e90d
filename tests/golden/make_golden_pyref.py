"""Generator of tests/golden/ref_reproject.npz and ref_filter_edges.npz -- run in the build container (CPU).

Runs the UNMODIFIED Python reference -- droid_slam/geom/projective_ops.py:96-125 (projective_transform, what
DepthVideo.reproject calls, depth_video.py:139-147) on top of the reference's vendored lietorch (built for the host by
oracle/build_lietorch.sh, staged by oracle/stage_pyref.sh, loaded by oracle/pyref.py) -- on the seeded geometry case of
tests/golden_cases.py with per-frame intrinsics, a frame pushed behind its neighbours (Z < 0.1 fallback, invalid
pixels) and a stereo edge (ii == jj), and stores coords / valid plus the motion features of FactorGraph.update
(factor_graph.py:203-204, evaluated with the reference's own two lines).  This is the reference-run pin of SURVEY 8f
N1: tests/test_oracle.py holds the restatement O.reproject to it on CPU, tests/test_gpu_parity.py the CUDA op.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (os.path.join(ROOT, "droid-slam_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import golden_cases as G  # noqa: E402
import pyref  # noqa: E402


def main():
    ns = pyref.load(backend=type(sys)("droid_backends_unused"), cpu=True)
    q = G.reproject_case()
    Gs = ns.lietorch.SE3(torch.from_numpy(q["poses"])[None])
    ii, jj = torch.from_numpy(q["ii"]), torch.from_numpy(q["jj"])
    coords, valid = ns.pops.projective_transform(Gs, torch.from_numpy(q["disps"])[None], torch.from_numpy(q["intr"])[None], ii, jj)
    ht, wd = q["disps"].shape[1:]
    # factor_graph.py:24-28,203-204
    coords0 = ns.pops.coords_grid(ht, wd)
    target = torch.from_numpy(q["target"])[None]
    motn = torch.cat([coords - coords0, target - coords], dim=-1)
    motn = motn.permute(0, 1, 4, 2, 3).clamp(-64.0, 64.0)
    np.savez_compressed(os.path.join(HERE, "ref_reproject.npz"), coords=coords[0].numpy(), valid=valid[0].numpy(),
                        motn=motn[0].numpy(), digest=G.digest([q["poses"], q["disps"], q["intr"], q["target"]]))
    print("wrote ref_reproject.npz", coords.shape, float(valid.mean()))

    # FactorGraph.__filter_repeated_edges (factor_graph.py:43-54), the unmodified method on a plain namespace as `self`
    import types
    FG = ns.factor_graph.FactorGraph
    store = {}
    for key in sorted(G.FILTER_CASES):
        c = G.filter_case(key)
        as_t = lambda a: torch.as_tensor(a, dtype=torch.long)  # noqa: E731
        half = len(c["ii_exist"]) // 2
        self = types.SimpleNamespace(ii=as_t(c["ii_exist"][:half]), jj=as_t(c["jj_exist"][:half]),
                                     ii_inac=as_t(c["ii_exist"][half:]), jj_inac=as_t(c["jj_exist"][half:]))
        ki, kj = FG._FactorGraph__filter_repeated_edges(self, as_t(c["ii"]), as_t(c["jj"]))
        store[key + "_ii"], store[key + "_jj"] = ki.numpy().astype(np.int64), kj.numpy().astype(np.int64)
        store[key + "_digest"] = G.digest([c["ii"], c["jj"], c["ii_exist"], c["jj_exist"]])
        print(key, len(c["ii"]), "->", len(ki))
    np.savez_compressed(os.path.join(HERE, "ref_filter_edges.npz"), **store)


if __name__ == "__main__":
    main()
