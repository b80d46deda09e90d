"""Small invocations of the hand-rolled mbarrier / TMEM / multi-stream kernels for compute-sanitizer
(tools/sanitize.sh): python tools/sanitize_cases.py <case>.  Cases: corr_build, ba_frontend (tensor-core Schur +
one-CTA solver), ba_chains (chains + separator Cholesky on side streams), chol_chains (the solver alone),
edges (proximity selection + filter), lookup (corr_index + alt-corr), geom."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "droid-slam_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import droid_backends as B  # noqa: E402
import synth  # noqa: E402

DEV = "cuda:0"
cu = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(DEV)  # noqa: E731


def ba(p, iters, lm, ep):
    po, di = cu(p["poses"]), cu(p["disps"])
    B.ba(po, di, cu(p["intrinsics"]), cu(p["disps_sens"]), cu(p["targets"]), cu(p["weights"]), cu(p["eta"]),
         cu(p["ii"]), cu(p["jj"]), p["t0"], p["t1"], iters, lm, ep, False)
    torch.cuda.synchronize()
    assert B.ba_status() == 0 and bool(torch.isfinite(po).all())


def main(case):
    if case == "corr_build":
        q = synth.make_corr_problem(3, 16, 32, 128, seed=1)
        lv = B.corr_build(cu(q["fmaps"][q["ii"]]), cu(q["fmaps"][q["jj"]]))
        q = synth.make_corr_problem(2, 48, 64, 128, seed=2)
        lv = B.corr_build(cu(q["fmaps"][q["ii"]]), cu(q["fmaps"][q["jj"]]))
        torch.cuda.synchronize()
        assert bool(torch.isfinite(lv[3].float()).all())
    elif case == "ba_frontend":
        ba(synth.make_ba_problem("C3"), 2, 1e-4, 0.1)
    elif case == "ba_chains":
        ba(synth.make_ba_problem("C4", n_frames=280), 1, 1e-5, 1e-2)
        assert len(B.ba_chains()) >= 2
    elif case == "chol_chains":
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        rng = np.random.default_rng(0)
        n, ends, bw = 1500, (400, 902), 90
        k = len(ends); b0 = [0] + list(ends) + [n]
        NONE = np.iinfo(np.int32).max
        firstc = np.full((k + 1, n), NONE, np.int64)
        A = np.zeros((n, n))
        for c in range(k):
            for r in range(b0[c], b0[c + 1]):
                f = max(b0[c], r - bw); firstc[c, r] = f; A[r, f:r] = rng.normal(size=r - f) * 0.1
        for r in range(b0[k], n):
            firstc[k, r] = b0[k]; A[r, b0[k]:r] = rng.normal(size=r - b0[k]) * 0.1
            for c in range(k):
                f = int(rng.integers(max(b0[c], b0[c + 1] - 3 * bw), b0[c + 1])); firstc[c, r] = f
                A[r, f:b0[c + 1]] = rng.normal(size=b0[c + 1] - f) * 0.1
        A = A + A.T; A += np.diag(np.abs(A).sum(1) * 0.6 + 1.0)
        b = rng.normal(size=n)
        x, info = B.chol_solve(cu(A), cu(b), cu(firstc.astype(np.int32)), chain_end=list(ends))
        assert int(info.item()) == 0 and np.abs(x.cpu().numpy() - np.linalg.solve(A, b)).max() < 1e-9
        x, info = B.chol_solve(cu(A[:700, :700]), cu(b[:700]))          # dense blocked path
        assert int(info.item()) == 0
    elif case == "edges":
        import proximity_cases as PC
        c = PC.make_case("backend_capped")
        ii, jj = B.proximity_edges(cu(c["d"]), c["t"], c["t0"], c["t1"], c["rad"], c["nms"], c["thresh"], c["max_factors"],
                                   c["stereo"], cu(c["ii_exist"]), cu(c["jj_exist"]))
        B.filter_repeated_edges(ii, jj, cu(c["ii_exist"]), cu(c["jj_exist"]))
    elif case == "lookup":
        q = synth.make_corr_problem(4, 16, 24, 128, seed=4, oob_frac=0.1)
        vols = [torch.randn((4, 16, 24, 16 >> l, 24 >> l), device=DEV).half() for l in range(4)]
        B.corr_index_pyramid(vols, cu(q["coords"]), 3)
        B.corr_index_forward(vols[0], cu(q["coords"]), 3)
        f = cu((q["fmaps"] / 4).astype(np.float32).transpose(0, 2, 3, 1))
        B.altcorr_forward(f[cu(q["ii"])].contiguous(), f[cu(q["jj"])].contiguous(),
                          cu(np.ascontiguousarray(q["coords"].transpose(0, 2, 3, 1)[:, None])), 3)
    elif case == "geom":
        p = synth.make_ba_problem("C1")
        P, D, K = cu(p["poses"]), cu(p["disps"]), cu(p["intrinsics"])
        ii, jj = cu(p["ii"]), cu(p["jj"])
        B.frame_distance(P, D, K, ii, jj, 0.3); B.projmap(P, D, K, ii, jj); B.iproj(P, D, K)
        B.depth_filter(P, D, K, cu(np.arange(8, dtype=np.int64)), cu(np.full(8, 0.05, np.float32)))
        B.reproject(P, D, cu(np.tile(p["intrinsics"][None], (8, 1))), ii, jj)
    else:
        raise SystemExit("unknown case " + case)
    torch.cuda.synchronize()
    print("case", case, "ok")


if __name__ == "__main__":
    main(sys.argv[1])
