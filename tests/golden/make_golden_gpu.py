"""Runs the UNMODIFIED reference `droid_backends` extension (oracle/_ref, built by oracle/build_ref.sh from
/root/reference/src where the sources lie) on the GPU box for the seeded cases of tests/golden_cases.py and
stores its outputs under tests/golden/ref_*.npz.  Also records how this repo's CUDA path and the CPU oracle
compare with the reference on the same inputs (gpurun_out/ref_compare.json).

Run on the GPU box:  python tests/golden/make_golden_gpu.py   (outputs also copied to gpurun_out/golden/)
"""
import json
import os
import shutil
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (os.path.join(ROOT, "droid-slam_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "oracle", "_ref"),
          os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import droid_backends as mine  # noqa: E402
import droid_backends_ref as ref  # noqa: E402
import droid_oracle as O  # noqa: E402
import golden_cases as G  # noqa: E402

DEV = "cuda:0"
report = {}


def cu(x):
    return torch.from_numpy(np.ascontiguousarray(x)).to(DEV)


def cmp(name, got, want):
    got = np.asarray(got, np.float64); want = np.asarray(want, np.float64)
    r = float(np.abs(got - want).max() / max(np.abs(want).max(), 1e-30))
    report[name] = dict(rel=r, bit_exact=bool(np.array_equal(got, want)), frac_diff=float(np.mean(got != want)))
    print("%-44s rel %.3e  exact %s" % (name, r, report[name]["bit_exact"]))


def save(name, **arrs):
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **arrs)


def main():
    for dt, tag in ((np.float32, "f32"), (np.float16, "f16")):
        vol, c, g = G.corr_case(dt)
        r, = ref.corr_index_forward(cu(vol), cu(c), 3)
        rb, = ref.corr_index_backward(cu(vol), cu(c), cu(g), 3)
        m, = mine.corr_index_forward(cu(vol), cu(c), 3)
        mb, = mine.corr_index_backward(cu(vol), cu(c), cu(g), 3)
        save("ref_corr_index_" + tag, fwd=r.cpu().numpy(), bwd=rb.cpu().numpy(), digest=G.digest([vol, c, g]))
        cmp("corr_index_forward_%s mine-vs-ref" % tag, m.float().cpu().numpy(), r.float().cpu().numpy())
        cmp("corr_index_backward_%s mine-vs-ref" % tag, mb.float().cpu().numpy(), rb.float().cpu().numpy())
        cmp("corr_index_forward_%s oracle-vs-ref" % tag, O.corr_index_forward(vol, c, 3).astype(np.float32), r.float().cpu().numpy())
    f1, f2, coords, g = G.altcorr_case()
    r, = ref.altcorr_forward(cu(f1), cu(f2), cu(coords), 3)
    rb = ref.altcorr_backward(cu(f1), cu(f2), cu(coords), cu(g), 3)
    m, = mine.altcorr_forward(cu(f1), cu(f2), cu(coords), 3)
    mb = mine.altcorr_backward(cu(f1), cu(f2), cu(coords), cu(g), 3)
    save("ref_altcorr", fwd=r.cpu().numpy(), g1=rb[0].cpu().numpy(), g2=rb[1].cpu().numpy(), gc=rb[2].cpu().numpy(),
         digest=G.digest([f1, f2, coords, g]))
    cmp("altcorr_forward mine-vs-ref", m.cpu().numpy(), r.cpu().numpy())
    cmp("altcorr_forward oracle-vs-ref", O.altcorr_forward(f1, f2, coords, 3), r.cpu().numpy())
    cmp("altcorr_backward.f1 mine-vs-ref", mb[0].cpu().numpy(), rb[0].cpu().numpy())
    cmp("altcorr_backward.f2 mine-vs-ref", mb[1].cpu().numpy(), rb[1].cpu().numpy())

    q = G.geom_case()
    P, D, K = cu(q["poses"]), cu(q["disps"]), cu(q["intr"])
    rc, rv = ref.projmap(P, D, K, cu(q["ii"]), cu(q["jj"]))
    mc, mv = mine.projmap(P, D, K, cu(q["ii"]), cu(q["jj"]))
    rp = ref.iproj(P, D, K); mp = mine.iproj(P, D, K)
    rf = ref.frame_distance(P, D, K, cu(q["fa"]), cu(q["fb"]), 0.3)
    mf = mine.frame_distance(P, D, K, cu(q["fa"]), cu(q["fb"]), 0.3)
    GP, GD = cu(q["gt_poses"]), cu(q["gt_disps"])
    rdf = ref.depth_filter(GP, GD, K, cu(q["ix"]), cu(q["thresh"]))
    mdf = mine.depth_filter(GP, GD, K, cu(q["ix"]), cu(q["thresh"]))
    save("ref_geom", coords=rc.cpu().numpy(), valid=rv.cpu().numpy(), points=rp[:2].cpu().numpy(),
         dist=rf.cpu().numpy(), counter=rdf.cpu().numpy(), digest=G.digest([q["poses"], q["disps"]]))
    cmp("projmap.coords mine-vs-ref", mc.cpu().numpy(), rc.cpu().numpy())
    cmp("projmap.valid mine-vs-ref", mv.cpu().numpy(), rv.cpu().numpy())
    cmp("iproj mine-vs-ref", mp.cpu().numpy(), rp.cpu().numpy())
    cmp("frame_distance mine-vs-ref", mf.cpu().numpy(), rf.cpu().numpy())
    cmp("frame_distance oracle-vs-ref", O.frame_distance(q["poses"], q["disps"], q["intr"], q["fa"], q["fb"], 0.3), rf.cpu().numpy())
    cmp("depth_filter mine-vs-ref", mdf.cpu().numpy(), rdf.cpu().numpy())
    cmp("depth_filter oracle-vs-ref", O.depth_filter(q["gt_poses"], q["gt_disps"], q["intr"], q["ix"], q["thresh"]), rdf.cpu().numpy())

    for key in G.BA_CASES:
        p, c = G.ba_case(key)

        def call(mod, poses0, disps0, iters):
            poses, disps = cu(poses0), cu(disps0)
            o = mod.ba(poses, disps, cu(p["intrinsics"]), cu(p["disps_sens"]), cu(p["targets"]), cu(p["weights"]),
                       cu(p["eta"]), cu(p["ii"]), cu(p["jj"]), p["t0"], p["t1"], iters, c["lm"], c["ep"], c["motion_only"])
            torch.cuda.synchronize()
            return (poses.cpu().numpy(), disps.cpu().numpy(), o[0].cpu().numpy(),
                    None if c["motion_only"] else o[1].cpu().numpy().reshape(-1, p["ht"] * p["wd"]))

        outs = {tag: call(mod, p["poses"], p["disps"], c["iters"]) for tag, mod in (("ref", ref), ("mine", mine))}
        rp_, rd_, rdx, rdz = outs["ref"]
        extra = {}
        if c["iters"] == 2:
            # the reference's state and updates after its FIRST iteration: lets every test compare iteration 2 from
            # identical inputs (north_star: <= 1e-4 per Gauss-Newton iteration)
            p1, d1, dx1, dz1 = call(ref, p["poses"], p["disps"], 1)
            extra = dict(poses_it1=p1, disps_it1=d1, dx_it1=dx1, dz_it1=(dz1 if dz1 is not None else np.zeros(0, np.float32)))
            p2, d2, dx2, dz2 = call(ref, p1, d1, 1)
            cmp(key + ".ref 1+1 iterations vs one call of 2 (dx)", dx2, rdx)
            m2 = call(mine, p1, d1, 1)
            cmp(key + ".dx iteration 2 from the reference's state mine-vs-ref", m2[2], dx2)
            if dz2 is not None:
                cmp(key + ".dz iteration 2 from the reference's state mine-vs-ref", m2[3], dz2)
        save("ref_" + key, poses=rp_, disps=rd_, dx=rdx, dz=(rdz if rdz is not None else np.zeros(0, np.float32)),
             digest=G.digest([p["poses"], p["disps"], p["targets"], p["weights"], p["eta"]]), **extra)
        mp_, md_, mdx, mdz = outs["mine"]
        cmp(key + ".dx mine-vs-ref", mdx, rdx)
        cmp(key + ".poses mine-vs-ref", mp_, rp_)
        if rdz is not None:
            cmp(key + ".dz mine-vs-ref", mdz, rdz)
            cmp(key + ".disps mine-vs-ref", md_, rd_)
        op, od = p["poses"].copy(), p["disps"].copy()
        odx, odz = O.ba(op, od, p["intrinsics"], p["disps_sens"], p["targets"], p["weights"], p["eta"], p["ii"], p["jj"],
                        p["t0"], p["t1"], c["iters"], c["lm"], c["ep"], c["motion_only"])
        cmp(key + ".dx oracle-vs-ref", odx, rdx)
        if rdz is not None:
            cmp(key + ".dz oracle-vs-ref", odz, rdz)
    out = os.path.join(ROOT, "gpurun_out", "golden")
    os.makedirs(out, exist_ok=True)
    for f in os.listdir(HERE):
        if f.endswith(".npz"):
            shutil.copy(os.path.join(HERE, f), out)
    json.dump(report, open(os.path.join(ROOT, "gpurun_out", "ref_compare.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
