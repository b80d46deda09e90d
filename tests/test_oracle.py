"""CPU tests (no GPU): the numpy oracle against the committed outputs of the UNMODIFIED reference extension
(tests/golden/ref_*.npz, produced on a B200 by tests/golden/make_golden_gpu.py), plus properties of the ops.
This is what pins the oracle: the reference ships no fixtures of its own."""
import os

import numpy as np
import pytest

import droid_oracle as O
import golden_cases as G
import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def gold(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


@pytest.mark.parametrize("tag,dtype", [("f32", np.float32), ("f16", np.float16)])
def test_corr_index_vs_reference(tag, dtype):
    vol, c, g = G.corr_case(dtype)
    ref = gold("ref_corr_index_" + tag)
    assert str(ref["digest"]) == G.digest([vol, c, g])
    fwd = O.corr_index_forward(vol, c, 3)
    bwd = O.corr_index_backward(vol.shape, dtype, c, g, 3)
    if dtype == np.float16:          # per-op half rounding reproduced exactly
        assert np.array_equal(fwd, ref["fwd"]) and np.array_equal(bwd, ref["bwd"])
    else:                            # the GPU contracts s*w + acc into FFMA
        assert rel(fwd, ref["fwd"]) <= 1e-6 and rel(bwd, ref["bwd"]) <= 1e-6


def test_altcorr_vs_reference():
    f1, f2, coords, g = G.altcorr_case()
    ref = gold("ref_altcorr")
    assert str(ref["digest"]) == G.digest([f1, f2, coords, g])
    assert rel(O.altcorr_forward(f1, f2, coords, 3), ref["fwd"]) <= 1e-6
    g1, g2, gc = O.altcorr_backward(f1, f2, coords, g, 3)
    assert rel(g1, ref["g1"]) <= 1e-5 and rel(g2, ref["g2"]) <= 1e-5
    assert np.abs(ref["gc"]).max() == 0 and np.abs(gc).max() == 0      # coords_grad is never written


def test_geometry_vs_reference():
    q = G.geom_case()
    ref = gold("ref_geom")
    assert str(ref["digest"]) == G.digest([q["poses"], q["disps"]])
    c, v = O.projmap(q["poses"], q["disps"], q["intr"], q["ii"], q["jj"])
    assert rel(c, ref["coords"]) <= 1e-6 and np.array_equal(v, ref["valid"])
    assert rel(O.iproj(q["poses"], q["disps"], q["intr"])[:2], ref["points"]) <= 1e-6
    d = O.frame_distance(q["poses"], q["disps"], q["intr"], q["fa"], q["fb"], 0.3)
    assert np.array_equal(d == 1000.0, ref["dist"] == 1000.0) and (d == 1000.0).any()
    assert rel(d, ref["dist"]) <= 1e-6
    cnt = O.depth_filter(q["gt_poses"], q["gt_disps"], q["intr"], q["ix"], q["thresh"])
    assert np.array_equal(cnt, ref["counter"])                         # integer counts: bit-exact


def test_reproject_oracle_pinned_through_projmap():
    """SURVEY 8f N1.  The Python path behind DepthVideo.reproject (geom/projective_ops.py:96-125) has no reference
    run that can travel (it needs lietorch), so its restatement is pinned through the reference extension's CUDA twin
    `projmap`, whose outputs are stored: wherever the point lies in front of the camera (Z > 0.25, valid in both) the
    two must give the same coordinates; the validity masks may differ only for 0.2 < Z <= 0.25 (MIN_DEPTH 0.2 vs
    0.25); behind the camera (Z < 0.1) the Python path projects with Z = 1 instead of returning the pixel itself."""
    q = G.geom_case()
    ref = gold("ref_geom")
    n = q["poses"].shape[0]
    intr = np.tile(q["intr"][None], (n, 1)).astype(np.float32)
    c, v = O.reproject(q["poses"], q["disps"], intr, q["ii"], q["jj"])
    both = (ref["valid"][..., 0] > 0) & (v[..., 0] > 0)
    assert both.mean() > 0.5
    assert np.abs(c - ref["coords"][..., :2])[both].max() <= 1e-4 * max(1.0, np.abs(ref["coords"]).max())
    assert not ((ref["valid"] > 0) & (v == 0)).any()                   # Z > 0.25 implies Z > 0.2
    # stereo edge: fixed baseline, identity rotation -> u shifts by -0.1 fx d, v unchanged
    ii = np.array([2], np.int64)
    cs, vs = O.reproject(q["poses"], q["disps"], intr, ii, ii)
    vg, ug = np.meshgrid(np.arange(cs.shape[1], dtype=np.float32), np.arange(cs.shape[2], dtype=np.float32), indexing="ij")
    assert np.abs(cs[0, ..., 0] - (ug - 0.1 * intr[2, 0] * q["disps"][2])).max() <= 1e-3
    assert np.abs(cs[0, ..., 1] - vg).max() <= 1e-3 and vs.min() == 1.0
    # motion features: clamp(cat(coords - grid, target - coords)), channel-first
    tgt = (c + 100.0).astype(np.float32)
    _, _, m = O.reproject(q["poses"], q["disps"], intr, q["ii"], q["jj"], target=tgt)
    assert m.shape == (len(q["ii"]), 4) + c.shape[1:3] and np.all(m[:, 2:] == 64.0)
    assert np.allclose(m[:, 0], np.clip(c[..., 0] - ug, -64, 64)) and np.allclose(m[:, 1], np.clip(c[..., 1] - vg, -64, 64))


def test_reproject_vs_python_reference():
    """SURVEY 8f N1, reference-run pin: O.reproject against stored outputs of the UNMODIFIED Python path
    (geom/projective_ops.py:96-125 on the reference's own lietorch, run on CPU by tests/golden/make_golden_pyref.py):
    per-frame intrinsics, Z < 0.1 fallback, stereo edge, motion features."""
    q = G.reproject_case()
    ref = gold("ref_reproject")
    assert str(ref["digest"]) == G.digest([q["poses"], q["disps"], q["intr"], q["target"]])
    c, v, m = O.reproject(q["poses"], q["disps"], q["intr"], q["ii"], q["jj"], target=q["target"])
    assert 0.0 < ref["valid"].mean() < 1.0 and (np.abs(ref["motn"]) == 64.0).any()
    assert rel(c, ref["coords"]) <= 1e-6 and np.array_equal(v, ref["valid"])
    assert np.abs(m - ref["motn"]).max() <= 1e-6 * np.abs(ref["coords"]).max()


@pytest.mark.parametrize("key", sorted(G.BA_CASES))
def test_ba_vs_reference(key):
    """<= 1e-4 relative PER Gauss-Newton iteration (north_star): iteration 1 from the initial state, iteration 2 from
    the state the reference extension stored after its first iteration (poses_it1 / disps_it1 in the golden file)."""
    p, c = G.ba_case(key)
    ref = gold("ref_" + key)
    assert str(ref["digest"]) == G.digest([p["poses"], p["disps"], p["targets"], p["weights"], p["eta"]])
    mo = c["motion_only"]

    def one(poses, disps):
        poses, disps = poses.copy(), disps.copy()
        dx, dz = O.ba(poses, disps, p["intrinsics"], p["disps_sens"], p["targets"], p["weights"], p["eta"], p["ii"],
                      p["jj"], p["t0"], p["t1"], 1, c["lm"], c["ep"], mo)
        return poses, disps, dx, dz
    poses, disps = p["poses"], p["disps"]
    if c["iters"] == 2:
        po, di, dx, dz = one(poses, disps)
        assert rel(dx, ref["dx_it1"]) <= 1e-4 and rel(po, ref["poses_it1"]) <= 1e-5
        if not mo:
            assert rel(dz, ref["dz_it1"]) <= 1e-4 and rel(di, ref["disps_it1"]) <= 1e-5
        poses, disps = ref["poses_it1"], ref["disps_it1"]
    po, di, dx, dz = one(poses, disps)
    assert rel(dx, ref["dx"]) <= 1e-4
    assert rel(po, ref["poses"]) <= 1e-5
    if not mo:
        assert rel(dz, ref["dz"]) <= 1e-4 and rel(di, ref["disps"]) <= 1e-5
    else:
        assert np.array_equal(di, p["disps"])


def test_corr_index_is_zero_padded_grid_sample():
    import torch
    import torch.nn.functional as F
    rng = np.random.default_rng(9)
    B, H1, W1, H2, W2, r = 2, 5, 6, 9, 12, 3
    vol = rng.normal(size=(B, H1, W1, H2, W2))
    c = rng.uniform(-2, 13, size=(B, 2, H1, W1)).astype(np.float32)
    out = O.corr_index_forward(vol, c, r)
    planes = torch.from_numpy(vol).view(B * H1 * W1, 1, H2, W2)
    xy = torch.from_numpy(c).double().permute(0, 2, 3, 1).reshape(-1, 1, 1, 2)
    d = torch.arange(-r, r + 1, dtype=torch.float64)
    g = xy + torch.stack(torch.meshgrid(d, d, indexing="ij"), -1).view(1, 7, 7, 2)
    g = torch.stack([2 * g[..., 0] / (W2 - 1) - 1, 2 * g[..., 1] / (H2 - 1) - 1], -1)
    s = F.grid_sample(planes, g, padding_mode="zeros", align_corners=True).view(B, H1, W1, 7, 7)
    assert np.abs(s.permute(0, 3, 4, 1, 2).numpy() - out).max() < 1e-6


def test_corr_index_backward_is_adjoint():
    rng = np.random.default_rng(1)
    vol = rng.normal(size=(1, 3, 4, 7, 9))
    c = rng.uniform(-2, 10, size=(1, 2, 3, 4)).astype(np.float32)
    g = rng.normal(size=(1, 7, 7, 3, 4))
    lhs = np.sum(O.corr_index_forward(vol, c, 3) * g)
    rhs = np.sum(vol * O.corr_index_backward(vol.shape, np.float64, c, g, 3))
    assert abs(lhs - rhs) < 1e-9 * max(1.0, abs(lhs))


def test_corr_build_pyramid_shapes_and_pooling():
    p = synth.make_corr_problem(2, 8, 16, 32, seed=1)
    f = p["fmaps"]
    lv = O.corr_build(f[p["ii"]], f[p["jj"]])
    assert [l.shape for l in lv] == [(2, 8, 16, 8 >> k, 16 >> k) for k in range(4)]
    a = f[p["ii"]][0].astype(np.float32).reshape(32, -1) / 4; b = f[p["jj"]][0].astype(np.float32).reshape(32, -1) / 4
    assert rel(lv[0][0].reshape(128, 128), a.T @ b) < 2e-3
    assert rel(lv[1][0, 3, 5, 1, 2], lv[0][0, 3, 5, 2:4, 4:6].astype(np.float32).mean()) < 2e-3


def test_ba_converges_and_fixes_gauge():
    p = synth.make_ba_problem("C1")
    def err(poses, disps):
        c, _ = O.projmap(poses, disps, p["intrinsics"], p["ii"], p["jj"])
        return float(np.abs(c[..., :2].transpose(0, 3, 1, 2) - p["targets"]).mean())
    poses, disps = p["poses"].copy(), p["disps"].copy()
    e0 = err(poses, disps)
    O.ba(poses, disps, p["intrinsics"], p["disps_sens"], p["targets"], p["weights"], p["eta"], p["ii"], p["jj"],
         p["t0"], p["t1"], 4, 1e-4, 0.1, False)
    assert err(poses, disps) < 0.4 * e0
    assert np.array_equal(poses[0], p["poses"][0])                     # poses < t0 stay fixed


def test_frame_distance_and_depth_filter_properties():
    q = G.geom_case()
    n = q["poses"].shape[0]
    same = np.arange(n, dtype=np.int64)
    assert np.all(O.frame_distance(q["poses"], q["disps"], q["intr"], same, same, 0.3) == 0)
    cnt = O.depth_filter(q["gt_poses"], q["gt_disps"], q["intr"], q["ix"], q["thresh"])
    assert cnt.min() >= 0 and cnt.max() <= 6 and cnt.max() >= 3 and np.all(cnt == np.round(cnt))


def test_retraction_small_and_large_angle():
    t, q = np.array([0.1, -0.2, 0.3], np.float32), np.array([0, 0, 0, 1], np.float32)
    for xi in (np.zeros(6, np.float32), np.array([1e-3, 0, 0, 1e-6, 0, 0], np.float32),
               np.array([0.1, 0.2, -0.1, 0.3, -0.2, 0.1], np.float32)):
        t1, q1 = O.retr_se3(xi, t, q)
        assert abs(np.linalg.norm(q1) - 1) < 1e-6
        R, tt = synth.se3_exp(xi.astype(np.float64))
        assert np.abs(t1 - (R @ t + tt)).max() < 1e-6


@pytest.mark.parametrize("key", sorted(__import__("proximity_cases").CASES))
def test_proximity_edges_vs_reference(key):
    """the restatement of FactorGraph.add_proximity_factors' edge selection against the edge lists the
    UNMODIFIED reference method produced on the same seeded inputs (tests/golden/make_golden_proximity.py): bit-exact,
    order included."""
    import proximity_cases as PC
    c = PC.make_case(key)
    ref = gold("ref_proximity")
    assert str(ref[key + "_digest"]) == PC.digest(c)
    ii, jj = O.proximity_edges(c["d"], c["t"], c["t0"], c["t1"], c["rad"], c["nms"], c["thresh"], c["max_factors"],
                               c["stereo"], c["ii_exist"], c["jj_exist"])
    assert np.array_equal(ii, ref[key + "_ii"]) and np.array_equal(jj, ref[key + "_jj"])


@pytest.mark.parametrize("key", sorted(G.FILTER_CASES))
def test_filter_repeated_edges_vs_reference(key):
    """SURVEY 8f N3, second half: the restatement of FactorGraph.__filter_repeated_edges against the output of the
    UNMODIFIED method (tests/golden/make_golden_pyref.py): bit-exact, order included."""
    c = G.filter_case(key)
    ref = gold("ref_filter_edges")
    assert str(ref[key + "_digest"]) == G.digest([c["ii"], c["jj"], c["ii_exist"], c["jj_exist"]])
    ii, jj = O.filter_repeated_edges(c["ii"], c["jj"], c["ii_exist"], c["jj_exist"])
    assert np.array_equal(ii, ref[key + "_ii"]) and np.array_equal(jj, ref[key + "_jj"])


@pytest.mark.parametrize("key", sorted(__import__("upsample_cases").CASES))
def test_cvx_upsample_vs_reference(key):
    """restatement of cvx_upsample against outputs of the UNMODIFIED reference function
    (droid_net.py:21-35, run on CPU by tests/golden/make_golden_upsample.py).  <= 1e-6 relative."""
    import upsample_cases as UC
    data, mask = UC.make_case(key)
    ref = gold("ref_cvx_upsample")
    assert str(ref[key + "_digest"]) == UC.digest(data, mask)
    assert rel(O.cvx_upsample(data, mask), ref[key]) <= 1e-6


def test_scatter_mean_restatement_against_torch_index_add():
    """torch_scatter is absent; its scatter_mean is restated (sum / clamp(count, 1)) and checked
    here against the same definition built from torch.index_add_ (the shim SURVEY 8c names)."""
    import torch
    rng = np.random.default_rng(4)
    src = rng.normal(size=(1, 23, 8, 3, 4)).astype(np.float32)
    idx = rng.integers(0, 6, size=23)
    idx[idx == 4] = 5                                               # an empty group in the middle
    out = O.scatter_mean(src, idx)
    s = torch.zeros((1, 6, 8, 3, 4), dtype=torch.float64).index_add_(1, torch.from_numpy(idx), torch.from_numpy(src).double())
    c = torch.zeros(6, dtype=torch.float64).index_add_(0, torch.from_numpy(idx), torch.ones(23, dtype=torch.float64))
    want = (s / c.clamp(min=1).view(1, 6, 1, 1, 1)).float().numpy()
    assert out.shape == (1, 6, 8, 3, 4) and rel(out, want) <= 1e-6 and np.all(out[0, 4] == 0)
