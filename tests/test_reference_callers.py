"""The drop-in claim, checked through the reference's OWN Python: the unmodified droid_slam/modules/corr.py
(CorrBlock, AltCorrBlock), depth_video.py (DepthVideo.distance / reproject / ba) and factor_graph.py (FactorGraph with a
random-init DroidNet update operator) are imported from the staged copy (oracle/_ref/pyref, oracle/stage_pyref.sh,
loaded by oracle/pyref.py) and run twice -- once with `droid_backends` bound to this repo's module, once bound to the
reference extension (oracle/_ref/droid_backends_ref) -- on identical seeded inputs.  Same Python, two back ends, same
numbers.  Reference: /root/reference/droid_slam/modules/corr.py:24-139, depth_video.py:139-193, factor_graph.py:84-237,
305-369.
"""
import numpy as np
import pytest
import torch

import synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def rel(a, b):
    a = a.double(); b = b.double()
    return float((a - b).abs().max() / b.abs().max().clamp(min=1e-30))


@pytest.fixture(scope="module")
def ref_py(backends, ref_ext):
    import pyref
    ok, why = pyref.available()
    if not ok:
        pytest.fail(why)
    torch.backends.cudnn.benchmark = False
    torch.backends.cudnn.deterministic = True
    return pyref.load(backends)


def both(ns, backends, ref_ext, fn):
    import pyref
    pyref.use(ns, backends)
    a = fn()
    pyref.use(ns, ref_ext)
    b = fn()
    pyref.use(ns, backends)
    return a, b


def test_corr_block_through_reference_python(backends, ref_ext, ref_py):
    """modules/corr.py CorrBlock: torch matmul + avg_pool pyramid, then droid_backends.corr_index_forward per level
    (float16 as in inference, float32 as in training); index-select / cat of the block included."""
    q = synth.make_corr_problem(12, 48, 64, 128, seed=3)
    for dt in (torch.float16, torch.float32):
        f1 = torch.from_numpy(q["fmaps"][q["ii"]]).to(DEV).to(dt)[None]
        f2 = torch.from_numpy(q["fmaps"][q["jj"]]).to(DEV).to(dt)[None]
        coords = torch.from_numpy(q["coords"]).to(DEV).permute(0, 2, 3, 1)[None].contiguous()

        def run():
            blk = ref_py.corr.CorrBlock(f1, f2)
            out = blk(coords)
            sub = blk[torch.arange(12, device=DEV) % 2 == 0]
            return out, sub(coords[:, ::2])
        (a, a2), (b, b2) = both(ref_py, backends, ref_ext, run)
        assert a.shape == (1, 12, 4 * 49, 48, 64) and a.dtype == dt
        assert torch.equal(a, b) and torch.equal(a2, b2)


def test_alt_corr_block_through_reference_python(backends, ref_ext, ref_py):
    """modules/corr.py AltCorrBlock (the backend's update_lowmem path): pyramid of pooled feature maps, per level a
    gather + .float() + droid_backends.altcorr_forward through the CorrLayer autograd function; backward too."""
    q = synth.make_corr_problem(10, 40, 64, 128, seed=8, oob_frac=0.02)
    fmaps = torch.from_numpy(q["fmaps"]).to(DEV)[None]                       # [1, frames, 128, H, W] half
    ii, jj = torch.from_numpy(q["ii"]).to(DEV), torch.from_numpy(q["jj"]).to(DEV)
    coords = torch.from_numpy(q["coords"]).to(DEV).permute(0, 2, 3, 1)[None].contiguous()   # [1, N, H, W, 2]

    def run():
        blk = ref_py.corr.AltCorrBlock(fmaps)
        return blk(coords, ii, jj)
    a, b = both(ref_py, backends, ref_ext, run)
    assert a.shape == (1, 10, 4 * 49, 40, 64)
    assert rel(a, b) <= 1e-5

    # the one-launch binding INTEGRATION.md proposes for corr_fn: the reference's own pyramid, every level at once
    import pyref
    pyref.use(ref_py, ref_ext)
    blk = ref_py.corr.AltCorrBlock(fmaps)
    want = blk(coords, ii, jj)                                                # unmodified reference, reference extension
    pyref.use(ref_py, backends)
    got = backends.altcorr_pyramid_indexed([p[0] for p in blk.pyramid], ii, jj, coords[0][:, None].contiguous(), 3)
    assert got.shape == (10, 1, 4 * 49, 40, 64)
    assert rel(got[:, 0], want[0]) <= 1e-5

    f1 = (fmaps[0, ii].float() / 4).permute(0, 2, 3, 1).contiguous().requires_grad_(True)
    f2 = (fmaps[0, jj].float() / 4).permute(0, 2, 3, 1).contiguous().requires_grad_(True)
    c5 = coords[0][:, None].contiguous()
    g = torch.randn((10, 1, 49, 40, 64), device=DEV, generator=torch.Generator(device=DEV).manual_seed(1))

    def grads():
        out = ref_py.corr.CorrLayer.apply(f1, f2, c5, 3)
        return torch.autograd.grad(out, (f1, f2), g)
    (a1, a2), (b1, b2) = both(ref_py, backends, ref_ext, grads)
    assert rel(a1, b1) <= 1e-5 and rel(a2, b2) <= 1e-5


def make_video(ns, p, buffer=None, fmaps=False, seed=0):
    """A reference DepthVideo filled with the synthetic problem `p` (synth.make_ba_problem)."""
    n = p["n"]
    ht, wd = p["ht"], p["wd"]
    v = ns.depth_video.DepthVideo(image_size=[8 * ht, 8 * wd], buffer=buffer or p["poses"].shape[0])
    v.poses[:p["poses"].shape[0]] = torch.from_numpy(p["poses"]).to(DEV)
    v.disps[:p["disps"].shape[0]] = torch.from_numpy(p["disps"]).to(DEV)
    v.intrinsics[:] = torch.from_numpy(p["intrinsics"]).to(DEV)
    v.counter.value = n
    if fmaps:
        g = torch.Generator(device="cpu").manual_seed(seed)
        v.fmaps[:n] = torch.randn((n, 1, 128, ht, wd), generator=g).to(DEV).half()
        v.nets[:n] = torch.tanh(torch.randn((n, 128, ht, wd), generator=g)).to(DEV).half()
        v.inps[:n] = torch.relu(torch.randn((n, 128, ht, wd), generator=g)).to(DEV).half()
    return v


def test_depth_video_through_reference_python(backends, ref_ext, ref_py):
    """depth_video.py DepthVideo.distance (frame_distance both directions, pair grid), .reproject (the lietorch path;
    also the reference-run check of this repo's fused `reproject` extension) and .ba (in place, lock,
    clamp) on the frontend window C3 and a 60-keyframe backend graph."""
    for name, kw, lm, ep in (("C3", {}, 1e-4, 0.1), ("C4", dict(n_frames=60), 1e-5, 1e-2)):
        p = synth.make_ba_problem(name, **kw)
        ii, jj = torch.from_numpy(p["ii"]).to(DEV), torch.from_numpy(p["jj"]).to(DEV)
        tgt = torch.from_numpy(p["targets"]).to(DEV)
        wgt = torch.from_numpy(p["weights"]).to(DEV)
        eta = torch.from_numpy(p["eta"]).to(DEV)
        vids = {}

        def run(tag):
            def f():
                v = make_video(ref_py, p)
                d_all = v.distance(beta=0.3)                                  # N x N matrix, bidirectional
                d_e = v.distance(ii, jj, beta=0.4, bidirectional=False)
                v.ba(tgt, wgt, eta, ii, jj, t0=p["t0"], t1=p["t1"], itrs=2, lm=lm, ep=ep, motion_only=False)
                vids[tag] = v
                return d_all, d_e, v.poses.clone(), v.disps.clone()
            return f
        import pyref
        pyref.use(ref_py, backends); a = run("mine")()
        pyref.use(ref_py, ref_ext); b = run("ref")()
        pyref.use(ref_py, backends)
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])           # frame_distance: bit-exact
        assert rel(a[2], b[2]) <= 1e-5 and rel(a[3], b[3]) <= 1e-5            # state after ba(itrs=2)
        assert float(a[3].min()) >= 0.001                                     # the caller's clamp ran
        # reproject: the reference's lietorch path vs this repo's one-launch extension, on the reference video's state
        v = vids["ref"]
        c_ref, m_ref = v.reproject(ii, jj)
        c, m = backends.reproject(v.poses, v.disps, v.intrinsics, ii, jj)
        scale = float(c_ref.abs().max())
        assert float((c - c_ref[0]).abs().max()) <= 1e-5 * scale
        assert float((m != m_ref[0]).float().mean()) <= 1e-4


def test_factor_graph_update_through_reference_python(backends, ref_ext, ref_py):
    """factor_graph.py FactorGraph with a random-init DroidNet update operator (the network itself stays PyTorch / cuDNN):
    add_neighborhood_factors -> CorrBlock build + reproject; update() x2 -> corr lookup, GRU, droid_backends.ba;
    add_proximity_factors -> DepthVideo.distance + the Python edge selection.  Both back ends must leave the same
    edges, poses and depth maps; this repo's device-side edge selection must pick the same proximity edges."""
    p = synth.make_ba_problem("C3")
    torch.manual_seed(0)
    net = ref_py.droid_net.DroidNet().to(DEV).eval()
    import pyref

    def run(backend):
        pyref.use(ref_py, backend)
        torch.manual_seed(1)
        v = make_video(ref_py, p, buffer=16, fmaps=True)
        g = ref_py.factor_graph.FactorGraph(v, net.update, device=DEV, corr_impl="volume", max_factors=48)
        with torch.no_grad():
            g.add_neighborhood_factors(0, p["n"], r=2)
            for _ in range(2):
                g.update(t0=1, t1=None, itrs=2, use_inactive=False)
            before = (g.ii.clone(), g.jj.clone())
            g.add_proximity_factors(t0=0, t1=0, rad=2, nms=2, beta=0.3, thresh=100.0, remove=False)
            g.update(t0=1, t1=None, itrs=2, use_inactive=False)
        return g, v, before
    ga, va, ba_ = run(backends)
    gb, vb, bb = run(ref_ext)
    pyref.use(ref_py, backends)
    assert torch.equal(ga.ii, gb.ii) and torch.equal(ga.jj, gb.jj) and ga.ii.numel() > ba_[0].numel()
    assert torch.isfinite(va.poses).all() and torch.isfinite(va.disps).all()
    ep, ed = rel(va.poses[:p["n"]], vb.poses[:p["n"]]), rel(va.disps[:p["n"]], vb.disps[:p["n"]])
    print("FactorGraph: poses rel %.2e, disps rel %.2e after 3 updates (6 BA iterations), %d edges" % (ep, ed, ga.ii.numel()))
    assert ep <= 1e-4 and ed <= 1e-4
    # the same proximity selection on the device, fed with the reference video's own distances
    t = p["n"]
    ix = torch.arange(0, t)
    ii, jj = torch.meshgrid(ix, ix, indexing="ij")
    d = vb.distance(ii.reshape(-1), jj.reshape(-1), beta=0.3)
    si, sj = backends.proximity_edges(d, t, 0, 0, 2, 2, 100.0, 48, False, bb[0], bb[1])
    new = backends.filter_repeated_edges(si, sj, bb[0], bb[1])
    want_i, want_j = gb.ii[bb[0].numel():], gb.jj[bb[0].numel():]
    assert torch.equal(new[0], want_i) and torch.equal(new[1], want_j)
