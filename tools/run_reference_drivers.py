"""The reference's own drivers of the hot path, run end to end on synthetic keyframes with three arms:

  reference   unmodified reference Python (oracle/_ref/pyref) + the unmodified reference extension (oracle/_ref)
  drop-in     the same unmodified Python with `droid_backends` bound to THIS repo's module (nothing else changes)
  integrated  drop-in + the bindings INTEGRATION.md proposes for the rows of SURVEY 8f (AltCorrBlock.corr_fn ->
              altcorr_pyramid_indexed, the edge selection of add_proximity_factors -> proximity_edges,
              __filter_repeated_edges -> filter_repeated_edges, DepthVideo.reproject -> reproject,
              CorrBlock.__init__ / __call__ -> corr_build / corr_index_pyramid)

backend:  droid_slam/droid_backend.py DroidBackend.__call__ (BASELINE configs[3]): FactorGraph(corr_impl="alt"),
          add_proximity_factors over all keyframe pairs, update_lowmem(steps): per step reproject, ~n/8 chunks of
          AltCorrBlock lookups + the update operator, one global ba(itrs=2, lm=1e-5, ep=1e-2).
frontend: droid_slam/droid_frontend.py DroidFrontend.__call__ (BASELINE configs[2] / the frontend half of configs[4]):
          initialisation on 8 keyframes, then per keyframe add_proximity_factors + 4..6 FactorGraph.update
          (CorrBlock lookup, update operator, ba on the window).

The network is a random-init DroidNet (PyTorch / cuDNN in every arm: north_star leaves it there); keyframes come from
synth.make_ba_problem (consistent poses / depth maps + noise) with random feature maps.  Every droid_backends call is
timed with CUDA events on the current stream, so the JSON shows where the time of each arm goes.

Usage (GPU box; TEST / MEASUREMENT TOOL, imports oracle/pyref like tests/test_reference_callers.py does):
  python tools/run_reference_drivers.py backend --frames 500 --steps 2 [--arms reference,drop-in,integrated]
  python tools/run_reference_drivers.py frontend --frames 60
"""
import argparse
import contextlib
import importlib
import io
import json
import os
import sys
import time
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("droid-slam_b200", "oracle", os.path.join("oracle", "_ref")):
    sys.path.insert(0, os.path.join(ROOT, p))

import droid_backends as B  # noqa: E402
import pyref  # noqa: E402
import synth  # noqa: E402

DEV = "cuda:0"


class Timed:
    """A `droid_backends` look-alike that forwards to `mod` and times every call with CUDA events."""

    def __init__(self, mod):
        self._mod = mod
        self._ev = {}

    def __getattr__(self, name):
        fn = getattr(self._mod, name)
        if not callable(fn):
            return fn

        def call(*a, **k):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            out = fn(*a, **k)
            e1.record()
            self._ev.setdefault(name, []).append((e0, e1))
            return out
        return call

    def report(self):
        torch.cuda.synchronize()
        out = {k: {"calls": len(v), "ms": round(sum(a.elapsed_time(b) for a, b in v), 3)} for k, v in self._ev.items()}
        self._ev = {}
        return out


def fill_video(ns, p, n, buffer, seed=0, stereo=False):
    ht, wd = p["ht"], p["wd"]
    v = ns.depth_video.DepthVideo(image_size=[8 * ht, 8 * wd], buffer=buffer, stereo=stereo)
    v.poses[:n] = torch.from_numpy(p["poses"][:n]).to(DEV)
    v.disps[:n] = torch.from_numpy(p["disps"][:n]).to(DEV)
    v.intrinsics[:] = torch.from_numpy(p["intrinsics"]).to(DEV)
    v.tstamp[:n] = torch.arange(n, device=DEV, dtype=torch.float)
    g = torch.Generator(device=DEV).manual_seed(seed)
    v.fmaps[:n] = torch.randn((n, 2 if stereo else 1, 128, ht, wd), generator=g, device=DEV).half()
    v.nets[:n] = torch.tanh(torch.randn((n, 128, ht, wd), generator=g, device=DEV)).half()
    v.inps[:n] = torch.relu(torch.randn((n, 128, ht, wd), generator=g, device=DEV)).half()
    v.counter.value = n
    return v


def integrate(ns, Bm):
    """The INTEGRATION.md bindings, applied to the loaded reference classes; returns an undo function."""
    corr, fg = ns.corr, ns.factor_graph
    old_corr_fn, old_prox = corr.AltCorrBlock.corr_fn, fg.FactorGraph.add_proximity_factors

    def corr_fn(self, coords, ii, jj):                      # modules/corr.py:106-125
        Bb, N, H, W, S, _ = coords.shape
        assert Bb == 1
        c = coords[0].permute(0, 3, 1, 2, 4).contiguous()   # [N, S, H, W, 2]
        out = Bm.altcorr_pyramid_indexed([p[0] for p in self.pyramid], ii, jj, c, self.radius)   # [N, S, L*49, H, W]
        return out.permute(0, 2, 3, 4, 1)[None]             # [1, N, L*49, H, W, S]

    def add_proximity_factors(self, t0=0, t1=0, rad=2, nms=2, beta=0.25, thresh=16.0, remove=False):   # factor_graph.py:305-369
        t = self.video.counter.value
        ix, jx = torch.arange(t0, t), torch.arange(t1, t)
        ii, jj = torch.meshgrid(ix, jx, indexing="ij")
        d = self.video.distance(ii.reshape(-1), jj.reshape(-1), beta=beta)
        ii, jj = Bm.proximity_edges(d, t, t0, t1, rad, nms, thresh, self.max_factors, self.video.stereo,
                                    torch.cat([self.ii, self.ii_bad, self.ii_inac]), torch.cat([self.jj, self.jj_bad, self.jj_inac]))
        self.add_factors(ii, jj, remove)

    def filter_repeated_edges(self, ii, jj):                # factor_graph.py:43-54
        return Bm.filter_repeated_edges(ii.contiguous(), jj.contiguous(), torch.cat([self.ii, self.ii_inac]),
                                        torch.cat([self.jj, self.jj_inac]))

    old_cb_init, old_cb_call = corr.CorrBlock.__init__, corr.CorrBlock.__call__

    def cb_init(self, fmap1, fmap2, num_levels=4, radius=3):   # modules/corr.py:24-38,63-71: matmul + 3 avg_pool2d
        H, W = fmap1.shape[-2:]
        if num_levels != 4 or fmap1.dtype != torch.float16 or H % 8 or W % 16 or W > 64:
            return old_cb_init(self, fmap1, fmap2, num_levels, radius)
        self.num_levels, self.radius = num_levels, radius
        self.corr_pyramid = Bm.corr_build(fmap1.contiguous(), fmap2.contiguous())

    def cb_call(self, coords):                              # modules/corr.py:40-50: 4 lookups + cat
        batch, num, ht, wd, _ = coords.shape
        c = coords.permute(0, 1, 4, 2, 3).contiguous().view(batch * num, 2, ht, wd)
        vols = [v.contiguous() for v in self.corr_pyramid]
        return Bm.corr_index_pyramid(vols, c, self.radius).view(batch, num, -1, ht, wd)

    corr.CorrBlock.__init__, corr.CorrBlock.__call__ = cb_init, cb_call
    dv = ns.depth_video
    old_reproject = dv.DepthVideo.reproject

    def reproject(self, ii, jj):                            # depth_video.py:139-147
        ii, jj = dv.DepthVideo.format_indicies(ii, jj)
        coords, valid = Bm.reproject(self.poses, self.disps, self.intrinsics, ii.contiguous(), jj.contiguous())
        return coords[None], valid[None]

    dv.DepthVideo.reproject = reproject
    old_filter = fg.FactorGraph._FactorGraph__filter_repeated_edges
    corr.AltCorrBlock.corr_fn = corr_fn
    fg.FactorGraph.add_proximity_factors = add_proximity_factors
    fg.FactorGraph._FactorGraph__filter_repeated_edges = filter_repeated_edges

    def undo():
        corr.AltCorrBlock.corr_fn = old_corr_fn
        fg.FactorGraph.add_proximity_factors = old_prox
        fg.FactorGraph._FactorGraph__filter_repeated_edges = old_filter
        dv.DepthVideo.reproject = old_reproject
        corr.CorrBlock.__init__, corr.CorrBlock.__call__ = old_cb_init, old_cb_call
    return undo


def run_arm(kind, arm, ns, net, p, args, ref_mod):
    mod = Timed(ref_mod if arm == "reference" else B)
    pyref.use(ns, mod)
    undo = integrate(ns, mod) if arm == "integrated" else (lambda: None)
    try:
        torch.manual_seed(1)
        quiet = io.StringIO()
        phases = {}

        def timed(name, fn):
            torch.cuda.synchronize()
            t = time.perf_counter()
            with contextlib.redirect_stdout(quiet):
                r = fn()
            torch.cuda.synchronize()
            phases[name] = round(time.perf_counter() - t, 4)
            return r
        if kind == "backend":
            n = args.frames
            v = fill_video(ns, p, n, n + 8, stereo=args.stereo)
            a = types.SimpleNamespace(beta=0.3, backend_thresh=22.0, backend_radius=2, backend_nms=3)
            with torch.no_grad():                          # the body of DroidBackend.__call__ (droid_backend.py:24-40), phase by phase
                if not v.stereo:                            # droid_backend.py:28-29
                    timed("normalize", v.normalize)
                g = timed("graph", lambda: ns.factor_graph.FactorGraph(v, net.update, corr_impl="alt", max_factors=16 * n))
                timed("add_proximity_factors", lambda: g.add_proximity_factors(rad=a.backend_radius, nms=a.backend_nms,
                                                                              thresh=a.backend_thresh, beta=a.beta))
                edges = int(g.ii.numel())
                timed("update_lowmem", lambda: g.update_lowmem(steps=args.steps))
            out = {"edges": edges, "keyframes": n, "steps": args.steps}
        else:
            n = args.frames
            fe_mod = importlib.import_module("droid_frontend")
            a = types.SimpleNamespace(warmup=8, beta=0.3, frontend_nms=1, keyframe_thresh=0.0, frontend_window=25,
                                      frontend_thresh=16.0, frontend_radius=2)
            full = fill_video(ns, p, n, n + 8, stereo=args.stereo)
            v = ns.depth_video.DepthVideo(image_size=[8 * p["ht"], 8 * p["wd"]], buffer=n + 8, stereo=args.stereo)
            v.intrinsics[:] = full.intrinsics
            fe = fe_mod.DroidFrontend(net, v, a)

            def stream():
                for k in range(n):                        # what MotionFilter.track appends for a new keyframe
                    v.tstamp[k] = float(k)
                    v.fmaps[k], v.nets[k], v.inps[k] = full.fmaps[k], full.nets[k], full.inps[k]
                    if k < 8:                             # the initial window starts from the synthetic state
                        v.poses[k], v.disps[k] = full.poses[k], full.disps[k]
                    v.counter.value = k + 1
                    fe()
            with torch.no_grad():
                timed("frontend", stream)
            out = {"keyframes": n, "keyframes_kept": int(v.counter.value), "edges_at_end": int(fe.graph.ii.numel()),
                   "keyframes_per_s": round((n - 8) / phases["frontend"], 2)}
        out["seconds"] = phases
        out["total_s"] = round(sum(phases.values()), 4)
        out["droid_backends_calls"] = mod.report()
        out["droid_backends_ms"] = round(sum(x["ms"] for x in out["droid_backends_calls"].values()), 3)
        out["finite"] = bool(torch.isfinite(v.poses[:n]).all() and torch.isfinite(v.disps[:n]).all())
        out["_state"] = (v.poses[:n].clone(), v.disps[:n].clone())
        return out
    finally:
        undo()
        pyref.use(ns, B)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("kind", choices=("backend", "frontend"))
    ap.add_argument("--frames", type=int, default=500)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--arms", default="reference,drop-in,integrated")
    ap.add_argument("--stereo", action="store_true", help="rig of two cameras (BASELINE configs[4]): fmaps of both, (i, i) edges")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    ok, why = pyref.available()
    if not ok:
        raise SystemExit(why)
    import droid_backends_ref as R
    torch.backends.cudnn.benchmark = False
    ns = pyref.load(B)
    torch.manual_seed(0)
    net = ns.droid_net.DroidNet().to(DEV).eval()
    p = synth.make_ba_problem("C4", n_frames=max(args.frames, 16), knn=16)      # EuRoC shape: 40x64 at 1/8
    res = {"kind": args.kind, "frames": args.frames, "stereo": bool(args.stereo), "image": "%dx%d at 1/8" % (p["ht"], p["wd"]), "arms": {}}
    for arm in args.arms.split(","):
        run_arm(args.kind, arm, ns, net, p, types.SimpleNamespace(frames=min(args.frames, 40), steps=1, stereo=args.stereo), R)   # warm-up (cuDNN, allocator)
        res["arms"][arm] = run_arm(args.kind, arm, ns, net, p, args, R)
    states = {arm: r.pop("_state") for arm, r in res["arms"].items()}
    ref = res["arms"].get("reference")
    rel = lambda a, b: float((a - b).abs().max() / b.abs().max().clamp(min=1e-30))   # noqa: E731
    for arm, r in res["arms"].items():
        if ref is not None and arm != "reference":
            r["vs_reference"] = {"poses": rel(states[arm][0], states["reference"][0]),
                                 "disps": rel(states[arm][1], states["reference"][1]),
                                 "speedup_total": round(ref["total_s"] / r["total_s"], 2),
                                 "speedup_droid_backends": round(ref["droid_backends_ms"] / max(r["droid_backends_ms"], 1e-9), 2)}
    s = json.dumps(res, indent=1)
    print(s)
    if args.out:
        open(args.out, "w").write(s + "\n")


if __name__ == "__main__":
    main()
