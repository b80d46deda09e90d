"""BASELINE configs[3] as ONE multi-GPU job: the backend's global update on N GPUs of one box.

Every rank holds the same synthetic keyframe video and the same factor graph (built by the reference's own
FactorGraph.add_proximity_factors, with the INTEGRATION.md bindings of tools/run_reference_drivers.py).  The body of
FactorGraph.update_lowmem (droid_slam/factor_graph.py:246-290) is then run SHARDED BY SOURCE FRAME: rank r
takes the source frames [bounds[r], bounds[r+1]) -- the alt-corr lookups and the update operator of their edges, no
communication -- and the step ends with ONE droid_backends.ba_sharded call, which gets only the rank's own rows of
targets / weights, sums the packed reduced camera system with one NCCL all-reduce per Gauss-Newton iteration and leaves
identical poses / depth maps on every rank.  Reported: seconds per update_lowmem(steps) as the max over ranks, the split
alt-corr + update operator / ba, and the difference to the unsharded run of the same code on rank 0.

Usage (GPU box):  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
                      tools/run_backend_sharded.py --frames 1000 --steps 2 [--out file.json]
MEASUREMENT TOOL: imports oracle/pyref (the staged, unmodified reference Python) like tests/test_reference_callers.py.
"""
import argparse
import contextlib
import io
import json
import os
import sys
import time
import types

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import run_reference_drivers as D  # noqa: E402  (sets the import paths, provides fill_video / integrate)

B, pyref, synth = D.B, D.pyref, D.synth


def update_lowmem_sharded(ns, g, steps, bounds, rank, itrs=2, EP=1e-7):
    """factor_graph.py:246-290 for the source frames of one rank; returns (seconds in lookups + update operator, in ba)."""
    v = g.video
    t = v.counter.value
    num, rig, ch, ht, wd = v.fmaps.shape
    corr_op = ns.corr.AltCorrBlock(v.fmaps.view(1, num * rig, ch, ht, wd))
    lo, hi = int(bounds[rank]), int(bounds[rank + 1])
    mine = (g.ii >= lo) & (g.ii < hi)
    t_net = t_ba = 0.0
    for _ in range(steps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        with torch.cuda.amp.autocast(enabled=False):
            coords1, mask = v.reproject(g.ii, g.jj)
            motn = torch.cat([coords1 - g.coords0, g.target - coords1], dim=-1)
            motn = motn.permute(0, 1, 4, 2, 3).clamp(-64.0, 64.0)
        s = 8
        for i in range(lo, hi, s):
            sel = (g.ii >= i) & (g.ii < min(i + s, hi))
            if not bool(sel.any()):
                continue
            iis, jjs = g.ii[sel], g.jj[sel]
            corr1 = corr_op(coords1[:, sel], rig * iis, rig * jjs + (iis == jjs).long())
            with torch.cuda.amp.autocast(enabled=True):
                net, delta, weight, damping, _ = g.update_op(g.net[:, sel], v.inps[None, iis], corr1, motn[:, sel], iis, jjs)
            g.net[:, sel] = net
            g.target[:, sel] = coords1[:, sel] + delta.float()
            g.weight[:, sel] = weight.float()
            g.damping[torch.unique(iis)] = damping
        damping = .2 * g.damping[torch.unique(g.ii)].contiguous() + EP          # rows of other ranks' frames are never read
        h, w = g.coords0.shape[0:2]
        target = g.target[:, mine].view(-1, h, w, 2).permute(0, 3, 1, 2).contiguous()   # this rank's rows only
        weight = g.weight[:, mine].view(-1, h, w, 2).permute(0, 3, 1, 2).contiguous()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        with v.get_lock():                                                       # depth_video.py:178-193
            B.ba_sharded(v.poses, v.disps, v.intrinsics[0], v.disps_sens, target, weight, damping, g.ii, g.jj, 1, t,
                         itrs, 1e-5, 1e-2, False, bounds=bounds)
            v.disps.clamp_(min=0.001)
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        t_net += t1 - t0
        t_ba += t2 - t1
    return t_net, t_ba


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=1000)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    D.DEV = "cuda:%d" % local
    dist.init_process_group("nccl", device_id=torch.device(D.DEV))
    ok, why = pyref.available()
    if not ok:
        raise SystemExit(why)
    torch.backends.cudnn.benchmark = False
    ns = pyref.load(B)
    pyref.use(ns, B)
    undo = D.integrate(ns, B)
    torch.manual_seed(0)
    net = ns.droid_net.DroidNet().to(D.DEV).eval()
    n = args.frames
    p = synth.make_ba_problem("C4", n_frames=max(n, 16), knn=16)
    quiet = io.StringIO()

    def build():
        torch.manual_seed(1)
        v = D.fill_video(ns, p, n, n + 8)
        with torch.no_grad(), contextlib.redirect_stdout(quiet):
            v.normalize()
            g = ns.factor_graph.FactorGraph(v, net.update, device=D.DEV, corr_impl="alt", max_factors=16 * n)
            g.add_proximity_factors(rad=2, nms=3, thresh=22.0, beta=0.3)
        return v, g

    res = {"kind": "backend, sharded by source frame", "world": world, "frames": n, "steps": args.steps}
    with torch.no_grad():
        # warm-up (cuDNN, allocator, NCCL) on a small graph
        pw = synth.make_ba_problem("C4", n_frames=48, knn=16)
        vw = D.fill_video(ns, pw, 48, 56)
        with contextlib.redirect_stdout(quiet):
            gw = ns.factor_graph.FactorGraph(vw, net.update, device=D.DEV, corr_impl="alt", max_factors=16 * 48)
            gw.add_proximity_factors(rad=2, nms=3, thresh=22.0, beta=0.3)
        update_lowmem_sharded(ns, gw, 1, B.shard_bounds(gw.ii, vw.disps.shape[0], world), rank)
        del vw, gw
        v, g = build()
        bounds = B.shard_bounds(g.ii, v.disps.shape[0], world)
        res["edges"] = int(g.ii.numel())
        res["edges_of_rank"] = int(((g.ii >= bounds[rank]) & (g.ii < bounds[rank + 1])).sum())
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        t_net, t_ba = update_lowmem_sharded(ns, g, args.steps, bounds, rank)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0, t_net, t_ba], device=D.DEV, dtype=torch.float64)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        res["seconds"] = {"update_lowmem": round(float(dt[0]), 4), "lookups_and_update_operator": round(float(dt[1]), 4),
                          "ba_sharded": round(float(dt[2]), 4)}
        # every rank must hold the same state
        ref_p, ref_d = v.poses[:n].clone(), v.disps[:n].clone()
        dist.broadcast(ref_p, 0)
        dist.broadcast(ref_d, 0)
        same = torch.tensor([float(torch.equal(ref_p, v.poses[:n]) and torch.equal(ref_d, v.disps[:n]))], device=D.DEV)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        res["ranks_hold_identical_state"] = bool(same.item() == 1.0)
        if rank == 0:                                  # the unsharded run of the same (integrated) code, for comparison
            v1, g1 = build()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            with contextlib.redirect_stdout(quiet):
                g1.update_lowmem(steps=args.steps)
            torch.cuda.synchronize()
            one = time.perf_counter() - t0
            rel = lambda a, b: float((a - b).abs().max() / b.abs().max().clamp(min=1e-30))   # noqa: E731
            res["single_gpu_seconds"] = round(one, 4)
            res["speedup"] = round(one / res["seconds"]["update_lowmem"], 2)
            res["vs_single_gpu"] = {"poses": rel(v.poses[:n], v1.poses[:n]), "disps": rel(v.disps[:n], v1.disps[:n])}
            res["finite"] = bool(torch.isfinite(v.poses[:n]).all() and torch.isfinite(v.disps[:n]).all())
    dist.barrier()
    undo()
    if rank == 0:
        s = json.dumps(res, indent=1)
        print(s)
        if args.out:
            open(args.out, "w").write(s + "\n")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
