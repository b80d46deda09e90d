#!/usr/bin/env python
"""bench.py -- global bundle-adjustment iterations/s at 1000 keyframes (BASELINE.json metric #1, configs[3]).

One "step" = one `droid_backends.ba(..., iterations=2, lm=1e-5, ep=1e-2)` call of the backend
(droid_slam/factor_graph.py:285-288) on the synthetic C4 problem of droid-slam_b200/synth.py
(1000 keyframes, 20 862 edges, 40x64 inverse-depth maps, EuRoC intrinsics) = 2 Gauss-Newton iterations.

  python bench.py [--gpus N] [--steps K] [--warmup W]        this repo's sm_100a path
  python bench.py --impl reference ...                        the UNMODIFIED reference extension (oracle/_ref)

Output: ONE JSON line on rank 0 (see the keys at the bottom).  Timing: CUDA events on the launching stream,
barrier + synchronize on both sides, max over ranks.  The per-step inputs (targets + weights = 855 MB) are far
larger than the 126 MB L2, so no explicit L2 flush is needed ("l2": "inputs>L2").
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (os.path.join(ROOT, "droid-slam_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "oracle", "_ref")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

LM, EP, ITERS = 1e-5, 1e-2, 2          # droid_slam/factor_graph.py:285-288 (update_lowmem -> video.ba)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=1000, help="keyframes of the C4 problem (default = BASELINE 1000)")
    ap.add_argument("--knn", type=int, default=16, help="proximity edges per keyframe (16 -> ~20k edges)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-frames", type=int, default=48, help="keyframes of the bounded CPU-baseline sample")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the live comparison with the reference extension")
    ap.add_argument("--no-extra", action="store_true", help="skip the corr-lookup / frontend-window side metrics")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
# clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc, self.thr = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
            return
        self.thr = threading.Thread(target=self._read, daemon=True)
        self.thr.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t_begin, t_end):
        if self.proc is None:
            return None
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            if ts < t_begin or ts > t_end + 0.1:
                continue
            f = [x.strip() for x in line.split(",")]
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); power.append(float(f[2]))
            except Exception:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": float(max(power))}


# ------------------------------------------------------------------------------------------------
def setup_dist(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # stdout carries exactly one JSON line: NCCL's own messages (this image prints "NCCL version ..." there) go to stderr
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        torch.cuda.set_device(0)
        local = 0
    return world, rank, local


def barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x, world, dev):
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(x, world, dev):
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def problem(frames, knn):
    import synth
    return synth.make_ba_problem("C4", n_frames=frames, knn=knn)


KEYS = ("poses", "disps", "intrinsics", "disps_sens", "targets", "weights", "eta", "ii", "jj")


def algorithmic(p):
    """Algorithmic (compulsory) bytes / flops per Gauss-Newton iteration, SURVEY.md 8(d) / DESIGN.md."""
    ne, hw = len(p["ii"]), p["ht"] * p["wd"]
    M, P = p["eta"].shape[0], p["t1"] - p["t0"]
    deg = np.bincount(p["ii"], minlength=p["n"])
    rows = 6 * (deg[deg > 0] + 1)
    pe = ne * hw                                                                      # pixel-edges
    return {
        # targets+weights once, disps/disps_sens/eta in, Q/W out, 90 fp64 values per edge into the system
        "ba_accum_bytes": pe * 16 + M * hw * 20 + ne * 90 * 8,
        # + the packed E_j (5 floats per pixel-edge, ba_lin.cuh) that this pass leaves behind for the next two
        "ba_accum_streamed_bytes": pe * 16 + M * hw * 20 + ne * 90 * 8 + pe * 20,
        # packed E_j + Q/W in, lower triangle of S out
        "ba_schur_bytes": pe * 20 + M * hw * 8 + int((rows * (rows + 1) // 2).sum()) * 8,
        "ba_schur_flops": int((rows.astype(np.int64) * (rows + 1)).sum()) * hw,       # lower-triangular Gram
        # what the tensor-core kernel issues: three 128x128 tf32 MMAs per 8 pixels and CTA; one CTA for frames with
        # <= 21 edges, three (one per pair of edge groups) for 22..31 edges
        "ba_schur_issued_flops": int(((deg > 0) & (deg <= 21)).sum() + 3 * ((deg > 21) & (deg <= 31)).sum()) * 3 * 2 * 128 * 128 * hw,
        "ba_solve_flops": (6 * P) ** 3 // 3 + 2 * (6 * P) ** 2,                        # dense fp64 Cholesky + 2 solves
        "ba_solve_bytes": (6 * P) ** 2 * 8 // 2,
        # packed E_j + Q/W + depth map in, dz + depth map out
        "ba_update_bytes": pe * 20 + M * hw * 20,
        "iter_bytes": pe * 16 + M * hw * 16 + (6 * P) ** 2 * 4,                        # 1.0 GB at C4
    }


# ------------------------------------------------------------------------------------------------
def run_b200(args, world, rank, local):
    import droid_backends as B
    dev = torch.device("cuda", local)
    p = problem(args.frames, args.knn)
    alg = algorithmic(p)
    host = {k: torch.from_numpy(np.ascontiguousarray(p[k])).pin_memory() for k in KEYS}
    d = {k: host[k].to(dev, non_blocking=True) for k in KEYS}
    poses0, disps0 = d["poses"].clone(), d["disps"].clone()
    t0, t1 = p["t0"], p["t1"]
    sharded = world > 1
    # source-frame shards balanced by work; a pure function of the edge list, computed once like the graph itself
    bounds = B.shard_bounds(torch.from_numpy(p["ii"]), p["disps"].shape[0], world) if sharded else None

    def step():
        d["poses"].copy_(poses0)
        d["disps"].copy_(disps0)
        if sharded:
            return B.ba_sharded(d["poses"], d["disps"], d["intrinsics"], d["disps_sens"], d["targets"], d["weights"],
                                d["eta"], d["ii"], d["jj"], t0, t1, ITERS, LM, EP, False, bounds=bounds)
        return B.ba(d["poses"], d["disps"], d["intrinsics"], d["disps_sens"], d["targets"], d["weights"],
                    d["eta"], d["ii"], d["jj"], t0, t1, ITERS, LM, EP, False)

    for _ in range(args.warmup):
        step()
    barrier(world)
    status = B.ba_status(dev)
    B.prof_enable(True)
    B.prof_read()
    clk = ClockSampler(local)
    if rank == 0:
        clk.start()
        time.sleep(0.25)
    n0 = B.launch_count()
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tb = time.time()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier(world)
    te = time.time()
    ms = max_over_ranks(e0.elapsed_time(e1), world, dev)
    launches = B.launch_count() - n0
    phases = B.prof_read()
    B.prof_enable(False)
    clocks = clk.stop(tb, te) if rank == 0 else None
    iters = args.steps * ITERS
    value = iters / (ms * 1e-3)

    # ---- end to end through the public API with HOST buffers --------------------------------------
    e2e = None
    if not args.no_e2e:
        out_h = torch.empty((t1 - t0, 6), dtype=torch.float32).pin_memory()
        pose_h = torch.empty_like(host["poses"]).pin_memory()
        h2d = sum(host[k].numel() * host[k].element_size() for k in KEYS)
        d2h = out_h.numel() * 4 + pose_h.numel() * 4

        # Two-deep pipeline: the pinned-host -> device copy of step s+1 (copy stream) overlaps the BA of step s
        # (compute stream).  Every step still pays its own full H2D of all inputs and its own D2H of dx + poses, and
        # the host reads the results of step s-1 while step s runs; all of it is inside the timed region.
        copy_stream = torch.cuda.Stream(device=dev)
        ekeys = KEYS
        if sharded:
            # every rank owns the edges whose source frame lies in its shard: it uploads only THEIR rows of targets /
            # weights (plus the edge -> row map, host-made like the shard itself); poses, depth maps, eta and the edge
            # list stay replicated.  h2d below counts this rank's bytes; the JSON reports the sum over ranks.
            lo, hi = bounds[rank], bounds[rank + 1]
            mask = (p["ii"] >= lo) & (p["ii"] < hi)
            host = dict(host)
            host["targets"] = torch.from_numpy(np.ascontiguousarray(p["targets"][mask])).pin_memory()
            host["weights"] = torch.from_numpy(np.ascontiguousarray(p["weights"][mask])).pin_memory()
            host["edge_row"] = torch.from_numpy(np.maximum(np.cumsum(mask) - 1, 0).astype(np.int32)).pin_memory()
            ekeys = KEYS + ("edge_row",)
            # per-frame inputs (depth maps, sensor depths, eta): a rank reads only the rows of its own source frames
            # (the depth maps of the other shards arrive through the all-gather at the end of the call), so it uploads
            # only those rows.  eta is indexed by slot in kx = unique(ii U [t0,t1)), which here is every frame in order.
            kx = np.unique(np.concatenate([np.arange(t0, t1), p["ii"]]))
            rows_ok = len(kx) == p["disps"].shape[0] and np.array_equal(kx, np.arange(len(kx)))
            row_keys = ("disps", "disps_sens", "eta") if rows_ok else ()
            h2d = sum(host[k].numel() * host[k].element_size() for k in ekeys if k not in row_keys)
            h2d += sum(host[k][lo:hi].numel() * host[k].element_size() for k in row_keys)
            h2d = int(sum_over_ranks(float(h2d), world, dev))
        bufs = [{k: host[k].to(dev) for k in ekeys} for _ in range(2)]
        outs = [(torch.empty((t1 - t0, 6), dtype=torch.float32).pin_memory(), torch.empty_like(host["poses"]).pin_memory())
                for _ in range(2)]
        ev_copied = [torch.cuda.Event() for _ in range(2)]
        ev_free = [torch.cuda.Event() for _ in range(2)]
        ev_out = [torch.cuda.Event() for _ in range(2)]
        main = torch.cuda.current_stream(dev)

        def enqueue_copy(s):
            i = s & 1
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(ev_free[i])
                for k in ekeys:
                    if sharded and k in row_keys:
                        bufs[i][k][lo:hi].copy_(host[k][lo:hi], non_blocking=True)
                    else:
                        bufs[i][k].copy_(host[k], non_blocking=True)
                ev_copied[i].record(copy_stream)

        def run_pipeline(nsteps):
            chk = 0.0
            for i in range(2):
                ev_free[i].record(main)
            enqueue_copy(0)
            for s in range(nsteps):
                i = s & 1
                if s + 1 < nsteps:
                    enqueue_copy(s + 1)
                main.wait_event(ev_copied[i])
                dd = bufs[i]
                if sharded:
                    dx, _ = B.ba_sharded(dd["poses"], dd["disps"], dd["intrinsics"], dd["disps_sens"], dd["targets"],
                                         dd["weights"], dd["eta"], dd["ii"], dd["jj"], t0, t1, ITERS, LM, EP, False,
                                         edge_row=dd["edge_row"], bounds=bounds)
                else:
                    dx, _ = B.ba(dd["poses"], dd["disps"], dd["intrinsics"], dd["disps_sens"], dd["targets"], dd["weights"],
                                 dd["eta"], dd["ii"], dd["jj"], t0, t1, ITERS, LM, EP, False)
                outs[i][0].copy_(dx, non_blocking=True)
                outs[i][1].copy_(dd["poses"], non_blocking=True)
                ev_out[i].record(main)
                ev_free[i].record(main)
                if s > 0:                         # host consumes the previous step's result
                    ev_out[1 - i].synchronize()
                    chk += float(outs[1 - i][0][0, 0])
            ev_out[(nsteps - 1) & 1].synchronize()
            chk += float(outs[(nsteps - 1) & 1][0][0, 0])
            return chk

        run_pipeline(2)
        barrier(world)
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        run_pipeline(args.steps)
        a1.record()
        barrier(world)
        ems = max_over_ranks(a0.elapsed_time(a1), world, dev)
        e2e = {"value": iters / (ems * 1e-3), "unit": "iters/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": ems / args.steps,
               "pipeline": "H2D of step s+1 overlaps BA of step s (2 device buffers, copy stream)",
               "inputs": ("every rank uploads the replicated poses / edge list and, of targets / weights / depth maps / "
                          "sensor depths / eta, the rows of ITS OWN edges and source frames; h2d_bytes_per_step is the "
                          "sum over ranks") if sharded else "all inputs"}

    parity = parity_check(B, p, d, poses0, disps0, dev, world, rank, bounds) if not args.no_parity else None
    if rank != 0:
        return None
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    per_phase = {k: {"ms_per_launch": v[0] / max(v[1], 1), "launches": v[1], "share": v[0] / ms}
                 for k, v in phases.items() if v[1] > 0}
    tf32_peak = float(peaks.get("bf16_tflops", 1590.0)) / 2.0          # tcgen05 kind::tf32 runs at half the bf16 rate
    if world > 1:                      # the algorithmic bytes / flops below are the whole job's: so is the roof
        hbm_peak *= world
        tf32_peak *= world
    ms_of = lambda k: per_phase[k]["ms_per_launch"] if k in per_phase else float("nan")   # noqa: E731
    # per-phase table: every throughput kernel against the roof that bounds it
    table = {}
    for k in ("ba_accum", "ba_update"):
        if k in per_phase:
            g = alg[k + "_bytes"] / (ms_of(k) * 1e-3) / 1e9
            table[k] = {"bound": "hbm", "ms": ms_of(k), "achieved": g, "peak": hbm_peak, "unit": "GB/s", "frac": g / hbm_peak,
                        "algorithmic_bytes": int(alg[k + "_bytes"])}
            if k + "_streamed_bytes" in alg:          # what the kernel moves by design (intermediate stream included)
                table[k]["streamed_GBps"] = alg[k + "_streamed_bytes"] / (ms_of(k) * 1e-3) / 1e9
                table[k]["streamed_bytes"] = int(alg[k + "_streamed_bytes"])
    if "ba_schur" in per_phase:
        # Which roof binds is decided by the kernel's arithmetic intensity, not by which unit it runs on: algorithmic
        # flops per algorithmic byte (the E_j stream it must read) against the machine balance tf32 peak / HBM peak.
        # At C4 that is ~41 flop/B against ~126: the HBM roof is the lower one (0.17 ms vs 0.05 ms), so `frac` is taken
        # there; the tensor-pipe figures are kept beside it.
        sec = ms_of("ba_schur") * 1e-3
        t = alg["ba_schur_flops"] / sec / 1e12
        g = alg["ba_schur_bytes"] / sec / 1e9
        intensity = alg["ba_schur_flops"] / max(alg["ba_schur_bytes"], 1)
        balance = tf32_peak * 1e12 / (hbm_peak * 1e9)
        hbm_bound = intensity < balance
        table["ba_schur"] = {"bound": "hbm" if hbm_bound else "tensor", "ms": ms_of("ba_schur"),
                             "achieved": g if hbm_bound else t, "peak": hbm_peak if hbm_bound else tf32_peak,
                             "unit": "GB/s" if hbm_bound else "TFLOP/s",
                             "frac": g / hbm_peak if hbm_bound else t / tf32_peak,
                             "algorithmic_bytes": int(alg["ba_schur_bytes"]), "algorithmic_flops": int(alg["ba_schur_flops"]),
                             "flop_per_byte": intensity, "machine_balance_tf32_flop_per_byte": balance,
                             "tensor_TFLOPs": t, "tensor_peak_TFLOPs": tf32_peak, "tensor_frac": t / tf32_peak,
                             "issued_flops": int(alg["ba_schur_issued_flops"]), "hbm_GBps": g,
                             "note": "3xTF32 tcgen05 Gram product of the edge rows (frames with <= 31 edges; > 21 edges: one "
                                     "CTA per pair of edge groups), source-pose rows derived from the Gram per frame; "
                                     "CUDA-core kernel only beyond 31 edges.  What limits it in practice are the builder "
                                     "warps that unpack, split (3xTF32) and re-pack the operand tiles (tensor pipe 47 % "
                                     "busy on 8x the algorithmic flops, DRAM 22 %)"}
    if "ba_solve" in per_phase:
        table["ba_solve"] = {"bound": "latency (serial pivot chain, fp64)", "ms": ms_of("ba_solve"),
                             "pivots": int(6 * (p["t1"] - p["t0"])), "ns_per_pivot": ms_of("ba_solve") * 1e6 / max(6 * (p["t1"] - p["t0"]), 1),
                             "dense_equivalent_tflops": alg["ba_solve_flops"] / (ms_of("ba_solve") * 1e-3) / 1e12,
                             "note": "envelope Cholesky of the re-ordered reduced camera system; ~1% of the dense flops "
                                     "remain, the time is the dependent chain of 6P pivots (panel -> update -> panel)"}
    # `roofline` itself: the largest THROUGHPUT phase (the solve is a latency chain, see table)
    # DRAM traffic per launch of the same kernels from the committed ncu --set full capture (same workload only)
    traffic = {}
    try:
        if p["n"] == 1000 and len(p["ii"]) == 20862 and world == 1:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
    except Exception:
        traffic = {}
    for k in table:
        if k in traffic:
            table[k]["traffic"] = int(traffic[k])
    # `roofline`: the dominant THROUGHPUT kernel (accumulation / Schur / depth update, whichever is longest in this run)
    # against the roof that bounds it.  The longest PHASE is the fp64 factorisation, a latency chain with no bandwidth or
    # flop roof to be a fraction of: it is reported beside it under roofline.longest_phase (ns per pivot) and in the table.
    dom = max((k for k in ("ba_accum", "ba_schur", "ba_update") if k in table), key=lambda k: table[k]["ms"])
    longest = max(table, key=lambda k: table[k]["ms"])
    iter_traffic = sum(int(traffic[k]) for k in ("ba_accum", "ba_schur", "ba_update") if k in traffic) if traffic else None
    roof = {"kernel": dom, "bound": table[dom]["bound"], "achieved": table[dom]["achieved"], "peak": table[dom]["peak"],
            "unit": table[dom]["unit"], "frac": table[dom]["frac"], "traffic": table[dom].get("traffic"),
            "traffic_source": traffic.get("source") if table[dom].get("traffic") else None,
            "peak_source": ("measured (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md)") +
                           ("; tf32 = bf16_tflops / 2" if table[dom]["bound"] == "tensor" or dom == "ba_schur" else "") +
                           ("; x %d GPUs (achieved is the whole job's)" % world if world > 1 else ""),
            "ms_per_launch": table[dom]["ms"],
            "longest_phase": {"kernel": longest, "ms": table[longest]["ms"], "share_of_iteration": table[longest]["ms"] / (ms / iters),
                              "bound": table[longest]["bound"],
                              **({"ns_per_pivot": table[longest]["ns_per_pivot"]} if "ns_per_pivot" in table[longest] else {})},
            "phases": table,
            # the whole Gauss-Newton iteration against SURVEY 8(d)'s compulsory traffic (1.0 GB at C4): what an ideal
            # single-pass BA would be bound by; the Schur Gram product and the fp64 factorisation keep it far from that.
            # dram_bytes = what the three streaming kernels really move per iteration (ncu), the packed E_j stream
            # written once and read twice included
            "iteration": {"algorithmic_bytes": int(alg["iter_bytes"]), "dram_bytes": iter_traffic, "ms": ms / iters,
                          "achieved": alg["iter_bytes"] / (ms / iters * 1e-3) / 1e9, "unit": "GB/s",
                          "frac": alg["iter_bytes"] / (ms / iters * 1e-3) / 1e9 / hbm_peak}}
    if "ba_schur" in table:
        table["ba_schur"]["issued_over_algorithmic_flops"] = alg["ba_schur_issued_flops"] / max(alg["ba_schur_flops"], 1)
    line = {
        "metric": "global BA iters/s @1000 KF", "value": value, "unit": "iters/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32 (reduced camera system assembled and factorised in f64)",
        "data": "synthetic",
        "config": {"workload": "C4 backend global BA: %d keyframes, %d edges, %dx%d, ba(iterations=2, lm=1e-5, ep=1e-2)"
                               % (p["n"], len(p["ii"]), p["ht"], p["wd"]),
                   "edges": int(len(p["ii"])), "keyframes": int(p["n"]), "iters_per_step": ITERS,
                   "parallelism": "1 GPU" if world == 1 else "work-balanced source-frame shards x%d + NCCL all-reduce of the packed reduced camera system" % world,
                   "l2": "inputs>L2 (855 MB of targets+weights per pass vs 126 MB L2)"},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "phases": per_phase,
        "ba_status": status, "impl": "b200", "parity": parity,
    }
    if status != 0:
        line["error"] = "device-side BA status word is %d (see DROID_BA_STATUS_* in include/droid_b200.h)" % status
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args)
    if world == 1 and not args.no_extra:
        try:
            line["extra"] = extra_metrics(dev, hbm_peak, peaks)
        except Exception as ex:                                    # noqa: BLE001
            line["extra"] = {"error": repr(ex)}
    return line


def parity_check(B, p, d, poses0, disps0, dev, world, rank, bounds):
    """Outside the timed region: the timed configuration itself (same inputs, ITERS iterations, same LM / EP) against
    the UNMODIFIED reference extension (oracle/_ref) when it is loadable -- relative errors max|a-b| / max|b| of dx,
    dz, poses and depth maps -- and, under torchrun, ba_sharded against single-GPU ba on rank 0."""
    t0, t1 = p["t0"], p["t1"]
    rel = lambda a, b: float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp(min=1e-30))  # noqa: E731

    def call(fn, **kw):
        po, di = poses0.clone(), disps0.clone()
        dx, dz = fn(po, di, d["intrinsics"], d["disps_sens"], d["targets"], d["weights"], d["eta"], d["ii"], d["jj"],
                    t0, t1, ITERS, LM, EP, False, **kw)
        torch.cuda.synchronize()
        return po, di, dx, dz.reshape(-1)

    out = {"tolerance": "north_star: <= 1e-4 relative per Gauss-Newton iteration (tests/test_gpu_parity.py asserts it per "
                        "iteration at this size); here ONE call of %d iterations, where iteration 2 also carries the "
                        "1e-6 state difference left by iteration 1" % ITERS}
    mine = call(B.ba_sharded, bounds=bounds) if world > 1 else call(B.ba)
    if world > 1:
        if rank == 0:
            single = call(B.ba)
            # ba_sharded returns the depth step of the rank's OWN source frames (the depth maps are all-gathered)
            kx = torch.unique(d["ii"])
            own = ((kx >= bounds[0]) & (kx < bounds[1]))[:, None]
            hw = mine[3].numel() // kx.numel()
            own = own.expand(-1, hw).reshape(-1)
            out["sharded_vs_single_gpu"] = {"dx": rel(mine[2], single[2]), "dz": rel(mine[3][own], single[3][own]),
                                            "poses": rel(mine[0], single[0]), "disps": rel(mine[1], single[1])}
            mine = single
    if rank != 0:
        return None
    try:
        import droid_backends_ref as R
        ref = call(R.ba)
        out["vs_reference_extension"] = {"dx": rel(mine[2], ref[2]), "dz": rel(mine[3], ref[3]),
                                         "poses": rel(mine[0], ref[0]), "disps": rel(mine[1], ref[1])}
        v = out["vs_reference_extension"]
        out["ok"] = bool(v["poses"] <= 1e-5 and v["disps"] <= 1e-5 and v["dx"] <= 5e-4 and v["dz"] <= 5e-4)
    except Exception as ex:                                        # noqa: BLE001
        out["vs_reference_extension"] = {"unavailable": repr(ex)}
        out["ok"] = None
    if "sharded_vs_single_gpu" in out:
        v = out["sharded_vs_single_gpu"]
        out["sharded_ok"] = bool(v["poses"] <= 1e-5 and v["disps"] <= 1e-5 and v["dx"] <= 5e-4 and v["dz"] <= 5e-4)
    return out


def extra_metrics(dev, hbm_peak, peaks):
    """The other two parts of BASELINE.json's metric, measured in the same run (not the headline `value`):
    corr build + 4-level radius-3 lookup at configs[1] (24 edges, 48x64x128 fp16) in GB/s against the measured HBM
    peak, and the frontend-window BA (configs[2]: 12 keyframes, 60 edges, 48x64, iterations=2) as calls/s."""
    import droid_backends as B
    import synth
    out = {}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timeit(fn, n=20, warm=3):
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(n):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        return float(np.median(ts)) * 1e-3

    # write-only HBM bandwidth of this GPU (a plain fill), measured here: the roof of kernels that only write.  On this
    # pool's B200s it is ~3.9 TB/s, well below the 6.55 TB/s that a copy (half reads, half writes) moves.
    fill = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
    tf = []
    for _ in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fill.zero_(); e1.record(); torch.cuda.synchronize(); tf.append(e0.elapsed_time(e1))
    write_peak = (1 << 30) / (sorted(tf)[len(tf) // 2] * 1e-3) / 1e9
    out["hbm_write_only_fill"] = {"GBps": write_peak, "workload": "torch fill of 1 GiB, median of 6"}
    del fill
    Bn, H, W, C, r = 24, 48, 64, 128, 3
    q = synth.make_corr_problem(Bn, H, W, C, seed=1234)
    f1 = torch.from_numpy(q["fmaps"][q["ii"]]).to(dev)
    f2 = torch.from_numpy(q["fmaps"][q["jj"]]).to(dev)
    coords = torch.from_numpy(q["coords"]).to(dev)
    HW = H * W
    t = timeit(lambda: B.corr_build(f1, f2))
    vols = B.corr_build(f1, f2)
    by = Bn * HW * (HW + HW // 4 + HW // 16 + HW // 64) * 2 + 2 * Bn * C * HW * 2
    fl = 2.0 * HW * HW * C * Bn
    tpeak = float(peaks.get("bf16_tflops", 1590.0))
    out["corr_build"] = {"us": t * 1e6, "GBps": by / t / 1e9, "hbm_frac": by / t / 1e9 / hbm_peak, "TFLOPs": fl / t / 1e12,
                         "tensor_frac": fl / t / 1e12 / tpeak, "bound": "hbm (write)", "algorithmic_bytes": by,
                         "write_GBps": (by - 2 * Bn * C * HW * 2) / t / 1e9,
                         "write_roof_frac": (by - 2 * Bn * C * HW * 2) / t / 1e9 / write_peak,
                         "workload": "24 edges, 48x64x128 fp16 -> 4-level pyramid", "l2": "flushed between calls"}
    # lookup: coords (8 B) + up to 64 taps in + 49 out per pixel per level, 2 B elements (SURVEY 8d upper bound)
    lb = 4 * Bn * HW * (8 + 64 * 2 + 49 * 2)
    t = timeit(lambda: B.corr_index_pyramid(vols, coords, r))
    out["corr_lookup_fused"] = {"us": t * 1e6, "GBps": lb / t / 1e9, "hbm_frac": lb / t / 1e9 / hbm_peak,
                                "algorithmic_bytes": lb, "workload": "4 levels, radius 3, 24 edges, one launch"}
    cl = [coords / 2 ** l for l in range(4)]
    t = timeit(lambda: [B.corr_index_forward(vols[l], cl[l], r) for l in range(4)])
    out["corr_lookup_4calls"] = {"us": t * 1e6, "GBps": lb / t / 1e9, "hbm_frac": lb / t / 1e9 / hbm_peak,
                                 "workload": "the reference call pattern: 4x corr_index_forward"}
    try:
        import droid_backends_ref as R
        t = timeit(lambda: [R.corr_index_forward(vols[l], cl[l], r) for l in range(4)])
        out["corr_lookup_4calls_reference_ext"] = {"us": t * 1e6, "GBps": lb / t / 1e9}
    except Exception:
        pass
    # reproject + motion features on the edge list of a 250-keyframe backend graph: one launch instead of
    # the ~15 lietorch / elementwise kernels of DepthVideo.reproject; depth in, coords + valid + 4 feature planes out
    try:
        pr = synth.make_ba_problem("C4", n_frames=250, knn=16)
        nb = pr["disps"].shape[0]
        Kf = torch.from_numpy(np.tile(pr["intrinsics"][None], (nb, 1)).astype(np.float32)).to(dev)
        rp, rd = torch.from_numpy(pr["poses"]).to(dev), torch.from_numpy(pr["disps"]).to(dev)
        ri, rj = torch.from_numpy(pr["ii"]).to(dev), torch.from_numpy(pr["jj"]).to(dev)
        rt = torch.from_numpy(np.ascontiguousarray(pr["targets"].transpose(0, 2, 3, 1))).to(dev)
        ne, hw = len(pr["ii"]), pr["ht"] * pr["wd"]
        t = timeit(lambda: B.reproject(rp, rd, Kf, ri, rj, rt), n=10)
        rb = ne * hw * (4 + 8 + 8 + 4 + 16)
        out["reproject_motion"] = {"us": t * 1e6, "GBps": rb / t / 1e9, "hbm_frac": rb / t / 1e9 / hbm_peak,
                                   "algorithmic_bytes": rb, "bound": "hbm",
                                   "workload": "%d edges, %dx%d: depth + target in, coords + valid + 4 motion planes out" % (ne, pr["ht"], pr["wd"])}
        del rt
    except Exception as ex:                                        # noqa: BLE001
        out["reproject_motion"] = {"error": repr(ex)}
    # convex upsampling: 64 inverse-depth maps 48x64 -> 384x512, float32 mask (2304 B in, 260 B out per pixel)
    try:
        nb_, hh, ww = 64, 48, 64
        dat = torch.rand((nb_, hh, ww, 1), device=dev) + 0.1
        msk = torch.randn((nb_, 576, hh, ww), device=dev)
        t = timeit(lambda: B.cvx_upsample(dat, msk), n=10)
        ub = nb_ * hh * ww * (576 * 4 + 4 + 256)
        out["cvx_upsample"] = {"us": t * 1e6, "GBps": ub / t / 1e9, "hbm_frac": ub / t / 1e9 / hbm_peak, "bound": "hbm",
                               "algorithmic_bytes": ub, "workload": "64 maps 48x64 -> 384x512, float32 mask"}
        del msk
    except Exception as ex:                                        # noqa: BLE001
        out["cvx_upsample"] = {"error": repr(ex)}
    # segmented mean of GraphAgg: 60 edges x 128 channels x 48x64 float16 -> 12 source frames
    try:
        ne_, m_ = 60, 12
        srcm = torch.randn((1, ne_, 128, 48, 64), device=dev).half()
        ixm = (torch.arange(ne_, device=dev) % m_).long()
        t = timeit(lambda: B.scatter_mean(srcm, ixm, dim_size=m_), n=10)
        sb = (ne_ + m_) * 128 * 48 * 64 * 2
        out["scatter_mean"] = {"us": t * 1e6, "GBps": sb / t / 1e9, "hbm_frac": sb / t / 1e9 / hbm_peak, "bound": "hbm",
                               "algorithmic_bytes": sb, "workload": "60 edges x 128 x 48x64 float16 -> 12 frames"}
        del srcm
    except Exception as ex:                                        # noqa: BLE001
        out["scatter_mean"] = {"error": repr(ex)}
    # edge selection at the backend's size: 1000 keyframes = 10^6 candidate pairs, on the device vs the
    # restatement of the reference's Python loops on one host core
    try:
        import droid_oracle as O
        rng = np.random.default_rng(11)
        tk = 1000
        sa = np.arange(tk) * (2 * np.pi / 250.0)
        pos = np.stack([10 * np.cos(sa), 10 * np.sin(sa), 0.002 * np.arange(tk)], -1) + rng.normal(0, 0.2, size=(tk, 3))
        gi, gj = np.meshgrid(np.arange(tk), np.arange(tk), indexing="ij")
        dd = (np.linalg.norm(pos[gi.reshape(-1)] - pos[gj.reshape(-1)], axis=-1) * 6.0
              + rng.uniform(0, 1e-3, size=tk * tk)).astype(np.float32)
        ddev = torch.from_numpy(dd).to(dev)
        none = torch.zeros(0, dtype=torch.long, device=dev)
        t = timeit(lambda: B.proximity_edges(ddev, tk, 0, 0, 2, 3, 22.0, 16 * tk, False, none, none), n=10)
        c0 = time.perf_counter()
        ri, _ = O.proximity_edges(dd, tk, 0, 0, 2, 3, 22.0, 16 * tk, False, np.zeros(0, np.int64), np.zeros(0, np.int64))
        cpu_s = time.perf_counter() - c0
        out["proximity_edges"] = {"us": t * 1e6, "edges": int(len(ri)), "cpu_port_s": cpu_s, "speedup_vs_cpu_port": cpu_s / t,
                                  "workload": "1000 keyframes, 10^6 candidate pairs, rad 2, nms 3, thresh 22 (backend settings)"}
    except Exception as ex:                                        # noqa: BLE001
        out["proximity_edges"] = {"error": repr(ex)}
    try:
        out.update(altcorr_and_geometry_metrics(dev, hbm_peak, timeit))
    except Exception as ex:                                        # noqa: BLE001
        out["altcorr_geometry_error"] = repr(ex)
    out["frontend_driver"] = frontend_driver_metric()
    p = synth.make_ba_problem("C3")
    d = {k: torch.from_numpy(np.ascontiguousarray(p[k])).to(dev) for k in KEYS}
    p0, d0 = d["poses"].clone(), d["disps"].clone()

    def fba(mod, **kw):
        d["poses"].copy_(p0); d["disps"].copy_(d0)
        mod.ba(d["poses"], d["disps"], d["intrinsics"], d["disps_sens"], d["targets"], d["weights"], d["eta"],
               d["ii"], d["jj"], p["t0"], p["t1"], 2, 1e-4, 0.1, False, **kw)
    t = timeit(lambda: fba(B), n=50)
    tn = timeit(lambda: fba(B, _validate=False), n=50)
    out["frontend_ba"] = {"us_per_call": t * 1e6, "calls_per_s": 1.0 / t, "us_per_call_unvalidated": tn * 1e6,
                          "note": "default call: the device-side input validation is read back (one host wait); "
                                  "unvalidated = DROID_BA_CHECK=0, fully stream-ordered like round 1",
                          "workload": "C3 frontend window: 12 keyframes, 60 edges, 48x64, ba(iterations=2, lm=1e-4, ep=0.1)"}
    try:
        import droid_backends_ref as R
        t = timeit(lambda: fba(R), n=20)
        out["frontend_ba_reference_ext"] = {"us_per_call": t * 1e6, "calls_per_s": 1.0 / t}
    except Exception:
        pass
    return out


def frontend_driver_metric(frames=48):
    """BASELINE's "frontend fps", through the reference's OWN DroidFrontend (unmodified droid_slam/*.py staged under
    oracle/_ref/pyref) on a synthetic stream of `frames` keyframes with a random-init DroidNet: the same Python with
    `droid_backends` bound to the reference extension, to this repo's module (drop-in) and to this repo's module plus the
    bindings INTEGRATION.md proposes for the rows of SURVEY 8f (integrated).  Runs tools/run_reference_drivers.py in a
    child process (a measurement tool: the only place besides cpu_baseline where bench.py touches oracle/); DESIGN.md 5.1
    holds the 100-keyframe run and the 1000-keyframe backend run of the same tool."""
    import subprocess
    tool = os.path.join(ROOT, "tools", "run_reference_drivers.py")
    try:
        r = subprocess.run([sys.executable, tool, "frontend", "--frames", str(frames), "--arms", "reference,drop-in,integrated"],
                           capture_output=True, text=True, timeout=420)
        if r.returncode != 0:
            return {"unavailable": (r.stderr or r.stdout).strip().splitlines()[-1][:300] if (r.stderr or r.stdout) else "rc %d" % r.returncode}
        res = json.loads(r.stdout[r.stdout.index("{"):])
        out = {"workload": "DroidFrontend.__call__ per keyframe (edge selection + 6 FactorGraph.update: corr lookup, update operator, "
                           "ba on the window), %d keyframes 40x64 at 1/8, random-init DroidNet in PyTorch" % frames}
        for arm, a in res["arms"].items():
            out[arm] = {"keyframes_per_s": a["keyframes_per_s"], "droid_backends_ms": a["droid_backends_ms"],
                        "ba_ms": a["droid_backends_calls"].get("ba", {}).get("ms"), "finite": a["finite"]}
            if "vs_reference" in a:
                out[arm]["vs_reference"] = a["vs_reference"]
        return out
    except Exception as ex:                                        # noqa: BLE001
        return {"unavailable": repr(ex)[:300]}


def altcorr_and_geometry_metrics(dev, hbm_peak, timeit):
    """The other half of BASELINE configs[3] and the geometry operators, L2 flushed between calls, the unmodified
    reference extension beside every figure where it is loadable:
      * alt-corr on ONE chunk of the backend's update_lowmem loop (droid_slam/factor_graph.py:262-270): the edges of
        8 consecutive source frames of a 250-keyframe graph (~160), 40x64, C = 128, 4 pyramid levels, radius 3 --
        the drop-in signature (float32 gathered copies, modules/corr.py:113-120) and the indexed float16-pyramid
        extension; flops = B HW 64 C 2 per level, bytes per SURVEY 8(d);
      * frame_distance on the 10^6 pairs of a 1000-keyframe video (depth_video.py:160-170), pairs/s;
      * projmap / iproj / depth_filter, one line each."""
    import droid_backends as B
    import synth
    try:
        import droid_backends_ref as R
    except Exception:                                              # noqa: BLE001
        R = None
    out = {}
    pr = synth.make_ba_problem("C4", n_frames=250, knn=16)
    H, W, C, r = pr["ht"], pr["wd"], 128, 3
    HW = H * W
    sel = np.nonzero((pr["ii"] >= 120) & (pr["ii"] < 128))[0]
    ii = torch.from_numpy(pr["ii"][sel]).to(dev)
    jj = torch.from_numpy(pr["jj"][sel]).to(dev)
    Bn = len(sel)
    g = torch.Generator(device="cpu").manual_seed(5)
    fm = (torch.randn((250, C, H, W), generator=g) * 0.5).to(dev)                # [frames, C, H, W] like video.fmaps / 4
    pyr = []
    f = fm
    for l in range(4):
        pyr.append(f.permute(0, 2, 3, 1).contiguous().half())
        f = torch.nn.functional.avg_pool2d(f, 2, stride=2)
    del fm, f
    coords = torch.from_numpy(np.ascontiguousarray(pr["targets"][sel].transpose(0, 2, 3, 1))).to(dev)[:, None].contiguous()
    cl = [(coords / 2 ** l).contiguous() for l in range(4)]
    f1 = pyr[0][ii].float().contiguous()
    f2 = [pyr[l][jj].float().contiguous() for l in range(4)]
    flops = sum(Bn * HW * 64 * C * 2 for l in range(4))
    by32 = sum(Bn * (HW + (H >> l) * (W >> l)) * C * 4 + Bn * HW * 8 + Bn * 49 * HW * 4 for l in range(4))
    by16 = sum(Bn * (HW + (H >> l) * (W >> l)) * C * 2 + Bn * HW * 8 + Bn * 49 * HW * 4 for l in range(4))

    def rec(t, by):
        return {"us": t * 1e6, "GFLOPs": flops / t / 1e9, "GBps": by / t / 1e9, "hbm_frac": by / t / 1e9 / hbm_peak,
                "algorithmic_bytes": int(by), "flops": int(flops)}
    wl = "%d edges (8 source frames of a 250-keyframe backend graph), %dx%d, C=128, 4 levels, radius 3" % (Bn, H, W)
    t = timeit(lambda: [B.altcorr_forward(f1, f2[l], cl[l], r) for l in range(4)], n=10)
    out["altcorr_forward"] = dict(rec(t, by32), workload=wl + "; drop-in signature, float32 gathered copies, 4 calls (fp16 copies + exactness flag, then the tensor-core kernel)")
    t = timeit(lambda: [B.altcorr_forward_indexed(pyr[0], pyr[l], ii, jj, cl[l], r) for l in range(4)], n=10)
    out["altcorr_forward_indexed"] = dict(rec(t, by16), workload=wl + "; float16 pyramid indexed by (ii, jj), no gathered copies, 4 calls")
    t = timeit(lambda: B.altcorr_pyramid_indexed(pyr, ii, jj, coords, r), n=10)
    out["altcorr_pyramid_indexed"] = dict(rec(t, by16), workload=wl + "; float16 pyramid, all 4 levels in ONE launch (tcgen05 GEMM of 128 "
                                          "source pixels against the target rows their windows touch, sampled from TMEM)",
                                          bound="latency per CTA (prologue -> TMA -> MMA -> TMEM read -> 196 stores per pixel); "
                                                "HBM floor = the 321 MB of output")
    t = timeit(lambda: [B.altcorr_forward(pyr[0][ii].float(), pyr[l][jj].float(), cl[l], r) for l in range(4)], n=10)
    out["altcorr_forward_with_gathers"] = {"us": t * 1e6, "workload": wl + "; as AltCorrBlock.corr_fn calls it: pyramid[l][:, jj].float() gathers + the op"}
    if R is not None:
        t = timeit(lambda: [R.altcorr_forward(f1, f2[l], cl[l], r) for l in range(4)], n=5)
        out["altcorr_forward_reference_ext"] = rec(t, by32)
        out["altcorr_forward"]["speedup_vs_reference_ext"] = out["altcorr_forward_reference_ext"]["us"] / out["altcorr_forward"]["us"]
        out["altcorr_forward_indexed"]["speedup_vs_reference_ext"] = out["altcorr_forward_reference_ext"]["us"] / out["altcorr_forward_indexed"]["us"]
        out["altcorr_pyramid_indexed"]["speedup_vs_reference_ext"] = out["altcorr_forward_reference_ext"]["us"] / out["altcorr_pyramid_indexed"]["us"]
    del f1, f2, pyr, coords, cl

    # ---- frame_distance: every pair of a 1000-keyframe video, one direction (the caller averages two calls)
    pk = synth.make_ba_problem("C4", n_frames=1000, knn=2)
    n = 1000
    P, D, K = (torch.from_numpy(pk[k]).to(dev) for k in ("gt_poses", "gt_disps", "intrinsics"))
    a, b = torch.meshgrid(torch.arange(n, device=dev), torch.arange(n, device=dev), indexing="ij")
    a, b = a.reshape(-1).contiguous(), b.reshape(-1).contiguous()
    t = timeit(lambda: B.frame_distance(P, D, K, a, b, 0.3), n=5, warm=1)
    fb = n * pk["ht"] * pk["wd"] * 4 + a.numel() * 20
    out["frame_distance"] = {"us": t * 1e6, "pairs_per_s": a.numel() / t, "pairs": int(a.numel()),
                             "pixel_pairs_per_s": a.numel() * pk["ht"] * pk["wd"] / t, "algorithmic_bytes": int(fb),
                             "workload": "10^6 pairs (meshgrid of 1000 keyframes), 40x64, beta 0.3, one direction",
                             "bound": "ALU / L2 (every pair re-projects 2560 pixels twice; depth maps are L2 resident)"}
    if R is not None:
        t2 = timeit(lambda: R.frame_distance(P, D, K, a, b, 0.3), n=3, warm=1)
        out["frame_distance_reference_ext"] = {"us": t2 * 1e6, "pairs_per_s": a.numel() / t2}
        out["frame_distance"]["speedup_vs_reference_ext"] = t2 / t
        out["frame_distance"]["bit_exact_vs_reference_ext"] = bool(torch.equal(B.frame_distance(P, D, K, a, b, 0.3),
                                                                                R.frame_distance(P, D, K, a, b, 0.3)))
    # ---- projmap / iproj / depth_filter
    ei = torch.from_numpy(pr["ii"]).to(dev); ej = torch.from_numpy(pr["jj"]).to(dev)
    ne = ei.numel()
    ix = torch.arange(n, device=dev)
    th = torch.full((n,), 0.05, device=dev)
    ops = {
        "projmap": (lambda M: M.projmap(P, D, K, ei, ej), ne * HW * (4 + 12 + 4), "%d edges, 40x64: depth in, coords[3] + valid out" % ne),
        "iproj": (lambda M: M.iproj(P, D, K), n * HW * (4 + 12), "1000 frames, 40x64: depth in, points[3] out"),
        "depth_filter": (lambda M: M.depth_filter(P, D, K, ix, th), n * HW * (4 + 4), "1000 frames x 6 temporal neighbours, 40x64 (depth maps L2 resident)"),
    }
    for name, (fn, by, wl_) in ops.items():
        t = timeit(lambda: fn(B), n=10)
        out[name] = {"us": t * 1e6, "GBps": by / t / 1e9, "hbm_frac": by / t / 1e9 / hbm_peak, "algorithmic_bytes": int(by), "workload": wl_}
        if R is not None:
            t2 = timeit(lambda: fn(R), n=5)
            out[name]["reference_ext_us"] = t2 * 1e6
            out[name]["speedup_vs_reference_ext"] = t2 / t
    return out


# ------------------------------------------------------------------------------------------------
def cpu_baseline(args, frames=None):
    """The reference's OWN CPU implementation of the path, timed on this box's host cores on a bounded sample:
    droid_slam/geom/ba.py:31-106 (BA) + geom/chol.py:46-73 + geom/projective_ops.py:96-125 on the reference's
    vendored lietorch (CPU build, oracle/build_lietorch.sh), the unmodified Python staged under oracle/_ref/pyref and
    loaded by oracle/pyref.py.  Two samples: BASELINE configs[0] as it is (8 frames, 24 edges, 48x64; all cores and
    1 core) and the headline problem cut to `frames` keyframes -- the dense Schur solve of geom/chol.py needs
    P*M*6*HW*4 bytes for E (61 GB at 1000 keyframes), so the full size cannot run; `value` scales the cut linearly by
    edges (optimistic for the CPU).  Falls back to the numpy port of the CUDA op when the staged reference is missing."""
    frames = frames or args.cpu_frames
    try:
        import pyref
        import synth
        ok, why = pyref.available()
        if not ok:
            raise ImportError(why)
        import droid_backends as B
        ns = pyref.load(B, cpu=True)
        SE3 = ns.lietorch.SE3

        def run_ba(p, nthreads, reps):
            torch.set_num_threads(nthreads)
            n = p["poses"].shape[0]
            poses = SE3(torch.from_numpy(p["poses"])[None])
            disps = torch.from_numpy(p["disps"])[None]
            intr = torch.from_numpy(np.tile(p["intrinsics"][None], (n, 1)))[None]
            ii, jj = torch.from_numpy(p["ii"]), torch.from_numpy(p["jj"])
            target = torch.from_numpy(p["targets"]).permute(0, 2, 3, 1)[None].contiguous()
            weight = torch.from_numpy(p["weights"]).permute(0, 2, 3, 1)[None].contiguous()
            kx = torch.unique(ii)
            eta = torch.from_numpy(p["eta"])[None][:, :len(kx)]
            ns.ba.BA(target, weight, eta, poses, disps, intr, ii, jj, fixedp=p["t0"])      # warm-up
            ts = []
            for _ in range(reps):
                t = time.perf_counter()
                ns.ba.BA(target, weight, eta, poses, disps, intr, ii, jj, fixedp=p["t0"])
                ts.append(time.perf_counter() - t)
            return float(np.median(ts))
        cores = os.cpu_count() or 1
        c1 = synth.make_ba_problem("C1")
        t_c1_all, t_c1_one = run_ba(c1, cores, 5), run_ba(c1, 1, 5)
        q = problem(frames, args.knn)
        threads = cores if t_c1_all <= t_c1_one else 1            # lietorch's OpenMP + torch threads can oversubscribe
        dt = run_ba(q, threads, 2)
        torch.set_num_threads(cores)
        full = problem_edges(args.frames, args.knn)
        scale = len(q["ii"]) / float(full)
        return {"value": (1.0 / dt) * scale, "unit": "iters/s", "cores": threads, "kind": "reference",
                "sample": "the reference's pure-PyTorch BA (droid_slam/geom/ba.py + chol.py + projective_ops.py on its own "
                          "lietorch, CPU) on the C4 problem cut to %d keyframes / %d edges: %.2f s per Gauss-Newton "
                          "iteration on %d thread(s); value = that rate scaled linearly by edges to %d edges (the full "
                          "size needs 61 GB for the dense E of geom/chol.py)" % (frames, len(q["ii"]), dt, threads, full),
                "sample_seconds": dt, "host_cores": cores,
                "config0": {"workload": "BASELINE configs[0]: 8 frames, 24 edges, 48x64, one BA iteration",
                            "ms_all_cores": t_c1_all * 1e3, "ms_one_core": t_c1_one * 1e3, "cores": cores}}
    except Exception as ex:                                        # noqa: BLE001
        note = "staged Python reference unavailable (%r): numpy port timed instead" % (ex,)
    import droid_oracle as O
    q = problem(frames, args.knn)
    t = time.perf_counter()
    O.ba(q["poses"].copy(), q["disps"].copy(), q["intrinsics"], q["disps_sens"], q["targets"], q["weights"],
         q["eta"], q["ii"], q["jj"], q["t0"], q["t1"], 1, LM, EP, False)
    dt = time.perf_counter() - t
    full = problem_edges(args.frames, args.knn)
    scale = len(q["ii"]) / float(full)
    return {"value": (1.0 / dt) * scale, "unit": "iters/s", "cores": 1, "kind": "port", "note": note,
            "sample": "1 Gauss-Newton iteration of the C4 problem cut to %d keyframes / %d edges took %.2f s on 1 thread "
                      "(numpy port of src/droid_kernels.cu); value = that rate scaled linearly by edges to %d edges"
                      % (frames, len(q["ii"]), dt, full),
            "sample_seconds": dt, "host_cores": os.cpu_count()}


def problem_edges(frames, knn):
    import synth
    Rs, ts = synth.trajectory(frames, loop=250)
    return len(synth.edges_global(Rs, ts, rad=2, knn=knn)[0])


# ------------------------------------------------------------------------------------------------
def run_reference(args, world, rank, local):
    """The reference's own implementation of the path: the UNMODIFIED droid_backends extension compiled from
    /root/reference/src by oracle/build_ref.sh (CUDA kernels recompiled for sm_100 + its host-side Eigen sparse
    LLT / O(P^2 deg^2) triple enumeration on one CPU thread, src/droid_kernels.cu:1117-1311).  Falls back to the
    CPU port when that extension is not loadable."""
    if rank != 0:
        return None
    try:
        import droid_backends_ref as R
    except Exception as ex:                                        # noqa: BLE001
        cb = cpu_baseline(args)
        v = cb["value"]
        return {"impl": "reference", "metric": "global BA iters/s @1000 KF", "value": v, "unit": "iters/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ITERS / v * 1e3,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "C4 backend global BA (bounded CPU sample, see cpu_baseline.sample)"},
                "cpu_baseline": cb, "note": "oracle/_ref not loadable (%s): CPU port timed instead" % type(ex).__name__,
                "e2e": {"value": v, "unit": "iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    dev = torch.device("cuda", 0)
    p = problem(args.frames, args.knn)
    d = {k: torch.from_numpy(np.ascontiguousarray(p[k])).to(dev) for k in KEYS}
    poses0, disps0 = d["poses"].clone(), d["disps"].clone()

    def step():
        d["poses"].copy_(poses0)
        d["disps"].copy_(disps0)
        R.ba(d["poses"], d["disps"], d["intrinsics"], d["disps_sens"], d["targets"], d["weights"], d["eta"],
             d["ii"], d["jj"], p["t0"], p["t1"], ITERS, LM, EP, False)

    # bound the run: the driver's K and W stand, but never more than ~4 minutes in total
    torch.cuda.synchronize()
    t = time.perf_counter(); step(); torch.cuda.synchronize()
    probe = time.perf_counter() - t
    steps, warm = args.steps, max(args.warmup - 1, 0)
    budget = 240.0
    note = None
    if probe * (steps + warm) > budget:
        warm = 0
        steps = max(1, int(budget / probe))
        note = "probe call took %.1f s; timed %d of the requested %d steps to stay within minutes" % (probe, steps, args.steps)
    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    clk = ClockSampler(0); clk.start(); time.sleep(0.25)
    tb = time.time()
    t = time.perf_counter()
    for _ in range(steps):
        step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t
    clocks = clk.stop(tb, time.time())
    v = steps * ITERS / dt
    return {"impl": "reference", "metric": "global BA iters/s @1000 KF", "value": v, "unit": "iters/s", "n_gpus": world,
            "steps": steps, "warmup": args.warmup, "ms_per_step": dt / steps * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32 (+ f64 host assembly and sparse LLT)", "data": "synthetic",
            "config": {"workload": "C4 backend global BA: %d keyframes, %d edges, %dx%d, ba(iterations=2, lm=1e-5, ep=1e-2)"
                                   % (p["n"], len(p["ii"]), p["ht"], p["wd"]),
                       "edges": int(len(p["ii"])), "keyframes": int(p["n"]), "iters_per_step": ITERS,
                       "parallelism": "reference is single-GPU (README.md:23); rank 0 only"},
            "cpu_baseline": {"value": v, "unit": "iters/s", "cores": 1, "kind": "reference",
                             "sample": "unmodified reference droid_backends.ba (oracle/_ref, sources compiled where they lie for "
                                       "sm_100) on the full workload: its CUDA kernels on GPU 0 plus its single-threaded host "
                                       "assembly / Eigen SimplicialLLT, wall clock"},
            "e2e": {"value": v, "unit": "iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "clocks": clocks, "note": note}


def main():
    args = parse()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for the product path)")
    if args.impl == "reference":
        # rank 0 alone runs the single-GPU reference; the other ranks exit 0 without work (no rendezvous)
        if int(os.environ.get("RANK", "0")) == 0:
            torch.cuda.set_device(0)
            line = run_reference(args, int(os.environ.get("WORLD_SIZE", "1")), 0, 0)
            print(json.dumps(line), flush=True)
        return
    world, rank, local = setup_dist(args)
    if True:
        line = run_b200(args, world, rank, local)
        if rank == 0:
            print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
