#!/usr/bin/env python
"""bench.py -- dense BA hot path on B200: edge-pixel residuals/s and ms per BA iteration.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config E] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

A "step" is one `droid_backends.ba` call (`iters` Gauss-Newton iterations, BASELINE.json's
backend setting 2) over one synthetic factor graph.  Default workload: config E (ScanNet-scale
global BA, 1000 keyframes / 20000 edges / 48x64), the configuration BASELINE.json's metric and
target are quoted on; it fits one GPU.  At N > 1 the SAME graph is edge-sharded by source keyframe
over the ranks (strong scaling) with one NCCL all-reduce of the packed reduced camera system per
iteration.

Printed JSON (one line, rank 0): the driver contract plus
  roofline     -- the per-edge HBM kernel (k_linearize): algorithmic bytes / its CUDA-event time
  stages       -- device ms per step of every stage (events on the launch stream, separate pass)
  cpu_baseline -- the CPU oracle (port of the reference's ba_cuda) on a bounded sample, host cores
  e2e          -- same metric through the public API with pinned HOST buffers, copies timed
`--impl reference` times the CPU port on the host cores (the reference's Python BA cannot travel
to the GPU box and its CUDA build has no CPU path).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "glorie-slam_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "edge-pixel residuals/sec (global BA; ms per BA iteration in ms_per_ba_iteration)"
UNIT = "edge-pixels/s"
# bounded CPU sample = config C (backend global BA, 200 keyframes / 3000 edges): a few seconds per run
CPU_SAMPLE = dict(n=200, e=3000, r=7, iters=2, lm=1e-5, ep=1e-2, seed=0)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback (B200_PROFILING.md)"


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.path)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons),
                       samples=len(sm))
        return out


def cpu_baseline(steps=1):
    """CPU oracle (C port of the reference's ba_cuda, OpenMP over edges / keyframes) on a bounded
    sample of the same kind of graph; edge-pixels/s on this box's host cores."""
    import synth
    from oracle import oracle

    g = synth.make_graph(**CPU_SAMPLE)
    best = None
    for _ in range(max(steps, 1)):
        t = time.perf_counter()
        oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  g.iters, g.lm, g.ep, precision="f32")
        dt = time.perf_counter() - t
        best = dt if best is None else min(best, dt)
    return {
        "value": g.n_edges * g.hw * g.iters / best, "unit": UNIT, "cores": oracle.num_threads(), "kind": "port",
        "sample": "%d keyframes / %d edges / %dx%d / %d iterations, best of %d (%.2f s each)" % (
            g.n_frames, g.n_edges, g.disps.shape[1], g.disps.shape[2], g.iters, max(steps, 1), best),
        "ms_per_ba_iteration": 1e3 * best / g.iters,
    }


def reference_cuda_side_by_side(dev, name="C", reps=5):
    """Same-box GPU baseline: the reference's own CUDA droid_backends.ba (oracle/_ref, rebuilt
    unmodified for sm_100 by oracle/build_ref.py) against libgba on config C, device resident,
    CUDA events, best of `reps`.  Config E is left out: the reference's host-side Schur loop
    (droid_kernels.cu:1257-1272) makes a single call take tens of seconds there."""
    import importlib.util

    import torch

    import droid_backends
    import synth

    so = os.path.join(ROOT, "oracle", "_ref", "ref_droid_backends.so")
    if not os.path.exists(so):
        return None
    try:
        spec = importlib.util.spec_from_file_location("ref_droid_backends", so)
        ref = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(ref)
        g = synth.make_config(name)
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
        base = [t(x) for x in (g.poses, g.disps)]
        rest = [t(x) for x in (g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj)]
        out = {}
        for label, fn in (("reference", ref.ba), ("libgba", droid_backends.ba)):
            best = None
            for _ in range(reps + 1):
                p_, d_ = base[0].clone(), base[1].clone()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                a.record()
                fn(p_, d_, *rest, g.t0, g.t1, g.iters, g.lm, g.ep, False, False)
                b.record()
                torch.cuda.synchronize()
                ms = a.elapsed_time(b)
                best = ms if best is None else min(best, ms)
            out[label + "_ms_per_ba_iteration"] = best / g.iters
        out["config"] = "config %s: %d keyframes / %d edges, %d iterations, device resident" % (
            name, g.n_frames, g.n_edges, g.iters)
        out["speedup"] = out["reference_ms_per_ba_iteration"] / out["libgba_ms_per_ba_iteration"]
        # frame_distance over all keyframe pairs of the same graph
        n = g.n_frames
        gi, gj = torch.meshgrid(torch.arange(n, device=dev), torch.arange(n, device=dev), indexing="ij")
        gi, gj = gi.reshape(-1).contiguous(), gj.reshape(-1).contiguous()
        for label, fn in (("reference", ref.frame_distance), ("libgba", droid_backends.frame_distance)):
            best = None
            for _ in range(reps + 1):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                a.record()
                fn(base[0], base[1], rest[0], gi, gj, 0.3)
                b.record()
                torch.cuda.synchronize()
                ms = a.elapsed_time(b)
                best = ms if best is None else min(best, ms)
            out[label + "_frame_distance_ms"] = best
        out["frame_distance_pairs"] = int(gi.numel())
        # depth_filter over every keyframe + the torch form of align_scale_and_shift
        import common

        ix = torch.arange(n, device=dev)
        th = (0.05 / base[1].mean(dim=(1, 2))).contiguous()

        def torch_align(p, t_, w):  # src/utils/common.py:402-438 verbatim semantics, the reference's own torch ops
            a00, a01, a11 = (w * p * p).sum((1, 2)), (w * p).sum((1, 2)), w.sum((1, 2))
            b0, b1 = (w * p * t_).sum((1, 2)), (w * t_).sum((1, 2))
            det = a00 * a11 - a01 * a01
            s_, q_ = (a11 * b0 - a01 * b1) / det, (-a01 * b0 + a00 * b1) / det
            e_ = ((s_[:, None, None] * p + q_[:, None, None] - t_).abs() * w).sum((1, 2)) / a11
            return s_, q_, e_

        wts = (base[1] > base[1].median()).float()
        for key, fn in (("reference_depth_filter_ms", lambda: ref.depth_filter(base[0], base[1], rest[0], ix, th)),
                        ("libgba_depth_filter_ms", lambda: droid_backends.depth_filter(base[0], base[1], rest[0], ix, th)),
                        ("torch_align_scale_shift_ms", lambda: torch_align(rest[1], base[1], wts)),
                        ("libgba_align_scale_shift_ms", lambda: common.align_scale_and_shift(rest[1], base[1], wts))):
            best = None
            for _ in range(reps + 1):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                a.record()
                fn()
                b.record()
                torch.cuda.synchronize()
                ms = a.elapsed_time(b)
                best = ms if best is None else min(best, ms)
            out[key] = best
        return out
    except Exception as e:  # the prebuilt reference is optional evidence
        return {"unavailable": repr(e)[:200]}


def run_reference(args, rank):
    if rank != 0:
        return
    for _ in range(args.warmup if args.warmup < 2 else 1):
        cpu_baseline(1)
    cb = cpu_baseline(args.steps)
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_ba_iteration"] * CPU_SAMPLE["iters"],
        "ms_per_ba_iteration": cb["ms_per_ba_iteration"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "config %s (global BA), timed on the bounded CPU sample in cpu_baseline.sample" % args.config},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="E", choices=["A", "B", "C", "E"])
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--skip-cpu", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist

    import _gba
    import droid_backends
    import sharded
    import synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libgba has no CPU path (use --impl reference for the CPU port)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    args.warmup = max(args.warmup, 3)

    # ---- workload -----------------------------------------------------------------------------
    g = synth.make_config(args.config)
    iters = g.iters
    N, ht, wd = g.disps.shape
    HW = ht * wd
    kx = sharded.frame_list(g.ii, g.t0, g.t1)
    owner = sharded.partition(g.ii, kx, world)
    sel = sharded.local_edges(g.ii, kx, owner, rank)
    E_total, E_local = g.n_edges, int(sel.sum())
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    host = dict(poses=pin(g.poses), disps=pin(g.disps), intr=pin(g.intrinsics), sens=pin(g.disps_sens),
                targets=pin(g.targets[sel]), weights=pin(g.weights[sel]), eta=pin(g.eta), ii=pin(g.ii[sel]),
                jj=pin(g.jj[sel]))
    d = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
    poses0, disps0 = d["poses"].clone(), d["disps"].clone()
    kx_t = torch.from_numpy(kx).to(dev)
    owned_t = torch.from_numpy((owner == rank).astype(np.uint8)).to(dev)
    ii_all, jj_all = torch.from_numpy(g.ii).to(dev), torch.from_numpy(g.jj).to(dev)
    droid_backends.RETURN_UPDATES = False  # GlORIE-SLAM discards [dx, dz]; skip the extra sync
    input_bytes = sum(v.numel() * v.element_size() for k, v in host.items())
    l2_flush = None
    if input_bytes < 400e6:  # inputs not much larger than the 126 MB L2: flush it between steps
        l2_flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def reset():
        d["poses"].copy_(poses0)
        d["disps"].copy_(disps0)

    def step():
        if world == 1:
            droid_backends.ba(d["poses"], d["disps"], d["intr"], d["sens"], d["targets"], d["weights"], d["eta"],
                              d["ii"], d["jj"], g.t0, g.t1, iters, g.lm, g.ep, False, False)
        else:
            sharded.ba_sharded(d["poses"], d["disps"], d["intr"], d["sens"], d["targets"], d["weights"], d["eta"],
                               d["ii"], d["jj"], g.t0, g.t1, iters, g.lm, g.ep, kx_t, owned_t, ii_all, jj_all)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(nsteps, fn):
        """sum of per-step CUDA-event times (ms) on the launch stream; state reset and the L2 flush
        stay outside the event pairs"""
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(nsteps)]
        barrier()
        for a, b in evs:
            reset()
            if l2_flush is not None:
                l2_flush.fill_(1)
            a.record()
            fn()
            b.record()
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # ---- warm-up, then the device-timed region ---------------------------------------------------
    for _ in range(args.warmup):
        reset()
        step()
    barrier()
    ctx = _gba.context(local_rank, N, max(E_local, E_total if world > 1 else 1, 1), ht, wd)
    info = _gba.query(ctx)
    launches_per_step = info.kernel_launches
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    total_ms = timed(args.steps, step)
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = total_ms / args.steps
    value = E_total * HW * iters / (ms_per_step * 1e-3)

    # ---- per-stage pass (events inside the library, on the launch stream) ------------------------
    _gba.set_timing(ctx, True)
    for _ in range(args.steps):
        reset()
        if l2_flush is not None:
            l2_flush.fill_(1)
        step()
    tm = _gba.get_timing(ctx)
    _gba.set_timing(ctx, False)
    stages = {s: {"ms_per_step": v[0] / args.steps, "launches_per_step": v[1] / args.steps} for s, v in tm.items()
              if v[1] > 0}
    pk, pk_src = peaks()
    K_local = int(owned_t.sum().item())
    # algorithmic bytes of ONE k_linearize launch on this rank: target + weight
    # 16 B per edge pixel, and per depth-keyframe pixel disparity 4 + eta 4 + sensor 4 + Q/w out 8
    alg_bytes = 16.0 * E_local * HW + 20.0 * K_local * HW
    lin_ms = stages.get("linearize", {}).get("ms_per_step", 0.0) / max(iters, 1)
    achieved = alg_bytes / (lin_ms * 1e-3) / 1e9 if lin_ms > 0 else 0.0
    roofline = {"bound": "hbm", "kernel": "k_linearize", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
                "frac": achieved / pk["hbm_gbs"], "traffic": None, "peak_source": pk_src,
                "algorithmic_bytes_per_launch": alg_bytes, "ms_per_launch": lin_ms}
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            roofline["traffic"] = json.load(f).get("k_linearize_%s_dram_bytes" % args.config)
    except Exception:
        pass

    # The Schur Gram (k_gram_tc) is the GEMM-shaped kernel of the path: G_k = A_k A_k^T per depth keyframe,
    # A_k = [6 deg_k + 1 rows] x [HW pixels].  Algorithmic flops = sum_k R_k (R_k + 1) HW (the symmetric
    # half, 2 flops per multiply-add); the tensor cores execute 3 tf32 MMAs of 128 x N_k x HW for it
    # (3xTF32 split, full square).  Peak: dense tf32 = half the measured dense bf16 rate.
    deg_local = np.bincount(np.searchsorted(kx, g.ii[sel]), minlength=len(kx))
    R = 6.0 * deg_local[deg_local > 0] + 1.0
    gram_ms = stages.get("gram", {}).get("ms_per_step", 0.0) / max(iters, 1)
    gram_alg = float((R * (R + 1.0)).sum()) * HW
    small = R <= 128
    gram_exec = float((3.0 * 2.0 * 128.0 * (np.ceil(R[small] / 16.0) * 16.0)).sum()) * HW
    tf32_peak = pk.get("bf16_tflops_sustained", pk.get("bf16_tflops", 0.0)) / 2.0
    roofline_gram = None
    if gram_ms > 0:
        roofline_gram = {"bound": "tensor", "kernel": "k_gram_tc (+ k_gram for out-degree > 21)", "unit": "TFLOP/s",
                         "achieved": gram_alg / (gram_ms * 1e-3) / 1e12, "peak": tf32_peak,
                         "frac": gram_alg / (gram_ms * 1e-3) / 1e12 / tf32_peak if tf32_peak else None,
                         "executed_tf32": gram_exec / (gram_ms * 1e-3) / 1e12, "ms_per_launch": gram_ms,
                         "note": "producer-bound: the CUDA cores rebuild the operand rows (about 170 instructions per "
                                 "edge pixel); the MMAs run underneath"}

    # The factorisation (k_chain, one launch, the largest single kernel) is bound by neither HBM nor a
    # throughput pipe but by its dependency chain: potrf -> trsm -> update per tile column.
    fac_ms = stages.get("factor", {}).get("ms_per_step", 0.0) / max(iters, 1)
    info2 = _gba.query(ctx)
    roofline_chain = {"bound": "dependency chain (fp64 tile Cholesky)", "kernel": "k_chain", "ms_per_launch": fac_ms,
                      "sequential_tile_columns": info2.chain_columns,
                      "us_per_column": 1e3 * fac_ms / max(info2.chain_columns, 1),
                      "note": "per column: potrf of a 64x64 fp64 tile (~19 us, two pivots per barrier) + one trsm "
                              "slab + one update + three flag hops; see DESIGN.md section 3"}

    # ---- end to end: pinned host buffers -> public API -> host, copies inside the timed region ----
    # Every step copies ITS inputs from pinned host memory and reads its result back.  The copies run
    # on a second stream into one of two device buffer sets, so the transfer of step s+1 overlaps the
    # compute of step s (ordinary input prefetch); nothing is reused between steps.
    e2e = None
    if not args.skip_e2e:
        names = ("poses", "disps", "intr", "sens", "targets", "weights", "eta", "ii", "jj")
        bufs = [d, {k: torch.empty_like(v) for k, v in d.items()}]
        outs = [(torch.empty_like(host["poses"]).pin_memory(), torch.empty_like(host["disps"]).pin_memory())
                for _ in range(2)]
        copy_stream = torch.cuda.Stream(device=dev)
        main = torch.cuda.current_stream(dev)

        def run_e2e(nsteps):
            ready = [torch.cuda.Event() for _ in range(nsteps)]
            done = [torch.cuda.Event() for _ in range(nsteps)]
            t0_, t1_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            t0_.record(main)
            copy_stream.wait_stream(main)
            for s_ in range(nsteps):
                bset = bufs[s_ % 2]
                with torch.cuda.stream(copy_stream):
                    if s_ >= 2:
                        copy_stream.wait_event(done[s_ - 2])  # that buffer set is free again
                    for k in names:
                        bset[k].copy_(host[k], non_blocking=True)
                    ready[s_].record(copy_stream)
                main.wait_event(ready[s_])
                if world == 1:
                    droid_backends.ba(bset["poses"], bset["disps"], bset["intr"], bset["sens"], bset["targets"],
                                      bset["weights"], bset["eta"], bset["ii"], bset["jj"], g.t0, g.t1, iters, g.lm,
                                      g.ep, False, False)
                else:
                    sharded.ba_sharded(bset["poses"], bset["disps"], bset["intr"], bset["sens"], bset["targets"],
                                       bset["weights"], bset["eta"], bset["ii"], bset["jj"], g.t0, g.t1, iters, g.lm,
                                       g.ep, kx_t, owned_t, ii_all, jj_all)
                outs[s_ % 2][0].copy_(bset["poses"], non_blocking=True)
                outs[s_ % 2][1].copy_(bset["disps"], non_blocking=True)
                done[s_].record(main)
            t1_.record(main)
            barrier()
            ms = t0_.elapsed_time(t1_)
            if world > 1:
                t = torch.tensor([ms], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            return ms / nsteps

        run_e2e(2)
        ne = max(args.steps, 4)
        e2e_ms = run_e2e(ne)
        e2e = {"value": E_total * HW * iters / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms,
               "h2d_bytes_per_step": int(input_bytes),
               "d2h_bytes_per_step": int(outs[0][0].numel() * 4 + outs[0][1].numel() * 4),
               "note": "H2D of step s+1 on a copy stream overlaps compute of step s (two device buffer sets)"}
        if world > 1:
            t = torch.tensor([e2e["h2d_bytes_per_step"]], dtype=torch.float64, device=dev)
            dist.all_reduce(t)
            e2e["h2d_bytes_per_step"] = int(t.item())

    # ---- layout hand-off, same workload, inputs resident ------------------
    # What GlORIE does today -- permute + contiguous copy of the update network's [E,ht,wd,2] tensors, then
    # ba -- against ba_interleaved on the tensors as they are.
    layout = None
    if world == 1 and not args.skip_cpu:
        tg_hwc = d["targets"].permute(0, 2, 3, 1).contiguous()
        wg_hwc = d["weights"].permute(0, 2, 3, 1).contiguous()

        def permute_then_ba():
            droid_backends.ba(d["poses"], d["disps"], d["intr"], d["sens"], tg_hwc.permute(0, 3, 1, 2).contiguous(),
                              wg_hwc.permute(0, 3, 1, 2).contiguous(), d["eta"], d["ii"], d["jj"], g.t0, g.t1, iters,
                              g.lm, g.ep, False, False)

        def interleaved():
            droid_backends.ba_interleaved(d["poses"], d["disps"], d["intr"], d["sens"], tg_hwc, wg_hwc, d["eta"], d["ii"],
                                          d["jj"], g.t0, g.t1, iters, g.lm, g.ep, False, False)

        layout = {}
        for key, fn in (("permute_then_ba_ms_per_step", permute_then_ba), ("ba_interleaved_ms_per_step", interleaved)):
            reset()
            fn()
            layout[key] = timed(args.steps, fn) / args.steps  # timed() resets the state outside the events
        del tg_hwc, wg_hwc

    cb = None
    ref_cuda = None
    if rank == 0 and world == 1 and not args.skip_cpu:
        cb = cpu_baseline(3)
        ref_cuda = reference_cuda_side_by_side(dev)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "ms_per_ba_iteration": ms_per_step / iters,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32 (fp64 reduced-system solve)",
            "data": "synthetic",
            "config": {"workload": "config %s: global BA, %d keyframes / %d edges / %dx%d, %d iterations per step"
                                   % (args.config, N, E_total, ht, wd, iters),
                       "lm": g.lm, "ep": g.ep, "sharding": "by source keyframe, %d rank(s)" % world,
                       "l2": "inputs (%.0f MB per rank) larger than L2" % (input_bytes / 1e6) if l2_flush is None
                             else "256 MB L2 flush between steps"},
            "roofline": roofline, "roofline_gram": roofline_gram, "roofline_chain": roofline_chain, "layout_handoff": layout, "stages": stages, "cpu_baseline": cb, "e2e": e2e, "reference_cuda": ref_cuda,
            "gpu_launches": int(launches_per_step * args.steps), "gpu_launches_per_step": int(launches_per_step),
            "clocks": clocks,
            "solver": {"tiles": info.n_tiles, "active_tiles": info.n_active_tiles, "chol_failures": info.chol_failures},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
