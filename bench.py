#!/usr/bin/env python
"""bench.py -- dense BA hot path on B200: edge-pixel residuals/s and ms per BA iteration.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config E] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

A "step" is one `droid_backends.ba` call (`iters` Gauss-Newton iterations, BASELINE.json's
backend setting 2) over one synthetic factor graph.  Default workload: config E (ScanNet-scale
global BA, 1000 keyframes / 20000 edges / 48x64), the configuration BASELINE.json's metric and
target are quoted on; it fits one GPU.  At N > 1 the SAME graph is edge-sharded by source keyframe
over the ranks (strong scaling): every rank owns the keyframes of its band segments of the ordered
reduced system, eliminates those segments itself, and one NCCL all-reduce per iteration sums the
closure block + rhs (GBA_REPLICATED_SOLVE=1: the older scheme, whole system all-reduced, solve replicated).

Printed JSON (one line, rank 0): the driver contract plus
  roofline            the dominant HBM kernel: algorithmic bytes of one launch / its CUDA-event time
  roofline_backsub    the second streaming kernel (k_backsub, weights only: 8 B per edge pixel), same form
  roofline_iteration  SURVEY 8(d): B_alg = 16 E HW + 16 K HW per iteration over the SUM of the HBM-streaming
                      stages of an iteration (every pass that re-reads the edge inputs counts against it)
  roofline_gram / roofline_chain   the tensor-core Gram and the fp64 tile Cholesky
  stages              device ms per step of every stage (events on the launch stream, separate pass);
                      at N > 1 also the all-reduce and the disparity merge
  dspo                BASELINE.json configs[3]: one BA_with_scale_shift iteration at config D (ms, HBM fraction
                      of 16 E HW + 24 M HW) next to the reference's own PyTorch implementation on the host cores
  cpu_baseline        the reference's own src/geom/ba.py::BA (staged copy, oracle/_ref/pyref) on the host cores on
                      configs A and B, plus the C/OpenMP port of ba_cuda on config C
  e2e                 same metric through the public API with pinned HOST buffers, copies timed
  parity_n_ranks      N > 1: the N-rank state against the 1-rank state on the same graph (untimed)
`--impl reference` times the reference's PyTorch BA on the host cores (bounded sample: config B).
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

# torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU arms time the host cores they report
_OMP_ENV = os.environ.pop("OMP_NUM_THREADS", None)

import numpy as np  # noqa: E402

METRIC = "edge-pixel residuals/sec (global BA; ms per BA iteration in ms_per_ba_iteration)"
UNIT = "edge-pixels/s"
# bounded sample of the C/OpenMP port = config C (backend global BA, 200 keyframes / 3000 edges)
CPU_SAMPLE = dict(n=200, e=3000, r=7, iters=2, lm=1e-5, ep=1e-2, seed=0)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback (B200_PROFILING.md)"


def traffic_of(kernel, config, world, what="dram_bytes"):
    """DRAM bytes per launch from an ncu --set full capture of THIS kernel, config and GPU count
    (profiles/traffic.json), else None -- never a number measured on another configuration."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get("%s_%s_%dgpu_%s" % (kernel, config, world, what))
    except Exception:
        return None


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


# ---------------------------------------------------------------------------------------------------
# CPU arms (rank 0 only).  The one place besides tests/ and smoke() that may execute oracle/.
# ---------------------------------------------------------------------------------------------------
def host_cores():
    return os.cpu_count() or 1


def cpu_port(steps=1):
    """C/OpenMP port of the reference's ba_cuda (oracle/ba_oracle.c) on config C, all host cores."""
    import synth
    from oracle import oracle

    cores = oracle.set_threads(host_cores())
    g = synth.make_graph(**CPU_SAMPLE)
    best = None
    for _ in range(max(steps, 1)):
        t = time.perf_counter()
        oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  g.iters, g.lm, g.ep, precision="f32")
        dt = time.perf_counter() - t
        best = dt if best is None else min(best, dt)
    return {
        "value": g.n_edges * g.hw * g.iters / best, "unit": UNIT, "cores": cores, "kind": "port",
        "sample": "C/OpenMP port of ba_cuda: config C, %d keyframes / %d edges / %dx%d / %d iterations, best of %d "
                  "(%.2f s each)" % (g.n_frames, g.n_edges, g.disps.shape[1], g.disps.shape[2], g.iters, max(steps, 1), best),
        "ms_per_ba_iteration": 1e3 * best / g.iters,
    }


def cpu_reference(names=("A", "B"), reps=3):
    """The reference's own PyTorch BA, src/geom/ba.py::BA (:33-121), imported from the staged copy under
    oracle/_ref/pyref (oracle/build_ref.py stage_python) and timed on the host cores in fp32.  Configs A
    (BASELINE.json configs[0], 4 iterations) and B; C needs ~15 GB and 10-20 s per iteration for its dense
    [P,M,6,HW] E tensor, E (73 GB) cannot run at all -- stated, not extrapolated."""
    import torch

    import synth
    from oracle import ref_harness

    if not ref_harness.available():
        return None
    torch.set_num_threads(host_cores())
    out = {}
    for name in names:
        g = synth.make_config(name)
        ref_harness.run_BA(g, iters=1)  # warm-up
        best = None
        for _ in range(reps):
            t = time.perf_counter()
            ref_harness.run_BA(g)
            dt = time.perf_counter() - t
            best = dt if best is None else min(best, dt)
        out[name] = {"value": g.n_edges * g.hw * g.iters / best, "ms_per_ba_iteration": 1e3 * best / g.iters,
                     "sample": "reference src/geom/ba.py::BA fp32: config %s, %d keyframes / %d edges / %dx%d / %d "
                               "iterations, best of %d (%.2f s each)" % (name, g.n_frames, g.n_edges, g.disps.shape[1],
                                                                        g.disps.shape[2], g.iters, reps, best)}
    return {"cores": torch.get_num_threads(), "configs": out}


def cpu_baseline(reps=3):
    """`cpu_baseline` object of the JSON line: the reference itself where it can run (kind "reference"),
    the port beside it; the port alone (kind "port") when the staged reference is missing."""
    port = cpu_port(reps)
    ref = None
    try:
        ref = cpu_reference(reps=reps)
    except Exception as e:  # noqa: BLE001 -- the staged reference is optional on a box without oracle/_ref/pyref
        port["reference_unavailable"] = repr(e)[:200]
    if ref is None:
        return port
    b = ref["configs"]["B"]
    return {"value": b["value"], "unit": UNIT, "cores": ref["cores"], "kind": "reference", "sample": b["sample"],
            "ms_per_ba_iteration": b["ms_per_ba_iteration"], "config_A": ref["configs"]["A"], "port_config_C": port,
            "note": "configs C / E do not fit the reference's dense formulation (E tensor 15 GB / 73 GB)"}


def run_reference(args, rank):
    """--impl reference: the reference's CPU implementation of the path on the host cores.  Rank 0 only."""
    if rank != 0:
        return
    cb = cpu_baseline(max(1, min(args.steps, 5)))
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": cb["ms_per_ba_iteration"] * 2,
        "ms_per_ba_iteration": cb["ms_per_ba_iteration"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "config %s (global BA); timed on the bounded CPU sample named in cpu_baseline.sample "
                               "(edge-pixels/s is size-normalised)" % args.config},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------
# side measurements on the GPU (rank 0, N = 1)
# ---------------------------------------------------------------------------------------------------
def reference_cuda_side_by_side(dev, name="C", reps=5):
    """Same-box GPU baseline: the reference's own CUDA droid_backends.ba (oracle/_ref, rebuilt
    unmodified for sm_100 by oracle/build_ref.py) against libgba on config C, device resident,
    CUDA events, best of `reps`."""
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

        def best_of(fn):
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
            return best

        out = {}
        for label, fn in (("reference", ref.ba), ("libgba", droid_backends.ba)):
            out[label + "_ms_per_ba_iteration"] = best_of(
                lambda: fn(base[0].clone(), base[1].clone(), *rest, g.t0, g.t1, g.iters, g.lm, g.ep, False, False)) / g.iters
        out["config"] = "config %s: %d keyframes / %d edges, %d iterations, device resident" % (
            name, g.n_frames, g.n_edges, g.iters)
        out["speedup"] = out["reference_ms_per_ba_iteration"] / out["libgba_ms_per_ba_iteration"]
        n = g.n_frames
        gi, gj = torch.meshgrid(torch.arange(n, device=dev), torch.arange(n, device=dev), indexing="ij")
        gi, gj = gi.reshape(-1).contiguous(), gj.reshape(-1).contiguous()
        for label, fn in (("reference", ref.frame_distance), ("libgba", droid_backends.frame_distance)):
            out[label + "_frame_distance_ms"] = best_of(lambda: fn(base[0], base[1], rest[0], gi, gj, 0.3))
        out["frame_distance_pairs"] = int(gi.numel())
        ix = torch.arange(n, device=dev)
        th = (0.05 / base[1].mean(dim=(1, 2))).contiguous()
        out["reference_depth_filter_ms"] = best_of(lambda: ref.depth_filter(base[0], base[1], rest[0], ix, th))
        out["libgba_depth_filter_ms"] = best_of(lambda: droid_backends.depth_filter(base[0], base[1], rest[0], ix, th))
        return out
    except Exception as e:  # noqa: BLE001 -- the prebuilt reference is optional evidence
        return {"unavailable": repr(e)[:200]}


def dspo_measurement(dev, steps, pk, with_cpu):
    """BASELINE.json configs[3]: DSPO stage 2 (BA_with_scale_shift, src/geom/ba.py:127-216) at config D."""
    import ctypes

    import torch

    import _gba
    import geom_ba
    import synth

    g = synth.make_config("D")
    N, ht, wd = g.disps.shape
    HW, E = ht * wd, g.n_edges
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    tg, wg = t(g.targets.transpose(0, 2, 3, 1)), t(g.weights.transpose(0, 2, 3, 1))
    M = int(len(np.unique(g.ii)))
    eta = t(g.eta[:M])
    poses, disps0, intr = t(g.poses), t(g.disps), t(g.intrinsics)
    mono, mask = t(g.mono_disps), t(g.valid_depth_mask.astype(np.uint8))
    sc0, sh0 = t(g.scales), t(g.shifts)
    ii, jj = t(g.ii), t(g.jj)
    ctx = _gba.context(dev.index, N, E, ht, wd)
    L = _gba.lib()
    stream = torch.cuda.current_stream(dev).cuda_stream
    disps, sc, sh = disps0.clone(), sc0.clone(), sh0.clone()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # 147 MB of inputs ~ L2: flush between steps

    def call():
        _gba.check(ctx, L.gba_dspo(ctx, poses.data_ptr(), disps.data_ptr(), intr.data_ptr(), tg.data_ptr(), wg.data_ptr(),
                                   eta.data_ptr(), ii.data_ptr(), jj.data_ptr(), mono.data_ptr(), sc.data_ptr(),
                                   sh.data_ptr(), mask.data_ptr(), N, E, M, 0, g.lm, g.ep, g.alpha, 1, stream))

    for _ in range(3):
        disps.copy_(disps0)
        call()
    torch.cuda.synchronize()
    _gba.set_timing(ctx, True)
    evs = []
    for _ in range(steps):
        disps.copy_(disps0); sc.copy_(sc0); sh.copy_(sh0)
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        call()
        b.record()
        evs.append((a, b))
    tm = _gba.get_timing(ctx)
    _gba.set_timing(ctx, False)
    call_ms = sum(a.elapsed_time(b) for a, b in evs) / steps  # incl. index build + the status read-back
    kern_ms = tm["dspo"][0] / steps
    alg = 16.0 * E * HW + 24.0 * M * HW
    # through the reference-facing mirror (functional: clones its outputs)
    args = (tg[None], wg[None], eta, poses[None], disps0[None], intr[None, None].repeat(1, N, 1), ii, jj, mono[None],
            sc0[None], sh0[None], t(g.valid_depth_mask)[None], 0, g.lm, g.ep)
    geom_ba.BA_with_scale_shift(*args, alpha=g.alpha)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        geom_ba.BA_with_scale_shift(*args, alpha=g.alpha)
    torch.cuda.synchronize()
    mirror_ms = 1e3 * (time.perf_counter() - t0) / steps
    out = {"config": "config D: DSPO stage 2, %d keyframes / %d edges / %dx%d, one iteration" % (N, E, ht, wd),
           "ms_per_iteration_kernels": kern_ms, "ms_per_call_c_abi": call_ms, "ms_per_call_python_mirror": mirror_ms,
           "edge_pixels_per_s": E * HW / (kern_ms * 1e-3), "launches_per_call": tm["dspo"][1] / steps + tm["index"][1] / steps,
           "roofline": {"bound": "hbm", "algorithmic_bytes": alg, "achieved": alg / (kern_ms * 1e-3) / 1e9,
                        "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": alg / (kern_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
                        "traffic": traffic_of("dspo", "D", 1), "formula": "16 E HW + 24 M HW (SURVEY 8(d))",
                        "l2": "256 MB L2 flush between steps"}}
    if with_cpu:
        try:
            import torch as _t

            from oracle import ref_harness

            if ref_harness.available():
                _t.set_num_threads(host_cores())
                t0 = time.perf_counter()
                ref_harness.run_DSPO(g, iters=1)
                dt = time.perf_counter() - t0
                out["cpu_reference"] = {"ms_per_iteration": 1e3 * dt, "edge_pixels_per_s": E * HW / dt,
                                        "cores": _t.get_num_threads(), "kind": "reference",
                                        "sample": "reference src/geom/ba.py::BA_with_scale_shift fp32, config D, one iteration"}
        except Exception as e:  # noqa: BLE001
            out["cpu_reference"] = {"unavailable": repr(e)[:200]}
    return out


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
        raise SystemExit("bench.py needs a CUDA device: libgba has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # stdout carries the one JSON line only
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    args.warmup = max(args.warmup, 3)

    # ---- workload -----------------------------------------------------------------------------
    g = synth.make_config(args.config)
    iters = g.iters
    N, ht, wd = g.disps.shape
    HW = ht * wd
    # ownership: by band segment of the ordered reduced system when the ordering offers enough segments (then only
    # the closure block crosses ranks), else greedy by out-degree with the whole system all-reduced
    plan = None
    if world > 1 and not os.environ.get("GBA_REPLICATED_SOLVE"):
        kx, plan = sharded.plan(g.ii, g.jj, g.t0, g.t1, N, ht, wd, world, dev)
    else:
        kx = sharded.frame_list(g.ii, g.t0, g.t1)
    owner = plan.owner if plan is not None else sharded.partition(g.ii, kx, world)
    segments = plan.seg_of_rank[rank] if plan is not None else None
    pose_owner_t = torch.from_numpy(plan.pose_owner).to(dev) if plan is not None else None
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
    input_bytes = sum(v.numel() * v.element_size() for k, v in host.items())
    l2_flush = None
    if input_bytes < 400e6:  # inputs not much larger than the 126 MB L2: flush it between steps
        l2_flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def reset():
        d["poses"].copy_(poses0)
        d["disps"].copy_(disps0)

    def step():
        if world == 1:  # the drop-in's default behaviour: in-place update, [dx, dz] returned
            droid_backends.ba(d["poses"], d["disps"], d["intr"], d["sens"], d["targets"], d["weights"], d["eta"],
                              d["ii"], d["jj"], g.t0, g.t1, iters, g.lm, g.ep, False, False)
        else:
            sharded.ba_sharded(d["poses"], d["disps"], d["intr"], d["sens"], d["targets"], d["weights"], d["eta"],
                               d["ii"], d["jj"], g.t0, g.t1, iters, g.lm, g.ep, kx_t, owned_t, ii_all, jj_all,
                               segments=segments, pose_owner_t=pose_owner_t)

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
    sharded.TIMERS = {} if world > 1 else None
    for _ in range(args.steps):
        reset()
        if l2_flush is not None:
            l2_flush.fill_(1)
        step()
    tm = _gba.get_timing(ctx)
    _gba.set_timing(ctx, False)
    stages = {s: {"ms_per_step": v[0] / args.steps, "launches_per_step": v[1] / args.steps} for s, v in tm.items()
              if v[1] > 0}
    if world > 1:
        torch.cuda.synchronize()
        for name, evs in sharded.TIMERS.items():
            stages[name] = {"ms_per_step": sum(a.elapsed_time(b) for a, b in evs) / args.steps,
                            "calls_per_step": len(evs) / args.steps}
        sharded.TIMERS = None
    pk, pk_src = peaks()
    K_local = int(owned_t.sum().item())
    st_ms = lambda s: stages.get(s, {}).get("ms_per_step", 0.0) / max(iters, 1)
    # algorithmic bytes of ONE launch of the HBM kernel on this rank: target + weight
    # 16 B per edge pixel, and per depth-keyframe pixel disparity 4 + eta 4 + sensor 4 + Q/w out 8
    alg_bytes = 16.0 * E_local * HW + 20.0 * K_local * HW
    lin_ms = st_ms("linearize")
    achieved = alg_bytes / (lin_ms * 1e-3) / 1e9 if lin_ms > 0 else 0.0
    roofline = {"bound": "hbm", "kernel": "k_linearize", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s",
                "frac": achieved / pk["hbm_gbs"], "traffic": traffic_of("k_linearize", args.config, world),
                "peak_source": pk_src, "algorithmic_bytes_per_launch": alg_bytes, "ms_per_launch": lin_ms}
    # the second streaming kernel: k_backsub reads the weights only (8 B per edge pixel: E . delta carries no residual)
    # and per depth-keyframe pixel disparity 4 + Q/W 8 in, disparity 4 + dz 4 out
    bs_bytes = 8.0 * E_local * HW + 20.0 * K_local * HW
    bs_ms = st_ms("backsub")
    roofline_backsub = {"bound": "hbm", "kernel": "k_backsub", "unit": "GB/s", "peak": pk["hbm_gbs"],
                        "achieved": bs_bytes / (bs_ms * 1e-3) / 1e9 if bs_ms > 0 else 0.0,
                        "frac": bs_bytes / (bs_ms * 1e-3) / 1e9 / pk["hbm_gbs"] if bs_ms > 0 else 0.0,
                        "traffic": traffic_of("k_backsub", args.config, world),
                        "algorithmic_bytes_per_launch": bs_bytes, "ms_per_launch": bs_ms}
    # SURVEY 8(d) per iteration: B_alg = 16 E HW + 16 K HW once, against EVERY stage that streams the edge inputs
    b_alg = 16.0 * E_local * HW + 16.0 * K_local * HW
    hbm_ms = st_ms("linearize") + st_ms("gram") + st_ms("backsub")
    roofline_iteration = {
        "bound": "hbm", "unit": "GB/s", "peak": pk["hbm_gbs"], "B_alg_bytes_per_iteration": b_alg,
        "hbm_stage_ms_per_iteration": hbm_ms, "stages_counted": ["linearize", "gram", "backsub"],
        "achieved": b_alg / (hbm_ms * 1e-3) / 1e9 if hbm_ms > 0 else 0.0,
        "frac": b_alg / (hbm_ms * 1e-3) / 1e9 / pk["hbm_gbs"] if hbm_ms > 0 else 0.0,
        "frac_linearize_plus_backsub": (b_alg / ((st_ms("linearize") + st_ms("backsub")) * 1e-3) / 1e9 / pk["hbm_gbs"])
        if st_ms("linearize") + st_ms("backsub") > 0 else 0.0,
        "traffic": traffic_of("iteration", args.config, world),
        "note": "the edge inputs are streamed by k_linearize (targets + weights), the Gram and k_backsub (weights "
                "only): 32 B per edge pixel against the 16 B of B_alg"}

    # The Schur Gram (k_gram_tc) is the GEMM-shaped kernel of the path: G_k = A_k A_k^T per depth keyframe,
    # A_k = [6 deg_k + 1 rows] x [HW pixels].  Algorithmic flops = sum_k R_k (R_k + 1) HW (the symmetric
    # half, 2 flops per multiply-add); the tensor cores execute 3 tf32 MMAs of 128 x N_k x HW for it
    # (3xTF32 split, full square).  Peak: dense tf32 = half the measured dense bf16 rate.
    deg_local = np.bincount(np.searchsorted(kx, g.ii[sel]), minlength=len(kx))
    R = 6.0 * deg_local[deg_local > 0] + 1.0
    gram_ms = st_ms("gram")
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
                         "traffic": traffic_of("k_gram_tc", args.config, world)}

    # The factorisation: fp64 tile Cholesky of the ordered reduced camera system.  Flops = what the tile tasks
    # execute (64^3 multiply-adds per tile product, 64^3 / 3 per diagonal tile), from the symbolic structure.
    fac_ms, sol_ms = st_ms("factor"), st_ms("solve")
    info2 = _gba.query(ctx)
    tmask = np.zeros((info2.n_tiles, info2.n_tiles), dtype=np.uint8)
    import ctypes
    _gba.check(ctx, _gba.lib().gba_get(ctx, _gba.GET_TILEMASK, tmask.ctypes.data_as(ctypes.c_void_p), tmask.nbytes))
    col = np.tril(tmask, -1).sum(axis=0).astype(np.float64)  # active tiles below the diagonal, per tile column
    nb3 = 64.0 ** 3
    chol_flops = 2.0 * nb3 * float((1.0 / 3.0 + col + col * (col + 1) / 2).sum())
    fp64_peak = 2.0 * 64 * 148 * (clocks["sm_mhz"] if clocks and clocks.get("sm_mhz") else 1965.0) * 1e6 / 1e12
    roofline_chain = {"bound": "dependency chain (fp64 tile Cholesky on DMMA)", "kernel": "k_chain",
                      "ms_per_launch": fac_ms, "solve_ms": sol_ms, "sequential_tile_columns": info2.chain_columns,
                      "us_per_column": 1e3 * fac_ms / max(info2.chain_columns, 1), "fp64_flops_executed": chol_flops,
                      "achieved": chol_flops / (fac_ms * 1e-3) / 1e12 if fac_ms > 0 else 0.0, "unit": "TFLOP/s",
                      "peak": fp64_peak, "peak_source": "64 fp64 FMA/clk/SM x 148 SMs x measured SM clock",
                      "frac": chol_flops / (fac_ms * 1e-3) / 1e12 / fp64_peak if fac_ms > 0 else 0.0,
                      "tensor_pipe_utilisation_pct": traffic_of("k_chain", args.config, world, "tensor_pipe_pct"),
                      "note": "fp64 tile products on the tensor pipe (mma.sync.m8n8k4.f64 = SASS DMMA), whose fp64 rate on "
                              "B200 equals the DFMA pipe's (64 FMA/clk/SM, profiles/dev); sm__pipe_tensor_cycles_active "
                              "from the ncu capture named in profiles/traffic.json, null for other configs"}

    # ---- N-rank state against the 1-rank state (untimed) -------------------------------------------
    parity_n = None
    if world > 1:
        reset()
        step()
        torch.cuda.synchronize()
        pn, dn = d["poses"].clone(), d["disps"].clone()
        if rank == 0:
            full = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in
                    dict(targets=g.targets, weights=g.weights, ii=g.ii, jj=g.jj).items()}
            p1, d1 = poses0.clone(), disps0.clone()
            c1 = _gba.lib()
            h1 = ctypes.c_void_p()
            assert c1.gba_create(ctypes.byref(h1), local_rank, N, E_total, ht, wd) == 0
            _gba.check(h1, c1.gba_ba(h1, p1.data_ptr(), d1.data_ptr(), d["intr"].data_ptr(), d["sens"].data_ptr(),
                                     full["targets"].data_ptr(), full["weights"].data_ptr(), d["eta"].data_ptr(),
                                     full["ii"].data_ptr(), full["jj"].data_ptr(), N, E_total, int(d["eta"].shape[0]),
                                     g.t0, g.t1, iters, g.lm, g.ep, 0, 0, 0, None, None,
                                     torch.cuda.current_stream(dev).cuda_stream))
            torch.cuda.synchronize()
            c1.gba_destroy(h1)
            relf = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm())
            parity_n = {"poses_rel": relf(pn, p1), "disps_rel": relf(dn, d1), "bar": 1e-5,
                        "what": "%d-rank NCCL result vs the 1-rank result on rank 0's GPU, same graph, %d iterations" % (world, iters)}
            parity_n["ok"] = parity_n["poses_rel"] < 1e-5 and parity_n["disps_rel"] < 1e-5
            del full
        dist.barrier()

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
                                       g.ep, kx_t, owned_t, ii_all, jj_all, segments=segments, pose_owner_t=pose_owner_t)
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

    cb = ref_cuda = dspo = None
    if rank == 0 and world == 1 and not args.skip_cpu:
        health = droid_backends.check_health(dev)  # raises if a kernel-internal watchdog fired during the run
        del d, host
        torch.cuda.empty_cache()
        dspo = dspo_measurement(dev, args.steps, pk, True)
        ref_cuda = reference_cuda_side_by_side(dev)
        cb = cpu_baseline(3)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "ms_per_ba_iteration": ms_per_step / iters,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32 (fp64 reduced-system solve)",
            "data": "synthetic",
            "config": {"workload": "config %s: global BA, %d keyframes / %d edges / %dx%d, %d iterations per step"
                                   % (args.config, N, E_total, ht, wd, iters),
                       "lm": g.lm, "ep": g.ep, "sharding": "by source keyframe, %d rank(s)" % world,
                       "solve": ("one rank" if world == 1 else
                                 "segmented: own band segments per rank, closure block all-reduced (%d segments)" % plan.n_segments
                                 if plan is not None else "replicated after an all-reduce of the whole system"),
                       "l2": "inputs (%.0f MB per rank) larger than L2" % (input_bytes / 1e6) if l2_flush is None
                             else "256 MB L2 flush between steps"},
            "roofline": roofline, "roofline_backsub": roofline_backsub, "roofline_iteration": roofline_iteration,
            "roofline_gram": roofline_gram,
            "roofline_chain": roofline_chain, "layout_handoff": layout, "stages": stages, "dspo": dspo,
            "cpu_baseline": cb, "e2e": e2e, "reference_cuda": ref_cuda, "parity_n_ranks": parity_n,
            "gpu_launches": int(launches_per_step * args.steps), "gpu_launches_per_step": int(launches_per_step),
            "clocks": clocks,
            "solver": {"tiles": info.n_tiles, "active_tiles": info.n_active_tiles, "chol_failures": info.chol_failures,
                       "watchdog_trips": info.watchdog_trips},
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
