"""Edge-sharded bundle adjustment across the GPUs of one node (one process per GPU).

The factor graph shards by SOURCE keyframe: all edges i -> * live on
the rank that owns keyframe i, so the per-pixel depth diagonal C, w, the Schur Gram blocks and the
disparity update of keyframe i are rank-local.  The only exchange per Gauss-Newton iteration is
the sum of the packed reduced camera system (fp64 tiles of A - S and the rhs) over ranks --
`torch.distributed.all_reduce` on a zero-copy view of the library's buffer, NCCL over
NVLink/NVSwitch -- after which every rank factorises the identical system redundantly (no
further communication).  After the last iteration the owned disparity rows are merged.

The reference has no multi-GPU path; this is new capability behind the
same `ba(...)` argument list.  Only the droid_cuda semantics are supported here (the geom_py
damping acts on partial sums).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist

import _gba


# optional device timing of the two exchanges (bench.py): {"allreduce": [(ev0, ev1), ...], "merge": [...]}
TIMERS = None


def _timed(name, fn):
    if TIMERS is None:
        return fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    out = fn()
    b.record()
    TIMERS.setdefault(name, []).append((a, b))
    return out


def frame_list(ii, t0, t1):
    """Sorted depth keyframes kx = unique([t0,t1) U ii) (reference droid_kernels.cu:1337-1345)."""
    ii = np.asarray(ii, dtype=np.int64)
    return np.unique(np.concatenate([np.arange(t0, t1, dtype=np.int64), ii]))


def partition(ii, kx, world):
    """Greedy longest-processing-time assignment of depth keyframes to ranks by out-degree.
    Returns owner[K] (int32).  Keyframes without out-edges are dealt round-robin."""
    ii = np.asarray(ii, dtype=np.int64)
    kx = np.asarray(kx, dtype=np.int64)
    pos = np.searchsorted(kx, ii)
    deg = np.bincount(pos, minlength=len(kx)).astype(np.int64)
    owner = np.zeros(len(kx), dtype=np.int32)
    load = np.zeros(world, dtype=np.int64)
    order = np.argsort(-deg, kind="stable")
    rr = 0
    for k in order:
        if deg[k] == 0:
            owner[k] = rr % world
            rr += 1
            continue
        r = int(np.argmin(load))
        owner[k] = r
        load[r] += deg[k]
    return owner


def local_edges(ii, kx, owner, rank):
    """Boolean mask of the edges whose source keyframe `rank` owns."""
    pos = np.searchsorted(np.asarray(kx), np.asarray(ii))
    return owner[pos] == rank


def merge_owned_rows(disps, kx_t, owned_t, group=None):
    """Every rank updated only the disparity rows it owns; sum the masked rows so all ranks end
    with the full state.  disps [N,ht,wd], kx_t [K] int64, owned_t [K] (same device)."""
    rows = disps[kx_t] * owned_t.to(disps.dtype)[:, None, None]
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(rows, op=dist.ReduceOp.SUM, group=group)
    disps[kx_t] = rows
    return disps


def allreduce_system(view, group=None):
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(view, op=dist.ReduceOp.SUM, group=group)
    return view


def ba_sharded(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep,
               kx_t, owned_t, ii_all, jj_all, group=None, merge=True):
    """`droid_backends.ba` for one rank of an edge-sharded run.

    poses / disps / intrinsics / disps_sens / eta: the full (replicated) tensors on this rank's GPU;
    targets / weights / ii / jj: THIS RANK's edges only; kx_t [K] int64 = frame_list(global ii),
    owned_t [K] uint8 = (owner == rank), ii_all / jj_all: the edge indices of the WHOLE graph
    (replicated; the pose ordering and the tile layout of the reduced system are derived from them
    so that all ranks pack the system identically).  poses and the owned rows of disps are updated in place;
    with merge=True all rows of disps are consistent on return.  Returns the kernel-launch count."""
    dev = poses.device
    N, ht, wd = disps.shape
    E = int(ii.shape[0])
    L = _gba.lib()
    with torch.cuda.device(dev):
        ctx = _gba.context(dev.index, N, max(E, int(ii_all.shape[0]), 1), ht, wd)
        stream = torch.cuda.current_stream(dev).cuda_stream
        nul = lambda t: t.data_ptr() if t.numel() else None
        _gba.check(ctx, L.gba_ba_begin(ctx, poses.data_ptr(), disps.data_ptr(), intrinsics.data_ptr(),
                                       disps_sens.data_ptr(), nul(targets), nul(weights), eta.data_ptr(), nul(ii),
                                       nul(jj), N, E, int(eta.shape[0]), int(t0), int(t1), float(lm), float(ep), 0, 0,
                                       _gba.MODE_DROID_CUDA, kx_t.data_ptr(), int(kx_t.shape[0]), owned_t.data_ptr(),
                                       ii_all.data_ptr(), jj_all.data_ptr(), int(ii_all.shape[0]), stream))
        view = _gba.system_view(ctx, dev)
        for _ in range(int(iterations)):
            _gba.check(ctx, L.gba_ba_build(ctx))
            _timed("allreduce", lambda: allreduce_system(view, group))
            _gba.check(ctx, L.gba_ba_solve(ctx))
        _gba.check(ctx, L.gba_ba_end(ctx, None, None))
        if merge:
            _timed("merge", lambda: merge_owned_rows(disps, kx_t, owned_t, group))
        # every rank factorises the same summed system, but fp64 atomics inside the factorisation are unordered
        # and a failed factorisation (zero update) is rank-local: rank 0's poses are the run's poses
        if dist.is_initialized() and dist.get_world_size(group) > 1:
            _timed("merge", lambda: dist.broadcast(poses[int(t0):int(t1)], src=dist.get_global_rank(group, 0) if group is not None else 0, group=group))
    return ctx
