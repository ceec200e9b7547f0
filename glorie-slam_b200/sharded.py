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


class Plan:
    """Ownership of the depth keyframes for a SEGMENTED sharded solve (see `plan`)."""

    def __init__(self, owner, seg_of_rank, pose_owner, n_segments):
        self.owner = owner              # [K] int32: rank that owns kx[k]
        self.seg_of_rank = seg_of_rank  # [world] (seg_lo, seg_n)
        self.pose_owner = pose_owner    # [t1 - t0] int32: rank whose solve moves that pose (closure poses: 0)
        self.n_segments = n_segments


def plan_from_perm(ii, jj, kx, t0, t1, perm, n_segments, segment_tiles, world):
    """Host half of `plan`: given the pose -> position map of the ordered reduced system, assign every depth
    keyframe to the rank that owns the ONE band segment its pose blocks fall into (its own pose and the
    destinations of its out-edges; closure poses are shared and do not count).  Returns None when some keyframe
    touches two segments or the segments do not divide among the ranks -- the caller then keeps the replicated solve."""
    ii, jj, kx, perm = (np.asarray(a, dtype=np.int64) for a in (ii, jj, kx, perm))
    if n_segments < 2 or n_segments % world != 0:
        return None
    P = t1 - t0
    seg_pos = segment_tiles * 64 // 6            # pose positions per segment (tile aligned: 32 poses = 3 tiles)
    band_end = n_segments * seg_pos
    seg_of_pose = np.where(perm < band_end, perm // seg_pos, -1)  # -1: closure block
    def pose_seg(frames):
        pr = frames - t0
        ok = (pr >= 0) & (pr < P)
        out = np.full(frames.shape, -1, dtype=np.int64)
        out[ok] = seg_of_pose[pr[ok]]
        return out
    lo = np.full(len(kx), n_segments, dtype=np.int64)
    hi = np.full(len(kx), -1, dtype=np.int64)
    k_of_edge = np.searchsorted(kx, ii)
    for segs, ks in ((pose_seg(jj), k_of_edge), (pose_seg(kx), np.arange(len(kx)))):
        m = segs >= 0
        np.minimum.at(lo, ks[m], segs[m])
        np.maximum.at(hi, ks[m], segs[m])
    if np.any((hi >= 0) & (lo != hi)):
        return None
    seg = np.where(hi >= 0, hi, -1)
    # keyframes that touch closure poses only: follow the nearest keyframe in time that has a segment
    if np.any(seg < 0):
        known = np.where(seg >= 0)[0]
        if len(known) == 0:
            return None
        nearest = known[np.abs(kx[known][None, :] - kx[:, None]).argmin(axis=1)]
        seg = np.where(seg >= 0, seg, seg[nearest])
    per = n_segments // world
    owner = (seg // per).astype(np.int32)
    pose_owner = np.where(seg_of_pose >= 0, seg_of_pose // per, 0).astype(np.int32)
    return Plan(owner, [(r * per, per) for r in range(world)], pose_owner, n_segments)


def plan(ii, jj, t0, t1, n_frames, ht, wd, world, device):
    """Decide which rank owns which source keyframe so that only the CLOSURE block of the reduced system is
    shared between ranks: runs the library's ordering on the edge indices (gba_plan), then `plan_from_perm`.
    Every rank computes the same plan from the same (replicated) edge list.  Returns (kx, Plan or None)."""
    import ctypes

    ii_t = torch.as_tensor(np.ascontiguousarray(ii), dtype=torch.int64, device=device)
    jj_t = torch.as_tensor(np.ascontiguousarray(jj), dtype=torch.int64, device=device)
    kx = frame_list(ii, t0, t1)
    L = _gba.lib()
    with torch.cuda.device(device):
        ctx = _gba.context(torch.device(device).index, n_frames, int(ii_t.shape[0]), ht, wd)
        _gba.check(ctx, L.gba_plan(ctx, ii_t.data_ptr(), jj_t.data_ptr(), int(ii_t.shape[0]), int(n_frames), int(t0),
                                   int(t1), torch.cuda.current_stream(device).cuda_stream))
        info = _gba.query(ctx)
        perm = np.zeros(t1 - t0, dtype=np.int32)
        _gba.check(ctx, L.gba_get(ctx, _gba.GET_PERM, perm.ctypes.data_as(ctypes.c_void_p), perm.nbytes))
    return kx, plan_from_perm(ii, jj, kx, t0, t1, perm, info.n_segments, info.segment_tiles, world)


def merge_poses(poses, t0, t1, pose_owner_t, rank, group=None):
    """Segmented solve: every rank moved the closure poses and the poses of its own band segments; sum the
    masked rows so that all ranks end with the full trajectory (closure poses are taken from rank 0)."""
    rows = poses[t0:t1] * (pose_owner_t == rank).to(poses.dtype)[:, None]
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(rows, op=dist.ReduceOp.SUM, group=group)
    poses[t0:t1] = rows
    return poses


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
               kx_t, owned_t, ii_all, jj_all, group=None, merge=True, segments=None, pose_owner_t=None):
    """`droid_backends.ba` for one rank of an edge-sharded run.

    poses / disps / intrinsics / disps_sens / eta: the full (replicated) tensors on this rank's GPU;
    targets / weights / ii / jj: THIS RANK's edges only; kx_t [K] int64 = frame_list(global ii),
    owned_t [K] uint8 = (owner == rank), ii_all / jj_all: the edge indices of the WHOLE graph
    (replicated; the pose ordering and the tile layout of the reduced system are derived from them
    so that all ranks pack the system identically).  poses and the owned rows of disps are updated in place;
    with merge=True all rows of disps are consistent on return.
    segments = (seg_lo, seg_n) from `plan` switches to the SEGMENTED solve: this rank eliminates its own band
    segments, only the closure block + rhs is all-reduced, the closure factorisation is replicated; pose_owner_t
    [t1 - t0] (rank per pose) is then needed for the final merge of the poses.  Returns the context."""
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
        rank = dist.get_rank(group) if dist.is_initialized() else 0
        if segments is not None:
            _gba.check(ctx, L.gba_ba_set_shard(ctx, int(segments[0]), int(segments[1]), int(rank == 0)))
        view = _gba.system_view(ctx, dev, 0 if segments is None else 1)
        for _ in range(int(iterations)):
            _gba.check(ctx, L.gba_ba_build(ctx))
            if segments is None:
                _timed("allreduce", lambda: allreduce_system(view, group))
                _gba.check(ctx, L.gba_ba_solve(ctx))
            else:
                _gba.check(ctx, L.gba_ba_eliminate(ctx))
                _timed("allreduce", lambda: allreduce_system(view, group))
                _gba.check(ctx, L.gba_ba_finish(ctx))
        _gba.check(ctx, L.gba_ba_end(ctx, None, None))
        if merge:
            _timed("merge", lambda: merge_owned_rows(disps, kx_t, owned_t, group))
        if segments is not None:
            if merge:
                _timed("merge", lambda: merge_poses(poses, int(t0), int(t1), pose_owner_t, rank, group))
            return ctx
        # every rank factorises the same summed system, but fp64 atomics inside the factorisation are unordered
        # and a failed factorisation (zero update) is rank-local: rank 0's poses are the run's poses
        if dist.is_initialized() and dist.get_world_size(group) > 1:
            _timed("merge", lambda: dist.broadcast(poses[int(t0):int(t1)], src=dist.get_global_rank(group, 0) if group is not None else 0, group=group))
    return ctx
