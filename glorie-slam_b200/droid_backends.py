"""Drop-in for the reference's `droid_backends` extension module -- `ba` only.

Reference: src/lib/droid.cpp:89-119 (wrapper + pybind export :239-241) calling
src/lib/droid_kernels.cu:1314-1437 (ba_cuda); the only call site is
src/depth_video.py:217-218.  Same 16 positional arguments, same in-place update of `poses`
and `disps`, same return value `[dx, dz]` of the last iteration.  Put this directory first on
sys.path and `import droid_backends` resolves here; depth_video.py stays unchanged.

The work is done by libgba.so (hand-written sm_100a CUDA behind the C ABI in include/gba.h);
torch only supplies device memory and the current stream.  Also provided: `frame_distance`,
`depth_filter`, `ba_interleaved` (rank 3), `iproj` and `projmap`
(rank 4).  The remaining functions of the reference module (corr_index_*, altcorr_*) are outside this
repository's scope; set GBA_REFERENCE_BACKENDS to a module name to forward them.
"""
from __future__ import annotations

import importlib
import os

import torch

import _gba

# "droid_cuda" reproduces ba_cuda (default for the drop-in); "geom_py" reproduces
# src/geom/ba.py::BA (SURVEY.md section 8.1 lists the differences).
SEMANTICS = os.environ.get("GBA_SEMANTICS", "droid_cuda")
# The reference returns [dx, dz]; GlORIE-SLAM discards them.  K (the row count of dz) is on the host
# once the index build has read its status back, so returning them costs no extra synchronisation;
# set False to skip the dz buffer altogether.
RETURN_UPDATES = True
# A pipeline watchdog that fires (GBA_GET_WATCHDOG) ends an iteration like a failed factorisation:
# zero pose update.  That must never happen in a healthy process, so it is not silent: the count is
# checked whenever the caller synchronises anyway (`check_health`), and raises.
_watchdog_seen = {}


def _chk(t, name, dtype):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError("%s must be a CUDA tensor" % name)
    if not t.is_contiguous():
        raise RuntimeError("%s must be contiguous" % name)  # CHECK_CONTIGUOUS, droid.cpp:85-86
    if t.dtype != dtype:
        raise RuntimeError("%s must have dtype %s, got %s" % (name, dtype, t.dtype))


def ba(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep,
       motion_only, depth_only=False, _interleaved=False):
    """Dense bundle adjustment over the factor graph (ii -> jj), in place on poses / disps.

    poses [N,7] f32, disps [N,ht,wd] f32, intrinsics [4], disps_sens [N,ht,wd],
    targets / weights [E,2,ht,wd], eta [K,ht,wd], ii / jj [E] int64, all CUDA + contiguous.
    Returns [dx [t1-t0,6], dz [K,ht*wd]] of the last iteration (dz is None when motion_only).
    """
    for t, n in ((poses, "poses"), (disps, "disps"), (intrinsics, "intrinsics"), (disps_sens, "disps_sens"),
                 (targets, "targets"), (weights, "weights")):
        _chk(t, n, torch.float32)
    _chk(eta, "eta", torch.float32)
    _chk(ii, "ii", torch.int64)
    _chk(jj, "jj", torch.int64)
    if poses.dim() != 2 or poses.shape[1] != 7 or disps.dim() != 3 or disps.shape[0] != poses.shape[0]:
        raise RuntimeError("poses must be [N,7] and disps [N,ht,wd]")
    N, ht, wd = disps.shape
    E = int(ii.shape[0])
    if targets.numel() != E * 2 * ht * wd or weights.numel() != E * 2 * ht * wd or jj.shape[0] != E:
        raise RuntimeError("targets/weights must be [E,2,ht,wd] for E = len(ii) = len(jj)")
    if _interleaved and (tuple(targets.shape[-3:]) != (ht, wd, 2) or tuple(weights.shape[-3:]) != (ht, wd, 2)):
        raise RuntimeError("ba_interleaved: targets/weights must be [E,ht,wd,2]")
    t0, t1, iterations = int(t0), int(t1), int(iterations)
    dev = poses.device
    with torch.cuda.device(dev):
        ctx = _gba.context(dev.index, N, E, ht, wd)
        stream = torch.cuda.current_stream(dev).cuda_stream
        P = t1 - t0
        dx = torch.zeros((max(P, 0), 6), dtype=torch.float32, device=dev)
        dz = None
        if RETURN_UPDATES and not motion_only:
            dz = torch.zeros((N, ht * wd), dtype=torch.float32, device=dev)
        eta_rows = int(eta.shape[0]) if eta.dim() > 0 else 0
        rc = _gba.lib().gba_ba(ctx, poses.data_ptr(), disps.data_ptr(), intrinsics.data_ptr(), disps_sens.data_ptr(),
                               targets.data_ptr(), weights.data_ptr(), eta.data_ptr(), ii.data_ptr(), jj.data_ptr(),
                               N, E, eta_rows, t0, t1, iterations, float(lm), float(ep), int(bool(motion_only)),
                               int(bool(depth_only)), _gba.MODES[SEMANTICS] | (_gba.LAYOUT_INTERLEAVED if _interleaved else 0),
                               dx.data_ptr(),
                               dz.data_ptr() if dz is not None else None, stream)
        _gba.check(ctx, rc)
        if dz is not None:
            dz = dz[: _gba.lib().gba_last_depth_frames(ctx)]
    return [dx, dz]


def check_health(device=None):
    """Synchronises and raises if a kernel-internal watchdog fired since the last check (a pose update
    was then silently zero, like a failed factorisation).  Returns gba_info of the device's context."""
    out = None
    for key, (ctx, _, _) in list(_gba._ctx.items()):
        if device is not None and key[0] != torch.device(device).index:
            continue
        info = _gba.query(ctx)
        seen = _watchdog_seen.get(key, 0)
        _watchdog_seen[key] = info.watchdog_trips
        if info.watchdog_trips > seen:
            raise RuntimeError("libgba: %d kernel-internal wait(s) exceeded the watchdog (last failure code %d); the "
                               "affected iterations applied a zero pose update" % (info.watchdog_trips - seen,
                                                                                    info.last_fail_code))
        out = info
    return out


def ba_interleaved(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep,
                   motion_only, depth_only=False):
    """`ba` for targets / weights in the layout the update network produces, [E,ht,wd,2] (or
    [1,E,ht,wd,2]): the caller skips the permute + contiguous copy of src/depth_video.py:215-216.
    Opt-in; no counterpart in the reference extension.  Same
    per-pixel arithmetic as `ba`."""
    if targets.dim() == 5 and targets.shape[0] == 1:
        targets, weights = targets[0], weights[0]
    return ba(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations, lm, ep,
              motion_only, depth_only, _interleaved=True)


def frame_distance(poses, disps, intrinsics, ii, jj, beta):
    """Drop-in for the reference's droid_backends.frame_distance (src/lib/droid.cpp:122-138,
    src/lib/droid_kernels.cu:518-657): mean induced flow of every keyframe pair (ii[n] -> jj[n]),
    1000 where fewer than 75 % of the pixels stay in front of the target camera.  Returns [num] f32."""
    for t, n in ((poses, "poses"), (disps, "disps"), (intrinsics, "intrinsics")):
        _chk(t, n, torch.float32)
    _chk(ii, "ii", torch.int64)
    _chk(jj, "jj", torch.int64)
    _, ht, wd = disps.shape
    num = int(ii.shape[0])
    dev = poses.device
    dist = torch.zeros((num,), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _gba.lib().gba_frame_distance(poses.data_ptr(), disps.data_ptr(), intrinsics.data_ptr(), ii.data_ptr(),
                                           jj.data_ptr(), num, ht, wd, float(beta), dist.data_ptr(),
                                           torch.cuda.current_stream(dev).cuda_stream)
        _gba.check(None, rc)
    return dist


def depth_filter(poses, disps, intrinsics, ix, thresh):
    """Drop-in for the reference's droid_backends.depth_filter (src/lib/droid.cpp:222-236,
    src/lib/droid_kernels.cu:661-775,1494-1518; caller src/depth_video.py:349): per listed keyframe and
    pixel, the number of neighbouring keyframes (ix-1,-2,-3,+3,+4,+5) whose depth agrees with the
    reprojection within thresh.  Returns counter [len(ix), ht, wd] f32."""
    for t, n in ((poses, "poses"), (disps, "disps"), (intrinsics, "intrinsics"), (thresh, "thresh")):
        _chk(t, n, torch.float32)
    _chk(ix, "ix", torch.int64)
    n_frames, ht, wd = disps.shape
    num = int(ix.shape[0])
    if int(thresh.shape[0]) != num:
        raise ValueError("thresh must hold one value per entry of ix")
    dev = poses.device
    counter = torch.empty((num, ht, wd), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _gba.lib().gba_depth_filter(poses.data_ptr(), disps.data_ptr(), intrinsics.data_ptr(), ix.data_ptr(),
                                         thresh.data_ptr(), n_frames, num, ht, wd, counter.data_ptr(),
                                         torch.cuda.current_stream(dev).cuda_stream)
        _gba.check(None, rc)
    return counter


def iproj(poses, disps, intrinsics):
    """Drop-in for the reference's droid_backends.iproj (src/lib/droid.cpp:159-168,
    src/lib/droid_kernels.cu:779-850,1520-1545; caller src/neural_point.py:156): every pixel of every
    keyframe back-projected and transformed by its pose.  Returns points [N,ht,wd,3] f32."""
    for t, n in ((poses, "poses"), (disps, "disps"), (intrinsics, "intrinsics")):
        _chk(t, n, torch.float32)
    N, ht, wd = disps.shape
    if poses.shape[0] < N:
        raise ValueError("poses must hold one row per disparity map")
    dev = disps.device
    points = torch.empty((N, ht, wd, 3), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _gba.lib().gba_iproj(poses.data_ptr(), disps.data_ptr(), intrinsics.data_ptr(), N, ht, wd, points.data_ptr(),
                                  torch.cuda.current_stream(dev).cuda_stream)
        _gba.check(None, rc)
    return points


def projmap(poses, disps, intrinsics, ii, jj):
    """Drop-in for the reference's droid_backends.projmap (src/lib/droid.cpp:141-156,
    src/lib/droid_kernels.cu:427-516,1466-1492): where the pixels of keyframe ii[n] land in keyframe
    jj[n].  Returns [coords [num,ht,wd,3], valid [num,ht,wd,1]] like the reference."""
    for t, n in ((poses, "poses"), (disps, "disps"), (intrinsics, "intrinsics")):
        _chk(t, n, torch.float32)
    _chk(ii, "ii", torch.int64)
    _chk(jj, "jj", torch.int64)
    _, ht, wd = disps.shape
    num = int(ii.shape[0])
    dev = poses.device
    coords = torch.empty((num, ht, wd, 3), dtype=torch.float32, device=dev)
    valid = torch.empty((num, ht, wd, 1), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _gba.lib().gba_projmap(poses.data_ptr(), disps.data_ptr(), intrinsics.data_ptr(), ii.data_ptr(), jj.data_ptr(),
                                    num, ht, wd, coords.data_ptr(), valid.data_ptr(),
                                    torch.cuda.current_stream(dev).cuda_stream)
        _gba.check(None, rc)
    return [coords, valid]


def __getattr__(name):
    fwd = os.environ.get("GBA_REFERENCE_BACKENDS")
    if fwd:
        return getattr(importlib.import_module(fwd), name)
    raise AttributeError(
        "droid_backends.%s is outside the B200 hot path (provided: ba, ba_interleaved, frame_distance, depth_filter, iproj, projmap); set GBA_REFERENCE_BACKENDS "
        "to the reference extension's module name to forward it" % name)
