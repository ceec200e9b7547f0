"""Mirror of the reference's src/geom/ba.py entry points on top of libgba.so.

* `BA_with_scale_shift` -- DSPO stage 2 (reference src/geom/ba.py:127-216), same signature and
  functional behaviour (returns new disps / wqs, poses passed through).  Monkey-patch it over
  `src.geom.ba.BA_with_scale_shift`; src/depth_video.py:266-273 stays unchanged.
* `BA` -- one Gauss-Newton iteration with the semantics of reference src/geom/ba.py:33-121
  (damping before the Schur complement, no back-substitution skip, clamp at 0), functional.
"""
from __future__ import annotations

import torch

import _gba


def _data(poses):
    return poses.data if hasattr(poses, "manifold_dim") else poses


def _rewrap(poses, data):
    return poses.__class__(data) if hasattr(poses, "manifold_dim") else data


def BA_with_scale_shift(target, weight, eta, poses, disps, intrinsics, ii, jj, mono_disps, scales=None, shifts=None,
                        valid_depth_mask=None, ignore_frames=0, lm=0.0001, ep=0.1, alpha=1.0, fixedp=1, rig=1):
    """target / weight [1,E,ht,wd,2] (interleaved), eta [M,ht,wd], poses SE3 [1,N] (or [1,N,7]),
    disps / mono_disps [1,N,ht,wd], intrinsics [1,N,4], scales / shifts [1,N],
    valid_depth_mask [1,N,ht,wd] bool.  Returns (poses, disps [1,N,ht,wd], wqs [1,N,2])."""
    if rig != 1:
        raise NotImplementedError("rig != 1 is not used by GlORIE-SLAM")
    pd = _data(poses)
    B, N, ht, wd = disps.shape
    if B != 1:
        raise RuntimeError("batch size must be 1 (as in the reference call site)")
    dev = disps.device
    E = int(ii.shape[0])
    f32 = lambda t: t.to(device=dev, dtype=torch.float32).contiguous()
    tg, wg = f32(target).view(E, ht, wd, 2), f32(weight).view(E, ht, wd, 2)
    disps_o = f32(disps[0]).clone()
    sc, sh = f32(scales[0]).clone(), f32(shifts[0]).clone()
    mono = f32(mono_disps[0])
    mask = valid_depth_mask[0].to(device=dev, dtype=torch.uint8).contiguous()
    intr = f32(intrinsics[0, 0])
    et = f32(eta)
    iic, jjc = ii.to(device=dev, dtype=torch.int64).contiguous(), jj.to(device=dev, dtype=torch.int64).contiguous()
    p7 = f32(pd[0])
    with torch.cuda.device(dev):
        ctx = _gba.context(dev.index, N, E, ht, wd)
        rc = _gba.lib().gba_dspo(ctx, p7.data_ptr(), disps_o.data_ptr(), intr.data_ptr(), tg.data_ptr(), wg.data_ptr(),
                                 et.data_ptr(), iic.data_ptr(), jjc.data_ptr(), mono.data_ptr(), sc.data_ptr(),
                                 sh.data_ptr(), mask.data_ptr(), N, E, int(et.shape[0]), int(ignore_frames),
                                 float(lm), float(ep), float(alpha), 1, torch.cuda.current_stream(dev).cuda_stream)
        _gba.check(ctx, rc)
    wqs = torch.stack([sc, sh], dim=-1)[None]
    return poses, disps_o[None], wqs


def BA(target, weight, eta, poses, disps, intrinsics, ii, jj, sensor_disps=None, lm=0.0001, ep=0.1, alpha=0.05,
       fixedp=1, rig=1):
    """One iteration of the reference's PyTorch BA semantics (src/geom/ba.py:33-121)."""
    if rig != 1 or alpha != 0.05:
        raise NotImplementedError("rig != 1 / alpha != 0.05 are not used by GlORIE-SLAM")
    pd = _data(poses)
    B, N, ht, wd = disps.shape
    dev = disps.device
    E = int(ii.shape[0])
    f32 = lambda t: t.to(device=dev, dtype=torch.float32).contiguous()
    tg = f32(target).view(E, ht, wd, 2).permute(0, 3, 1, 2).contiguous()
    wg = f32(weight).view(E, ht, wd, 2).permute(0, 3, 1, 2).contiguous()
    p7, d = f32(pd[0]).clone(), f32(disps[0]).clone()
    sens = f32(sensor_disps[0]) if sensor_disps is not None else torch.zeros_like(d)
    et = f32(eta)
    intr = f32(intrinsics[0, 0])
    iic, jjc = ii.to(device=dev, dtype=torch.int64).contiguous(), jj.to(device=dev, dtype=torch.int64).contiguous()
    with torch.cuda.device(dev):
        ctx = _gba.context(dev.index, N, E, ht, wd)
        rc = _gba.lib().gba_ba(ctx, p7.data_ptr(), d.data_ptr(), intr.data_ptr(), sens.data_ptr(),
                               tg.data_ptr(), wg.data_ptr(), et.data_ptr(), iic.data_ptr(), jjc.data_ptr(), N, E,
                               int(et.shape[0]), int(fixedp), N, 1, float(lm), float(ep), 0, 0, _gba.MODE_GEOM_PY,
                               None, None, torch.cuda.current_stream(dev).cuda_stream)
        _gba.check(ctx, rc)
    return _rewrap(poses, p7[None]), d[None]


def MoBA(target, weight, eta, poses, disps, intrinsics, ii, jj, fixedp=1, rig=1):
    """Motion-only bundle adjustment with the reference's semantics (src/geom/ba.py:223-271:
    block_solve with its default damping ep = 0.1, lm = 1e-4, src/geom/chol.py:44-55)."""
    if rig != 1:
        raise NotImplementedError("rig != 1 is not used by GlORIE-SLAM")
    pd = _data(poses)
    B, N, ht, wd = disps.shape
    dev = disps.device
    E = int(ii.shape[0])
    f32 = lambda t: t.to(device=dev, dtype=torch.float32).contiguous()
    tg = f32(target).view(E, ht, wd, 2).permute(0, 3, 1, 2).contiguous()
    wg = f32(weight).view(E, ht, wd, 2).permute(0, 3, 1, 2).contiguous()
    p7, d = f32(pd[0]).clone(), f32(disps[0]).clone()
    sens = torch.zeros_like(d)
    intr = f32(intrinsics[0, 0])
    dummy_eta = torch.zeros((1, ht, wd), dtype=torch.float32, device=dev)
    iic, jjc = ii.to(device=dev, dtype=torch.int64).contiguous(), jj.to(device=dev, dtype=torch.int64).contiguous()
    with torch.cuda.device(dev):
        ctx = _gba.context(dev.index, N, E, ht, wd)
        rc = _gba.lib().gba_ba(ctx, p7.data_ptr(), d.data_ptr(), intr.data_ptr(), sens.data_ptr(), tg.data_ptr(),
                               wg.data_ptr(), dummy_eta.data_ptr(), iic.data_ptr(), jjc.data_ptr(), N, E, 1,
                               int(fixedp), N, 1, 1e-4, 0.1, 1, 0, _gba.MODE_GEOM_PY, None, None,
                               torch.cuda.current_stream(dev).cuda_stream)
        _gba.check(ctx, rc)
    return _rewrap(poses, p7[None])
