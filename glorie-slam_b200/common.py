"""Host-side mirror of the one function of the reference's src/utils/common.py that sits next to
the BA hot path: align_scale_and_shift (src/utils/common.py:402-438), the per-keyframe weighted
scale/shift fit of the mono-depth prior that seeds and follows DSPO (src/depth_video.py:233,308).
One CUDA launch through the C-ABI (gba_align_scale_shift) instead of ~15 torch reductions."""
from __future__ import annotations

import torch

import _gba


def align_scale_and_shift(prediction, target, weights):
    """prediction, target, weights: [B,H,W] (or [H,W]) float32 CUDA tensors; weights may be None or
    bool.  Returns (scale [B], shift [B], avg_error [B]) like the reference."""
    if weights is None:
        weights = torch.ones_like(prediction)
    if prediction.dim() < 3:
        prediction, target, weights = prediction.unsqueeze(0), target.unsqueeze(0), weights.unsqueeze(0)
    if not prediction.is_cuda:
        raise RuntimeError("align_scale_and_shift: the B200 path needs CUDA tensors (there is no CPU fallback)")
    p, t, w = (x.to(torch.float32).contiguous() for x in (prediction, target, weights))
    if p.shape != t.shape or p.shape != w.shape:
        raise ValueError("prediction, target and weights must have the same shape")
    n, ht, wd = p.shape
    dev = p.device
    out = torch.empty((3, n), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _gba.lib().gba_align_scale_shift(p.data_ptr(), t.data_ptr(), w.data_ptr(), n, ht, wd, out[0].data_ptr(),
                                              out[1].data_ptr(), out[2].data_ptr(),
                                              torch.cuda.current_stream(dev).cuda_stream)
        _gba.check(None, rc)
    return out[0], out[1], out[2]
