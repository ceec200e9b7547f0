"""Device versions of the Python loops that surround the BA calls in GlORIE-SLAM:

* `backend_proximity_edges`  -- FactorGraph.add_backend_proximity_factors (reference src/factor_graph.py:386-462)
  up to the edge list it hands to `add_factors`: all-pairs bidirectional frame distance, thresholds, local-window
  edges, sorted greedy selection with non-maximum suppression.  The reference walks the sorted candidates in
  Python with one `.item()` per candidate.
* `valid_depth_mask`         -- DepthVideo.update_valid_depth_mask (src/depth_video.py:326-361).
* `mono_filter`              -- the mono_thres edge / eta filter of DepthVideo.dspo (src/depth_video.py:245-261).

All work is done by libgba.so (csrc/gba_select.cu, csrc/gba_geom.cu) through the C ABI; torch supplies memory.
"""
from __future__ import annotations

import torch

import _gba
import droid_backends


def _stream(dev):
    return torch.cuda.current_stream(dev).cuda_stream


def proximity_select(d, t_start, t_end, nms, radius, thresh, max_factors, t_start_loop=None, loop=False):
    """The selection alone, on a given distance vector d [ilen*jlen] (float32 CUDA).  Returns (ii, jj) int64 edge
    tensors in the reference's order, or None where the reference returns 0 without adding edges."""
    if t_start_loop is None or not loop:
        t_start_loop = t_start
    ilen, jlen = t_end - t_start_loop, t_end - t_start
    dev = d.device
    d = d.to(torch.float32).contiguous()
    if d.numel() != ilen * jlen:
        raise ValueError("d must hold (t_end - t_start_loop) x (t_end - t_start) distances")
    L = _gba.lib()
    n_local = 2 * sum(min(radius + 1, q) for q in range(t_start_loop, t_end))
    cap = n_local + min(int(max_factors), (9 if loop else 2) * ilen * jlen) + 16
    scratch = torch.empty(L.gba_proximity_scratch_bytes(ilen, jlen), dtype=torch.uint8, device=dev)
    out = torch.empty((2, cap), dtype=torch.int32, device=dev)
    out_n = torch.zeros(2, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = L.gba_proximity_select(d.data_ptr(), int(t_start), int(t_start_loop), int(t_end), int(nms), int(radius),
                                    float(thresh), int(max_factors), int(bool(loop)), scratch.data_ptr(), scratch.numel(),
                                    out[0].data_ptr(), out[1].data_ptr(), cap, out_n.data_ptr(), _stream(dev))
        _gba.check(None, rc)
    n, status = (int(v) for v in out_n.tolist())  # the caller needs the edge count on the host anyway
    if status == 2:
        raise RuntimeError("proximity_select: edge capacity exceeded")
    if status == 1:
        return None
    return out[0, :n].to(torch.int64), out[1, :n].to(torch.int64)


def backend_proximity_edges(poses, disps, intrinsics, t_start, t_end, nms, radius, thresh, max_factors, beta,
                            t_start_loop=None, loop=False):
    """poses [*,7], disps [*,ht,wd], intrinsics [4] (CUDA float32).  Returns (ii, jj) or None, see proximity_select."""
    lo = t_start if (t_start_loop is None or not loop) else t_start_loop
    dev = poses.device
    ix = torch.arange(lo, t_end, device=dev)
    jx = torch.arange(t_start, t_end, device=dev)
    ii, jj = torch.meshgrid(ix, jx, indexing="ij")
    ii, jj = ii.reshape(-1).contiguous(), jj.reshape(-1).contiguous()
    d1 = droid_backends.frame_distance(poses, disps, intrinsics, ii, jj, beta)  # src/depth_video.py:181-188
    d2 = droid_backends.frame_distance(poses, disps, intrinsics, jj, ii, beta)
    return proximity_select(0.5 * (d1 + d2), t_start, t_end, nms, radius, thresh, max_factors, t_start_loop, loop)


def valid_depth_mask(poses, disps, intrinsics, index, mv_thresh, visible_num):
    """update_valid_depth_mask for the keyframes in `index` (int64): depth_filter votes, then the median rule.
    Returns (mask [n,ht,wd] bool, valid_cnt [n] int32, mean_disp [n] float32)."""
    dev = disps.device
    sel = torch.index_select(disps, 0, index).contiguous()
    depths_mean = (1.0 / sel).mean(dim=[1, 2])
    thresh = (mv_thresh * depths_mean).to(torch.float32).contiguous()
    count = droid_backends.depth_filter(poses, disps, intrinsics, index, thresh)
    return mask_from_count(sel, count, visible_num)


def mask_from_count(disps_sel, count, visible_num):
    n, ht, wd = disps_sel.shape
    dev = disps_sel.device
    mask = torch.empty((n, ht, wd), dtype=torch.uint8, device=dev)
    cnt = torch.empty(n, dtype=torch.int32, device=dev)
    mean = torch.empty(n, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _gba.lib().gba_valid_depth_mask(disps_sel.contiguous().data_ptr(), count.contiguous().data_ptr(), n, ht, wd,
                                             int(visible_num), mask.data_ptr(), cnt.data_ptr(), mean.data_ptr(), _stream(dev))
        _gba.check(None, rc)
    return mask.bool(), cnt, mean


def mono_filter(error, avg_disp, scale, valid_cnt, ht, wd, mono_thres, ii, jj, n_frames):
    """Returns (invalid [n] bool, keep [E] bool, src_kept [n_frames] bool); eta rows follow unique(ii), so
    `eta[src_kept[unique(ii)]]` is the reference's `eta[valid_eta_mask]`."""
    dev = error.device
    n, E = int(error.shape[0]), int(ii.shape[0])
    invalid = torch.zeros(max(n, 1), dtype=torch.uint8, device=dev)
    keep = torch.zeros(max(E, 1), dtype=torch.uint8, device=dev)
    src = torch.zeros(n_frames, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        rc = _gba.lib().gba_mono_filter(error.float().contiguous().data_ptr(), avg_disp.float().contiguous().data_ptr(),
                                        scale.float().contiguous().data_ptr(), valid_cnt.to(torch.int32).contiguous().data_ptr(),
                                        n, ht, wd, float(mono_thres), ii.contiguous().data_ptr(), jj.contiguous().data_ptr(), E,
                                        int(n_frames), invalid.data_ptr(), keep.data_ptr(), src.data_ptr(), _stream(dev))
        _gba.check(None, rc)
    return invalid[:n].bool(), keep[:E].bool(), src.bool()
