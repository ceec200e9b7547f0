"""NumPy restatement of the Python loops around the BA calls (TEST INFRASTRUCTURE ONLY; see oracle/oracle.py).

* proximity_select   FactorGraph.add_backend_proximity_factors, reference src/factor_graph.py:386-462
* valid_depth_mask   DepthVideo.update_valid_depth_mask after depth_filter, src/depth_video.py:349-361
* mono_filter        the mono_thres edge / eta filter of DepthVideo.dspo, src/depth_video.py:245-261
Pinned by tests/golden/prox.npz (tests/golden/make_golden_prox.py runs the reference's own function / statements).
"""
import numpy as np


def proximity_select(d, t_start, t_end, nms, radius, thresh, max_factors, t_start_loop=None, loop=False):
    if t_start_loop is None or not loop:  # :387-388
        t_start_loop = t_start
    ilen, jlen = t_end - t_start_loop, t_end - t_start
    d = np.array(d, dtype=np.float32).reshape(ilen, jlen)
    rawd = d.copy()
    ii, jj = np.meshgrid(np.arange(t_start_loop, t_end), np.arange(t_start, t_end), indexing="ij")
    d[ii - radius < jj] = np.inf  # :402
    d[d > thresh] = np.inf        # :403
    es = []
    for i in range(t_start_loop, t_end):  # :408-413
        for j in range(max(i - radius - 1, 0), i):
            es += [(i, j), (j, i)]
            d[i - t_start_loop, j - t_start] = np.inf  # a negative column wraps, as in the reference
    flat = d.reshape(-1)
    order = np.argsort(flat, kind="stable")  # :416
    order = order[flat[order] <= thresh]
    loop_edges = 0
    for k in order:  # :424-452
        di, dj = k // jlen, k % jlen
        if d[di, dj] > thresh:
            continue
        if len(es) > max_factors:
            break
        i, j = int(ii.reshape(-1)[k]), int(jj.reshape(-1)[k])
        if loop:
            sub = []
            for si in range(max(i - 1, t_start_loop), min(i + 2, t_end)):
                for sj in range(max(j - 1, t_start), min(j + 2, t_end)):
                    if rawd[si - t_start_loop, sj - t_start] <= thresh and si != sj and si - sj > 20:
                        sub.append((si, sj))
            es += sub
            loop_edges += len(sub)
        else:
            es += [(i, j), (j, i)]
        d[max(0, di - nms):min(ilen, di + nms + 1), max(0, dj - nms):min(jlen, dj + nms + 1)] = np.inf
    if len(es) < 3 or (loop and loop_edges == 0):  # :455-456
        return None
    e = np.asarray(es, dtype=np.int64)
    return e[:, 0].copy(), e[:, 1].copy()


def valid_depth_mask(disps, count, visible_num):
    """Returns (mask [n,ht,wd] bool, median [n] float32).  torch.nanmedian = the lower middle of the non-NaN values."""
    disps = np.asarray(disps, dtype=np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        depths = (np.float32(1.0) / disps).astype(np.float32)
    depths = np.where(np.asarray(count) >= visible_num, depths, np.float32(np.nan))
    n = depths.shape[0]
    med = np.full(n, np.nan, dtype=np.float32)
    for f in range(n):
        v = depths[f].reshape(-1)
        v = np.sort(v[~np.isnan(v)])
        if len(v):
            med[f] = v[(len(v) - 1) // 2]
    with np.errstate(invalid="ignore"):
        mask = depths < (np.float32(3.0) * med)[:, None, None]
    return mask, med


def mono_filter(error, avg_disp, scale, valid_cnt, hw, mono_thres, ii, jj):
    """Returns (invalid [n] bool, keep [E] bool, eta_keep [len(unique(ii))] bool)."""
    error, avg_disp, scale = (np.asarray(a, dtype=np.float32) for a in (error, avg_disp, scale))
    with np.errstate(invalid="ignore", divide="ignore"):
        invalid = (error / avg_disp > np.float32(mono_thres)) | np.isnan(error) | (scale < 0) | \
                  (np.asarray(valid_cnt) < hw * 0.5)
    bad = np.where(invalid)[0]
    drop = np.isin(ii, bad) | np.isin(jj, bad)
    eta_keep = np.isin(np.unique(ii), np.unique(np.asarray(ii)[~drop]))
    return invalid, ~drop, eta_keep
