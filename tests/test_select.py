"""Proximity edge selection, valid-depth mask and mono_thres filter: integer / boolean
outputs, bit-exact.  tests/golden/prox.npz holds what the reference's own
FactorGraph.add_backend_proximity_factors (src/factor_graph.py:386-462) and the statements of
src/depth_video.py:245-261,349-361 produce on CPU (tests/golden/make_golden_prox.py)."""
import os

import numpy as np
import pytest

import helpers
from oracle import select_oracle


@pytest.fixture(scope="module")
def prox():
    return np.load(os.path.join(helpers.GOLDEN, "prox.npz"))


def _case(prox, name):
    n, hw, seed, t_start, t_loop, nms, radius, thresh, maxf, loop = helpers.PROX_CASES[name]
    return prox[name + "_d"], dict(t_start=t_start, t_end=n, nms=nms, radius=radius, thresh=thresh, max_factors=maxf,
                                   t_start_loop=t_loop, loop=loop)


@pytest.mark.parametrize("name", list(helpers.PROX_CASES))
def test_oracle_selection_equals_the_reference(name, prox):
    d, kw = _case(prox, name)
    ii, jj = select_oracle.proximity_select(d, **kw)
    np.testing.assert_array_equal(ii, prox[name + "_ii"])
    np.testing.assert_array_equal(jj, prox[name + "_jj"])
    n_local = 2 * sum(min(kw["radius"] + 1, q) for q in range(kw["t_start"] if not kw["loop"] else kw["t_start_loop"], kw["t_end"]))
    assert len(ii) > n_local  # the selection added edges beyond the local window


def test_fixture_inputs_are_reproducible(prox):
    """the stored distance vectors come from the seeded scene through the frame_distance oracle"""
    for name, c in helpers.PROX_CASES.items():
        _, _, _, d = helpers.prox_distances(c[0], c[1], c[2], c[3], c[4], c[9])
        np.testing.assert_array_equal(d, prox[name + "_d"])


def test_oracle_mask_and_filter_equal_the_reference_statements(prox):
    disps, count, error, scale, ii, jj = helpers.mask_case()
    mask, med = select_oracle.valid_depth_mask(disps, count, 2)
    np.testing.assert_array_equal(mask, prox["mask"])
    np.testing.assert_array_equal(np.isnan(med), np.isnan(prox["mask_median"]))
    np.testing.assert_array_equal(med[~np.isnan(med)], prox["mask_median"][~np.isnan(med)])
    assert np.isnan(med[3]) and not mask[3].any()
    inv, keep, eta_keep = select_oracle.mono_filter(error, prox["filter_avg"], scale, mask.sum(axis=(1, 2)),
                                                    mask[0].size, 0.1, ii, jj)
    np.testing.assert_array_equal(inv, prox["filter_invalid"])
    np.testing.assert_array_equal(keep, prox["filter_keep"])
    np.testing.assert_array_equal(eta_keep, prox["filter_eta"])


# ------------------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", list(helpers.PROX_CASES))
def test_gpu_selection_equals_the_reference(name, prox):
    import torch

    import video_ops

    d, kw = _case(prox, name)
    ii, jj = video_ops.proximity_select(torch.from_numpy(d).cuda(), **kw)
    np.testing.assert_array_equal(ii.cpu().numpy(), prox[name + "_ii"])
    np.testing.assert_array_equal(jj.cpu().numpy(), prox[name + "_jj"])


@pytest.mark.gpu
def test_gpu_selection_through_frame_distance_and_degenerate_cases(prox):
    """the full entry point (all-pairs frame distance on the GPU, then the selection) against the NumPy restatement
    fed with the GPU's own distances; a window with nothing to select returns None like the reference returns 0"""
    import torch

    import droid_backends
    import video_ops

    n, hw, seed, t_start, t_loop, nms, radius, thresh, maxf, loop = helpers.PROX_CASES["global"]
    poses, disps, intr, _ = helpers.prox_distances(n, hw, seed, t_start, t_loop, loop)
    tp, td, ti = (torch.from_numpy(a).cuda() for a in (poses, disps, intr))
    got = video_ops.backend_proximity_edges(tp, td, ti, t_start, n, nms, radius, thresh, maxf, 0.25)
    gi, gj = torch.meshgrid(torch.arange(t_start, n), torch.arange(t_start, n), indexing="ij")
    gi, gj = gi.reshape(-1).cuda(), gj.reshape(-1).cuda()
    d = 0.5 * (droid_backends.frame_distance(tp, td, ti, gi, gj, 0.25) + droid_backends.frame_distance(tp, td, ti, gj, gi, 0.25))
    want = select_oracle.proximity_select(d.cpu().numpy(), t_start, n, nms, radius, thresh, maxf)
    np.testing.assert_array_equal(got[0].cpu().numpy(), want[0])
    np.testing.assert_array_equal(got[1].cpu().numpy(), want[1])
    assert video_ops.proximity_select(torch.full((4,), 5.0).cuda(), 0, 2, 1, 2, 1.0, 100) is None  # 2 local edges only
    big = helpers.prox_distances(150, (8, 12), 9, 0, None)[3]  # more candidates than one sort tile
    w = select_oracle.proximity_select(big, 0, 150, 1, 2, 40.0, 10 ** 6)
    g = video_ops.proximity_select(torch.from_numpy(big).cuda(), 0, 150, 1, 2, 40.0, 10 ** 6)
    assert (big <= 40.0).sum() > 4096
    np.testing.assert_array_equal(g[0].cpu().numpy(), w[0])
    np.testing.assert_array_equal(g[1].cpu().numpy(), w[1])


@pytest.mark.gpu
def test_gpu_mask_and_filter_equal_the_reference_statements(prox):
    import torch

    import video_ops

    disps, count, error, scale, ii, jj = helpers.mask_case()
    mask, cnt, mean = video_ops.mask_from_count(torch.from_numpy(disps).cuda(), torch.from_numpy(count).cuda(), 2)
    np.testing.assert_array_equal(mask.cpu().numpy(), prox["mask"])
    np.testing.assert_array_equal(cnt.cpu().numpy(), prox["mask"].sum(axis=(1, 2)))
    np.testing.assert_allclose(mean.cpu().numpy(), prox["filter_avg"], rtol=1e-6)
    n, ht, wd = disps.shape
    inv, keep, src = video_ops.mono_filter(torch.from_numpy(error).cuda(), torch.from_numpy(prox["filter_avg"]).cuda(),
                                           torch.from_numpy(scale).cuda(), cnt, ht, wd, 0.1, torch.from_numpy(ii).cuda(),
                                           torch.from_numpy(jj).cuda(), n + 2)
    np.testing.assert_array_equal(inv.cpu().numpy(), prox["filter_invalid"])
    np.testing.assert_array_equal(keep.cpu().numpy(), prox["filter_keep"])
    np.testing.assert_array_equal(src.cpu().numpy()[np.unique(ii)], prox["filter_eta"])
