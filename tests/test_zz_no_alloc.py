"""gba_edge_blocks_static: the one entry point that allocates nothing.

It serves callers that cannot call cudaMalloc when they need the per-edge pose blocks (a tracking
thread that must not wait on the driver's allocator, a process whose device memory is already
owned by another allocator), so its scratch is a fixed array inside libgba.so.  This test drives
it through the C ABI on two calls in a row that differ in edge count and image size -- the second
call finds the first call's leftovers in the shared array -- and checks each result against what
the ordinary path (gba_ba + gba_get(GBA_GET_EDGE_HV)) computes for the same inputs.

The file sorts last on purpose: with `pytest -x` everything else has already run if it fails.
It must also pass on its own in a fresh process (the first call then performs the per-device
address look-up).
"""
import ctypes

import numpy as np
import pytest
import torch

import synth
from helpers import rel

pytestmark = pytest.mark.gpu

import _gba  # noqa: E402
from gba_run import get_array, run_ba, to_dev  # noqa: E402

MAX_EDGES, MAX_PIXELS = 4096, 4800  # GBA_STATIC_MAX_EDGES / GBA_STATIC_MAX_PIXELS in include/gba.h


def static_blocks(g, mode=0):
    d = to_dev(g)
    N, ht, wd = g.disps.shape
    out = torch.full((g.n_edges, 32), float("nan"), dtype=torch.float32, device="cuda:0")
    rc = _gba.lib().gba_edge_blocks_static(
        d.poses.data_ptr(), d.disps.data_ptr(), d.intr.data_ptr(), d.targets.data_ptr(), d.weights.data_ptr(),
        d.ii.data_ptr(), d.jj.data_ptr(), N, g.n_edges, ht, wd, mode, out.data_ptr(),
        torch.cuda.current_stream().cuda_stream)
    assert rc == 0, _gba.lib().gba_last_error(None)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def existing_blocks(g, mode="droid_cuda"):
    r = run_ba(g, mode=mode, iters=1, motion_only=True)  # linearises at the input state
    return get_array(r.ctx, _gba.GET_EDGE_HV, (g.n_edges, 32), np.float32)


def test_two_calls_in_a_row_match_the_existing_entry_point():
    big = synth.make_config("B")                                    # 250 edges, 48 x 64: 3 CTAs per edge
    small = synth.make_graph(6, 14, 2, seed=3, ht=30, wd=40, t0=2)  # 14 edges, 30 x 40: 2 CTAs per edge
    for g in (big, small):  # the second call reuses (part of) the scratch the first one dirtied
        got = static_blocks(g)
        want = existing_blocks(g)
        assert np.isfinite(got).all()
        assert rel(got[:, :21], want[:, :21]) < 1e-5   # H_jj, same tolerance as the field tests
        assert rel(got[:, 21:27], want[:, 21:27]) < 1e-5  # v_j
        np.testing.assert_array_equal(got[:, 27:], 0)
    # and once more on the first problem: counters were left clean
    again = static_blocks(big)
    assert rel(again[:, :27], existing_blocks(big)[:, :27]) < 1e-5


def test_geom_validity_mode_and_out_of_range_edges():
    g = synth.make_graph(5, 8, 1, seed=2, ht=8, wd=12)
    assert rel(static_blocks(g, mode=1)[:, :27], existing_blocks(g, mode="geom_py")[:, :27]) < 1e-5
    g.jj = g.jj.copy()
    g.jj[3] = 99  # not a keyframe: that edge's block must come back as zeros, the rest untouched
    got = static_blocks(g)
    np.testing.assert_array_equal(got[3], 0)
    assert np.abs(got[[0, 1, 2, 4]]).sum() > 0


def test_sizes_beyond_the_built_in_scratch_are_refused():
    L = _gba.lib()
    x = torch.zeros(64, device="cuda:0")
    i = torch.zeros(8, dtype=torch.int64, device="cuda:0")
    args = [x.data_ptr()] * 5 + [i.data_ptr()] * 2
    rc = L.gba_edge_blocks_static(*args, 4, MAX_EDGES + 1, 48, 64, 0, x.data_ptr(), None)
    assert rc == -3 and b"exceeds the built-in scratch" in L.gba_last_error(None)
    rc = L.gba_edge_blocks_static(*args, 4, 8, 80, 80, 0, x.data_ptr(), None)  # 6400 pixels > 4800
    assert rc == -3 and b"4800" in L.gba_last_error(None)
    rc = L.gba_edge_blocks_static(*args, 4, 8, 48, 64, 7, x.data_ptr(), None)
    assert rc == -1
