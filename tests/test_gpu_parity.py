"""GPU parity tests proper: libgba.so (through the C ABI) against the CPU oracle on the same seeded
inputs, against the golden vectors produced by the reference itself, and -- at full size --
through size-independent properties.  Tolerance: BASELINE.json:north_star, 1e-4 relative on the
state (poses, disparities) after a fixed iteration count, fp32 arithmetic; index structures exact."""
import ctypes

import numpy as np
import pytest
import torch

import synth
import helpers
from helpers import STATE_RTOL, load_golden, rel, sign_normalise, small_graphs
from oracle import oracle

pytestmark = pytest.mark.gpu

import _gba  # noqa: E402
from gba_run import get_array, run_ba, segmented, split_phase, to_dev  # noqa: E402


# ------------------------------------------------------------------------------------------
# index structures: bit-exact (droid_kernels.cu:948-984,1337-1345)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["A", "R", "tiny", "B"])
def test_index_structures_exact(name):
    g = small_graphs()[name]
    r = run_ba(g, iters=1)
    kx = np.unique(np.concatenate([np.arange(g.t0, g.t1), g.ii]))
    K = len(kx)
    assert r.info.n_depth_frames == K
    np.testing.assert_array_equal(get_array(r.ctx, _gba.GET_KX, (K,), np.int32), kx)
    order = np.argsort(g.ii, kind="stable")  # accum_cuda: argsort + CSR by source keyframe
    ptr = np.searchsorted(g.ii[order], kx, side="left")
    ptr = np.concatenate([ptr, [g.n_edges]]) if kx[-1] >= g.ii.max() else None
    eptr = get_array(r.ctx, _gba.GET_EPTR, (K + 1,), np.int32)
    eidx = get_array(r.ctx, _gba.GET_EIDX, (g.n_edges,), np.int32)
    counts = np.array([(g.ii == f).sum() for f in kx])
    np.testing.assert_array_equal(np.diff(eptr), counts)
    np.testing.assert_array_equal(eidx, order)
    assert r.info.max_out_degree == counts.max()


def test_index_structures_shuffled_edges_and_sparse_window():
    """edges in random order, a window keyframe without out-edges, edges from fixed keyframes"""
    g = synth.make_graph(9, 20, 2, iters=1, seed=11, ht=8, wd=12, t0=3)
    rng = np.random.default_rng(0)
    keep = g.ii != 5  # keyframe 5 keeps no out-edge but stays in the window
    perm = rng.permutation(int(keep.sum()))
    g.ii, g.jj = g.ii[keep][perm], g.jj[keep][perm]
    g.targets, g.weights = g.targets[keep][perm], g.weights[keep][perm]
    kx = np.unique(np.concatenate([np.arange(g.t0, g.t1), g.ii]))
    g.eta = np.resize(g.eta, (len(kx),) + g.eta.shape[1:]).astype(np.float32)
    r = run_ba(g, iters=1)
    np.testing.assert_array_equal(get_array(r.ctx, _gba.GET_KX, (len(kx),), np.int32), kx)
    np.testing.assert_array_equal(get_array(r.ctx, _gba.GET_EIDX, (g.n_edges,), np.int32),
                                  np.argsort(g.ii, kind="stable"))
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  1, g.lm, g.ep, precision="f64")
    assert rel(sign_normalise(r.poses), sign_normalise(o.poses)) < STATE_RTOL
    assert rel(r.disps, o.disps) < STATE_RTOL


# ------------------------------------------------------------------------------------------
# first-iteration fields against the oracle
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["A", "R", "B"])
def test_linearisation_and_reduced_system(name):
    g = small_graphs()[name]
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  1, g.lm, g.ep, precision="f64", debug=True)
    s = split_phase(g, mode=0, iters=1)
    # reduced camera matrix A - S (lower triangle) and rhs b - vS
    n = 6 * (g.t1 - g.t0)
    Mo = np.tril(o.M)
    assert rel(s.M, Mo) < 2e-5
    assert rel(s.rhs, o.rhs) < 2e-5
    # pose update and state after one iteration
    assert rel(sign_normalise(s.poses), sign_normalise(o.poses)) < STATE_RTOL
    assert rel(s.disps, o.disps) < STATE_RTOL
    # per-edge H_jj / v_j and the depth diagonal through the one-shot entry point
    r = run_ba(g, iters=1)
    hv = get_array(r.ctx, _gba.GET_EDGE_HV, (g.n_edges, 32), np.float32)
    iu = np.triu_indices(6)
    Hjj = o.Hs[:, 3][:, iu[0], iu[1]]
    assert rel(hv[:, :21], Hjj) < 1e-5
    assert rel(hv[:, 21:27], o.vs[:, 1]) < 1e-5
    K = len(o.kx)
    # Q, W of the call are from its only iteration
    Q = get_array(r.ctx, _gba.GET_Q, (K, g.hw), np.float32)
    W = get_array(r.ctx, _gba.GET_W, (K, g.hw), np.float32)
    assert rel(1.0 / Q, o.C) < 1e-5
    assert rel(W, o.w) < 1e-5
    assert rel(r.dx, o.dx) < 1e-3  # dx itself is small; the bar is on the state
    assert rel(r.dz, o.dz) < 1e-3


# ------------------------------------------------------------------------------------------
# state after the configured iteration count, both semantics modes
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["A", "R", "tiny", "B"])
@pytest.mark.parametrize("mode", ["droid_cuda", "geom_py"])
def test_state_matches_oracle(name, mode):
    g = small_graphs()[name]
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  g.iters, g.lm, g.ep, mode=oracle.MODE_DROID_CUDA if mode == "droid_cuda" else oracle.MODE_GEOM_PY,
                  precision="f64")
    r = run_ba(g, mode=mode)
    assert r.info.chol_failures == 0 and o.fail == 0
    assert rel(sign_normalise(r.poses), sign_normalise(o.poses)) < STATE_RTOL
    assert rel(r.disps, o.disps) < STATE_RTOL
    np.testing.assert_array_equal(r.poses[: g.t0], g.poses[: g.t0])  # fixed poses never move


@pytest.mark.parametrize("name", ["A", "R", "near", "step6"])
def test_geom_mode_matches_reference_golden(name):
    """golden vectors = the reference's own src/geom/ba.py::BA run on CPU (tests/golden/make_golden.py)"""
    g, z, kw = load_golden(name)
    r = run_ba(g, mode="geom_py")
    for ref in ("f64", "f32"):
        assert rel(sign_normalise(r.poses), sign_normalise(z["poses_" + ref])) < STATE_RTOL
        assert rel(r.disps, z["disps_" + ref].reshape(r.disps.shape)) < STATE_RTOL


def test_geom_ba_mirror_is_functional():
    import geom_ba

    g, z, kw = load_golden("A")
    d = to_dev(g)
    target = d.targets.permute(0, 2, 3, 1).contiguous()[None]
    weight = d.weights.permute(0, 2, 3, 1).contiguous()[None]
    poses, disps = d.poses[None].clone(), d.disps[None].clone()
    intr = d.intr[None, None].repeat(1, g.n_frames, 1)
    for _ in range(g.iters):
        poses, disps = geom_ba.BA(target, weight, d.eta, poses, disps, intr, d.ii, d.jj, lm=g.lm, ep=g.ep, fixedp=g.t0)
    assert rel(sign_normalise(poses[0].cpu().numpy()), sign_normalise(z["poses_f64"])) < STATE_RTOL
    assert rel(disps[0].cpu().numpy(), z["disps_f64"]) < STATE_RTOL
    np.testing.assert_array_equal(d.poses.cpu().numpy(), g.poses)  # inputs untouched


# ------------------------------------------------------------------------------------------
# flags, contracts, errors
# ------------------------------------------------------------------------------------------
def test_motion_only_and_depth_only():
    g = small_graphs()["R"]
    for kw in (dict(motion_only=True), dict(depth_only=True)):
        o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0,
                      g.t1, 2, g.lm, g.ep, precision="f64", **kw)
        r = run_ba(g, iters=2, **kw)
        assert rel(sign_normalise(r.poses), sign_normalise(o.poses)) < STATE_RTOL
        assert rel(r.disps, o.disps) < STATE_RTOL
        if "motion_only" in kw:
            np.testing.assert_array_equal(r.disps, g.disps)
            assert r.dz is None
        else:
            np.testing.assert_array_equal(r.poses, g.poses)
            assert rel(r.dx, o.dx) < 1e-3


def test_sensor_depth_prior():
    g = small_graphs()["tiny"]
    rng = np.random.default_rng(4)
    g.disps_sens = (g.disps_gt * (rng.uniform(size=g.disps.shape) < 0.5)).astype(np.float32)
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  2, g.lm, g.ep, precision="f64")
    r = run_ba(g, iters=2)
    assert rel(r.disps, o.disps) < STATE_RTOL
    assert rel(sign_normalise(r.poses), sign_normalise(o.poses)) < STATE_RTOL


def test_non_spd_system_gives_zero_update():
    """failed factorisation -> zero pose update, no exception (droid_kernels.cu:1207-1210)"""
    g = small_graphs()["tiny"]
    r = run_ba(g, iters=1, motion_only=True, lm=-2.0, ep=0.0)
    assert r.info.chol_failures == 1
    np.testing.assert_array_equal(r.dx, 0)
    np.testing.assert_array_equal(r.poses, g.poses)


def test_non_spd_at_scale_does_not_hang_and_zeroes_the_update():
    """the failure path runs through the fused tile-column kernels (flags still get raised)"""
    g = synth.make_config("C")
    r = run_ba(g, iters=2, lm=-2.0, ep=0.0, motion_only=True)
    assert r.info.chol_failures == 2
    np.testing.assert_array_equal(r.dx, 0)
    np.testing.assert_array_equal(r.poses, g.poses)
    # and the context recovers: a normal call afterwards is fine
    r2 = run_ba(g)
    assert r2.info.chol_failures == 0 and np.isfinite(r2.poses).all()


def test_hub_keyframe_with_large_out_degree():
    """one keyframe with 40 out-edges: several Gram passes, producer batches > GRAM_PB, hub frame"""
    n = 48
    g = synth.make_graph(n, 2 * (n - 1), 1, iters=2, seed=9, ht=16, wd=24)
    rng = np.random.default_rng(3)
    hub = 5
    extra_j = np.array([j for j in range(n) if abs(j - hub) > 1][:38], dtype=np.int64)
    extra_i = np.full_like(extra_j, hub)
    ii = np.concatenate([g.ii, extra_i, extra_j])
    jj = np.concatenate([g.jj, extra_j, extra_i])
    tg = synth.project(g.poses_gt, g.disps_gt, g.intrinsics, ii, jj) + rng.normal(0, 0.5, (len(ii), 2, 16, 24)).astype(np.float32)
    g.ii, g.jj, g.targets = ii, jj, tg.astype(np.float32)
    g.weights = rng.uniform(0, 1, tg.shape).astype(np.float32)
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  g.iters, g.lm, g.ep, precision="f64")
    r = run_ba(g)
    assert r.info.max_out_degree >= 40
    assert rel(sign_normalise(r.poses), sign_normalise(o.poses)) < STATE_RTOL
    assert rel(r.disps, o.disps) < STATE_RTOL


def _mixed_degree_graph(n, ht, wd, seed):
    """temporal chain plus out-degree classes <= 6, 7..14, 15..21 and > 21 (one hub)"""
    g = synth.make_graph(n, 2 * (n - 1), 1, iters=2, seed=seed, ht=ht, wd=wd, lm=1e-4, ep=0.1)
    rng = np.random.default_rng(seed + 1)
    ei, ej = [g.ii], [g.jj]
    for i in range(0, n, 3):
        want = [3, 9, 17, 19][(i // 3) % 4] if i != 6 else 26
        cand = np.array([j for j in range(n) if abs(j - i) > 1])
        js = rng.choice(cand, size=min(want, len(cand)), replace=False)
        ei.append(np.full(len(js), i)); ej.append(js)
    ii, jj = np.concatenate(ei).astype(np.int64), np.concatenate(ej).astype(np.int64)
    tg = synth.project(g.poses_gt, g.disps_gt, g.intrinsics, ii, jj) + rng.normal(0, 0.5, (len(ii), 2, ht, wd))
    g.ii, g.jj, g.targets = ii, jj, tg.astype(np.float32)
    g.weights = rng.uniform(0, 1, tg.shape).astype(np.float32)
    return g


def _gram_blocks(r):
    K = r.info.n_depth_frames
    deg = np.diff(get_array(r.ctx, _gba.GET_EPTR, (K + 1,), np.int32))
    Rp = 8 * ((6 * deg + 1 + 7) // 8)
    flat = get_array(r.ctx, _gba.GET_GRAM, (int((Rp.astype(np.int64) ** 2).sum()),), np.float32)
    out, off = [], 0
    for k in range(K):
        R = 6 * deg[k] + 1
        blk = np.triu(flat[off:off + Rp[k] ** 2].reshape(Rp[k], Rp[k])[:R, :R]).astype(np.float64)
        blk[R - 1, R - 1] = 0.0  # sum Q w^2: not part of the contract (never read)
        out.append(blk)
        off += Rp[k] ** 2
    return deg, out


@pytest.mark.parametrize("n,ht,wd", [(40, 16, 24), (190, 30, 40)])
def test_tensor_core_gram_matches_cuda_core_gram(n, ht, wd, monkeypatch):
    """k_gram_tc (tcgen05, 3xTF32, drained accumulators) against k_gram (fp32 FFMA) entry by entry, on
    out-degrees below 14 (idle producer warps), 15-21 (two edges per warp), above 21 (k_gram fallback),
    a last tile that is not full (30 x 40 pixels) and more keyframes than SMs; then end to end
    against the fp64 oracle."""
    g = _mixed_degree_graph(n, ht, wd, seed=n)
    monkeypatch.setenv("GBA_GRAM_TC", "0")
    r0 = run_ba(g, iters=1)
    deg0, G0 = _gram_blocks(r0)
    monkeypatch.setenv("GBA_GRAM_TC", "1")
    r1 = run_ba(g, iters=1)
    deg1, G1 = _gram_blocks(r1)
    np.testing.assert_array_equal(deg0, deg1)
    assert deg0.max() > 21 and (deg0 <= 6).any() and ((deg0 > 14) & (deg0 <= 21)).any()
    for k in range(len(G0)):
        scale = np.abs(G0[k]).max() + 1e-30
        # 870 random graphs: worst 4.8e-6 (both sides carry fp32 rounding of ~3000-term sums)
        assert np.abs(G1[k] - G0[k]).max() / scale < 1e-5, (k, deg0[k])
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  g.iters, g.lm, g.ep, precision="f64")
    r = run_ba(g)
    assert rel(sign_normalise(r.poses), sign_normalise(o.poses)) < STATE_RTOL
    assert rel(r.disps, o.disps) < STATE_RTOL
    assert get_array(r.ctx, _gba.GET_WATCHDOG, (256,), np.int32)[0] == 0  # no pipeline wait ever expired


@pytest.mark.parametrize("n,e,r", [(60, 700, 3), (260, 4200, 8), (640, 3860, 3)])
def test_task_graph_factorisation_matches_oracle(n, e, r):
    """k_chain (the whole factorisation as one dependency graph of tile tasks, several band segments and a
    closure block in the larger graph) and k_back against the dense fp64 Cholesky of the oracle"""
    g = synth.make_graph(n, e, r, iters=2, seed=n + e, ht=24, wd=32, lm=1e-5, ep=0.01)
    r1 = run_ba(g)
    assert r1.info.chol_failures == 0 and r1.info.watchdog_trips == 0
    assert r1.info.chain_columns <= r1.info.n_tiles
    if n > 200:
        assert r1.info.chain_columns < r1.info.n_tiles  # nested dissection shortened the chain
        assert r1.info.n_segments > 1
    assert r1.info.kernel_launches < 60
    print("n %d: %d tiles, %d segments x %d tiles, chain of %d tile columns" % (n, r1.info.n_tiles, r1.info.n_segments, r1.info.segment_tiles, r1.info.chain_columns))
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  g.iters, g.lm, g.ep, precision="f64")
    assert rel(sign_normalise(r1.poses), sign_normalise(o.poses)) < STATE_RTOL
    assert rel(r1.disps, o.disps) < STATE_RTOL


@pytest.mark.parametrize("name", ["A", "B", "mixed"])
def test_interleaved_layout_matches_planar(name):
    """targets / weights as [E,ht,wd,2] (what the update network emits) through
    ba_interleaved against the planar [E,2,ht,wd] call: same per-pixel arithmetic --
    on the tensor-core Gram path (out-degree <= 21), the CUDA-core fallback (hub) and an image whose
    last 64-pixel slot is partial (30 x 40)."""
    import droid_backends

    g = _mixed_degree_graph(70, 30, 40, seed=5) if name == "mixed" else small_graphs()[name]
    outs = []
    for il in (False, True):
        d = to_dev(g)
        if il:
            tg = d.targets.permute(0, 2, 3, 1).contiguous()
            wg = d.weights.permute(0, 2, 3, 1).contiguous()
            droid_backends.ba_interleaved(d.poses, d.disps, d.intr, d.sens, tg.unsqueeze(0), wg.unsqueeze(0), d.eta, d.ii,
                                          d.jj, g.t0, g.t1, g.iters, g.lm, g.ep, False, False)
        else:
            droid_backends.ba(d.poses, d.disps, d.intr, d.sens, d.targets, d.weights, d.eta, d.ii, d.jj, g.t0, g.t1,
                              g.iters, g.lm, g.ep, False, False)
        torch.cuda.synchronize()
        outs.append((d.poses.cpu().numpy(), d.disps.cpu().numpy()))
    # per-pixel arithmetic is bit-identical; the sums are not bit-reproducible run to run (fp32 atomics
    # merge the pixel splits of the Gram, fp64 atomics the closure tiles), hence a tolerance well
    # below the parity bound instead of equality
    assert rel(outs[1][0], outs[0][0]) < 1e-5 and rel(outs[1][1], outs[0][1]) < 1e-5
    with pytest.raises(RuntimeError, match="ht,wd,2"):
        d = to_dev(g)
        droid_backends.ba_interleaved(d.poses, d.disps, d.intr, d.sens, d.targets, d.weights, d.eta, d.ii, d.jj, g.t0,
                                      g.t1, 1, g.lm, g.ep, False, False)


def test_argument_errors():
    import droid_backends

    g = small_graphs()["tiny"]
    d = to_dev(g)
    args = [d.poses, d.disps, d.intr, d.sens, d.targets, d.weights, d.eta, d.ii, d.jj, g.t0, g.t1, 1, 1e-4, 0.1, False]
    with pytest.raises(RuntimeError, match="contiguous"):
        bad = list(args)
        bad[4] = d.targets.permute(0, 1, 3, 2)
        droid_backends.ba(*bad)
    with pytest.raises(RuntimeError, match="eta"):
        bad = list(args)
        bad[6] = d.eta[:-1].contiguous()
        droid_backends.ba(*bad)
    with pytest.raises(RuntimeError, match="outside"):
        bad = list(args)
        bad[8] = d.jj + 100
        droid_backends.ba(*bad)
    with pytest.raises(RuntimeError):
        droid_backends.ba(*[a.cpu() if isinstance(a, torch.Tensor) else a for a in args])
    # 16th positional argument (depth_only) is accepted like the reference (depth_video.py:217-218)
    droid_backends.ba(*(args + [False]))


# ------------------------------------------------------------------------------------------
# DSPO stage 2
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["A", "R"])
def test_dspo_matches_reference_golden(name):
    import geom_ba

    g, z, kw = load_golden(name)
    synth.with_mono_prior(g, seed=kw["seed"] + 1)
    d = to_dev(g)
    target = d.targets.permute(0, 2, 3, 1).contiguous()[None]
    weight = d.weights.permute(0, 2, 3, 1).contiguous()[None]
    disps = d.disps[None]
    poses = d.poses[None]
    intr = d.intr[None, None].repeat(1, g.n_frames, 1)
    mono = torch.from_numpy(g.mono_disps).cuda()[None]
    mask = torch.from_numpy(g.valid_depth_mask).cuda()[None]
    scales = torch.from_numpy(g.scales).cuda()[None]
    shifts = torch.from_numpy(g.shifts).cuda()[None]
    d0 = disps.clone()
    for _ in range(2):
        poses, disps, wqs = geom_ba.BA_with_scale_shift(target, weight, d.eta, poses, disps, intr, d.ii, d.jj, mono,
                                                        scales, shifts, mask, 0, g.lm, g.ep, alpha=g.alpha)
        scales, shifts = wqs[..., 0], wqs[..., 1]
    assert torch.equal(d0, d.disps[None])  # functional: caller tensors untouched
    assert rel(disps[0].cpu().numpy(), z["dspo_disps_f64"]) < STATE_RTOL
    assert rel(wqs[0].cpu().numpy(), z["dspo_wqs_f64"]) < STATE_RTOL


def test_dspo_against_oracle_config_b_with_ignored_frames():
    import geom_ba

    g = synth.with_mono_prior(synth.make_config("B"), seed=5)
    tg = np.ascontiguousarray(g.targets.transpose(0, 2, 3, 1))
    wg = np.ascontiguousarray(g.weights.transpose(0, 2, 3, 1))
    o = oracle.dspo(g.poses, g.disps, g.intrinsics, tg, wg, g.eta, g.ii, g.jj, g.mono_disps, g.scales, g.shifts,
                    g.valid_depth_mask, 3, g.lm, g.ep, g.alpha, precision="f64")
    d = to_dev(g)
    _, disps, wqs = geom_ba.BA_with_scale_shift(
        torch.from_numpy(tg).cuda()[None], torch.from_numpy(wg).cuda()[None], d.eta, d.poses[None], d.disps[None],
        d.intr[None, None].repeat(1, g.n_frames, 1), d.ii, d.jj, torch.from_numpy(g.mono_disps).cuda()[None],
        torch.from_numpy(g.scales).cuda()[None], torch.from_numpy(g.shifts).cuda()[None],
        torch.from_numpy(g.valid_depth_mask).cuda()[None], 3, g.lm, g.ep, alpha=g.alpha)
    assert rel(disps[0].cpu().numpy(), o.disps) < STATE_RTOL
    assert rel(wqs[0].cpu().numpy(), np.stack([o.scales, o.shifts], -1)) < STATE_RTOL


def test_dspo_config_d_size_against_oracle():
    """config D of BASELINE.json: DSPO with a synthetic mono prior, 200 keyframes / 3000 edges"""
    import geom_ba

    g = synth.make_config("D")
    tg = np.ascontiguousarray(g.targets.transpose(0, 2, 3, 1))
    wg = np.ascontiguousarray(g.weights.transpose(0, 2, 3, 1))
    o = oracle.dspo(g.poses, g.disps, g.intrinsics, tg, wg, g.eta, g.ii, g.jj, g.mono_disps, g.scales, g.shifts,
                    g.valid_depth_mask, 0, g.lm, g.ep, g.alpha, precision="f64")
    d = to_dev(g)
    _, disps, wqs = geom_ba.BA_with_scale_shift(
        torch.from_numpy(tg).cuda()[None], torch.from_numpy(wg).cuda()[None], d.eta, d.poses[None], d.disps[None],
        d.intr[None, None].repeat(1, g.n_frames, 1), d.ii, d.jj, torch.from_numpy(g.mono_disps).cuda()[None],
        torch.from_numpy(g.scales).cuda()[None], torch.from_numpy(g.shifts).cuda()[None],
        torch.from_numpy(g.valid_depth_mask).cuda()[None], 0, g.lm, g.ep, alpha=g.alpha)
    assert rel(disps[0].cpu().numpy(), o.disps) < STATE_RTOL
    assert rel(wqs[0].cpu().numpy(), np.stack([o.scales, o.shifts], -1)) < STATE_RTOL


# ------------------------------------------------------------------------------------------
# edge sharding (two emulated ranks on one GPU; the NCCL plumbing itself is covered on CPU/gloo)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["A", "B", "C"])
def test_sharded_build_sums_to_single_rank(name):
    """C is large enough (19+ tile rows, band + closure ordering) that ranks only agree on the layout
    of the packed system because ordering and tile structure come from the whole graph."""
    g = small_graphs()[name] if name != "C" else synth.make_config("C")
    one = split_phase(g, iters=2)
    K = len(one.kx)
    m0 = (np.arange(K) % 2 == 0).astype(np.uint8)
    two = split_phase(g, owned_masks=[m0, 1 - m0], iters=2)
    assert rel(two.M, one.M) < 1e-6
    assert rel(two.rhs, one.rhs) < 1e-6
    assert rel(sign_normalise(two.poses), sign_normalise(one.poses)) < 1e-5
    assert rel(two.disps, one.disps) < 1e-5


@pytest.mark.parametrize("world", [2, 4])
def test_segmented_sharded_solve_matches_single_rank(world):
    """every rank eliminates its own band segments, only the closure block + rhs is summed across ranks, the
    closure factorisation is replicated: same state as the single-rank call"""
    g = synth.make_graph(420, 6400, 8, iters=2, seed=77, ht=24, wd=32, lm=1e-5, ep=0.01)
    one = run_ba(g)
    seg = segmented(g, world)
    assert seg is not None and seg.plan.n_segments >= world
    assert all(i.chol_failures == 0 and i.watchdog_trips == 0 for i in seg.infos)
    assert seg.shared_bytes[0] < 0.6 * seg.shared_bytes[1]  # the exchanged range is a fraction of the system
    assert rel(sign_normalise(seg.poses), sign_normalise(one.poses)) < 1e-5
    assert rel(seg.disps, one.disps) < 1e-5
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  g.iters, g.lm, g.ep, precision="f64")
    assert rel(sign_normalise(seg.poses), sign_normalise(o.poses)) < STATE_RTOL
    assert rel(seg.disps, o.disps) < STATE_RTOL


# ------------------------------------------------------------------------------------------
# full BASELINE sizes: properties instead of the (too slow) oracle
# ------------------------------------------------------------------------------------------
def reprojection_cost(g, poses, disps):
    """robust (median) reprojection error in pixels; long-range edges leave the image and would
    dominate a mean"""
    pred = synth.project(poses, disps, g.intrinsics, g.ii, g.jj)
    return float(np.median(np.abs(g.targets - pred)))


@pytest.mark.parametrize("name", ["C"])
def test_full_size_properties(name):
    g = synth.make_config(name)
    c0 = reprojection_cost(g, g.poses, g.disps)
    r = run_ba(g)
    assert r.info.chol_failures == 0
    assert np.isfinite(r.poses).all() and np.isfinite(r.disps).all()
    c1 = reprojection_cost(g, r.poses, r.disps)
    assert c1 < 0.6 * c0, (c0, c1)  # two Gauss-Newton iterations must cut the reprojection error
    assert rel(r.poses[1:, :3], g.poses_gt[1:, :3]) < rel(g.poses[1:, :3], g.poses_gt[1:, :3])
    np.testing.assert_allclose(np.linalg.norm(r.poses[:, 3:], axis=1), 1.0, atol=1e-4)
    # determinism of the state up to atomics-order noise, and edge-permutation invariance
    r2 = run_ba(g)
    assert rel(r2.poses, r.poses) < 1e-5 and rel(r2.disps, r.disps) < 1e-5
    perm = np.random.default_rng(1).permutation(g.n_edges)
    g.ii, g.jj, g.targets, g.weights = g.ii[perm], g.jj[perm], g.targets[perm], g.weights[perm]
    r3 = run_ba(g)
    assert rel(r3.poses, r.poses) < 1e-5 and rel(r3.disps, r.disps) < 1e-5


# ------------------------------------------------------------------------------------------
# frame_distance
# ------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def geomfn():
    """outputs of the reference's own CUDA frame_distance / depth_filter / iproj / projmap, written on a
    B200 by tests/golden/make_golden_cuda.py"""
    import os

    return np.load(os.path.join(helpers.GOLDEN, "geomfn.npz"))


@pytest.mark.parametrize("name,hw", [("A", None), ("R", None), ("C", None)])
def test_frame_distance_matches_oracle_and_reference(name, hw, geomfn):
    import droid_backends

    g, poses, ii, jj = helpers.frame_distance_case(name)  # one keyframe looks the other way: its pairs return 1000
    n = g.n_frames
    want = oracle.frame_distance(poses, g.disps, g.intrinsics, ii, jj, 0.3, precision="f64")
    d = to_dev(g)
    tp = torch.from_numpy(poses).cuda()
    ti, tj = torch.from_numpy(ii).cuda(), torch.from_numpy(jj).cuda()
    for a, b in ((ti, tj), (tj, ti)):  # the reference calls it both ways (depth_video.py:181-188)
        got = droid_backends.frame_distance(tp, d.disps, d.intr, a, b, 0.3).cpu().numpy()
        w = want if a is ti else oracle.frame_distance(poses, g.disps, g.intrinsics, jj, ii, 0.3, precision="f64")
        np.testing.assert_array_equal(got == 1000.0, w == 1000.0)
        assert (w == 1000.0).sum() >= n - 1
        ok = w != 1000.0
        np.testing.assert_allclose(got[ok], w[ok], rtol=1e-4, atol=1e-4)
        r = geomfn["fd_%s_%s" % (name, "ij" if a is ti else "ji")]  # the reference's CUDA kernel, same inputs
        np.testing.assert_array_equal(got == 1000.0, r == 1000.0)
        ok = r != 1000.0
        np.testing.assert_allclose(got[ok], r[ok], rtol=1e-4, atol=1e-4)


# ------------------------------------------------------------------------------------------
# depth_filter + align_scale_and_shift
# ------------------------------------------------------------------------------------------
# The count is a thresholded decision on float32 geometry: an implementation that rounds the
# reprojection differently (FMA contraction) may flip a pixel that sits exactly on the threshold or
# on a pixel boundary.  Bound: at most 0.2 % of the pixels differ, and never by more than one vote.
DEPTH_FILTER_MAX_FLIPS = 2e-3


@pytest.mark.parametrize("n,ht,wd,noise", helpers.DEPTH_FILTER_CASES)
def test_depth_filter_matches_oracle_and_reference(n, ht, wd, noise, geomfn):
    import droid_backends

    poses, disps, intr, inds, th = helpers.depth_filter_case(n, ht, wd, noise)
    want = oracle.depth_filter(poses, disps, intr, inds, th)
    tp, td, ti = (torch.from_numpy(a).cuda() for a in (poses, disps, intr))
    tix, tth = torch.from_numpy(inds).cuda(), torch.from_numpy(th).cuda()
    got = droid_backends.depth_filter(tp, td, ti, tix, tth).cpu().numpy()
    assert got.shape == want.shape and got.dtype == np.float32
    assert want.max() >= 5 and 0.1 < (want > 0).mean()
    diff = np.abs(got - want)
    assert diff.max() <= 1 and (diff > 0).mean() <= DEPTH_FILTER_MAX_FLIPS
    # the poses buffer of the reference is longer than the keyframe count: only disps bounds the neighbours
    tp_long = torch.cat([tp, tp[:5]], 0)
    got2 = droid_backends.depth_filter(tp_long, td, ti, tix, tth).cpu().numpy()
    np.testing.assert_array_equal(got2, got)
    r = geomfn["df_%d" % n].astype(np.float32)  # the reference's CUDA kernel, same inputs
    d = np.abs(got - r)
    assert d.max() <= 1 and (d > 0).mean() <= DEPTH_FILTER_MAX_FLIPS


def test_depth_filter_empty_and_errors():
    import droid_backends
    from helpers import planar_scene

    poses, disps, intr = planar_scene(4, 8, 12)
    tp, td, ti = (torch.from_numpy(a).cuda() for a in (poses, disps, intr))
    out = droid_backends.depth_filter(tp, td, ti, torch.zeros(0, dtype=torch.int64).cuda(), torch.zeros(0).cuda())
    assert out.shape == (0, 8, 12)
    with pytest.raises(ValueError):
        droid_backends.depth_filter(tp, td, ti, torch.zeros(2, dtype=torch.int64).cuda(), torch.zeros(1).cuda())
    with pytest.raises((TypeError, ValueError, RuntimeError)):
        droid_backends.depth_filter(tp.cpu(), td, ti, torch.zeros(1, dtype=torch.int64).cuda(), torch.zeros(1).cuda())


# ------------------------------------------------------------------------------------------
# iproj + projmap
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["A", "C"])
def test_iproj_and_projmap_match_oracle_and_reference(name, geomfn):
    import droid_backends

    g, poses, ks, es = helpers.reproj_case(name)  # keyframe 1 looks the other way: its pairs land behind the camera
    tp, td, ti = (torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (poses, g.disps, g.intrinsics))
    tii, tjj = torch.from_numpy(g.ii).cuda(), torch.from_numpy(g.jj).cuda()
    want = oracle.iproj(poses, g.disps, g.intrinsics, precision="f64")
    got = droid_backends.iproj(tp, td, ti).cpu().numpy()
    assert got.shape == want.shape and got.dtype == np.float32
    np.testing.assert_allclose(got, want, rtol=2e-5, atol=2e-5)
    wc, wv = oracle.projmap(poses, g.disps, g.intrinsics, g.ii, g.jj, precision="f64")
    gc, gv = droid_backends.projmap(tp, td, ti, tii, tjj)
    gc, gv = gc.cpu().numpy(), gv.cpu().numpy()
    assert gc.shape == wc.shape and gv.shape == wv.shape
    assert 0 < (wv == 0).sum() < wv.size  # both branches of the validity test are exercised
    same = gv == wv
    assert same.mean() > 0.9999  # a pixel exactly on Z = 0.25 may round either way
    ok = np.broadcast_to(same & (np.abs(wc[..., :1]) < 1e4), wc.shape)
    np.testing.assert_allclose(gc[ok], wc[ok], rtol=1e-4, atol=2e-3)
    np.testing.assert_array_equal(gc[..., 2], 0.0)
    # the reference's CUDA kernels on the same inputs (every ks-th keyframe / es-th edge is stored)
    np.testing.assert_allclose(got[::ks], geomfn["iproj_%s" % name], rtol=2e-5, atol=2e-5)
    rc, rv = geomfn["projmap_%s_coords" % name], geomfn["projmap_%s_valid" % name].astype(np.float32)
    same = gv[::es] == rv
    assert same.mean() > 0.9999
    ok = np.broadcast_to(same & (np.abs(rc[..., :1]) < 1e4), rc.shape)
    np.testing.assert_allclose(gc[::es][ok], rc[ok], rtol=1e-4, atol=2e-3)
    assert droid_backends.projmap(tp, td, ti, tii[:0], tjj[:0])[0].shape == (0,) + g.disps.shape[1:] + (3,)


def _torch_align(prediction, target, weights):
    """float64 torch restatement of src/utils/common.py:402-438"""
    p, t, w = prediction.double(), target.double(), weights.double()
    a00, a01, a11 = (w * p * p).sum((1, 2)), (w * p).sum((1, 2)), w.sum((1, 2))
    b0, b1 = (w * p * t).sum((1, 2)), (w * t).sum((1, 2))
    det = a00 * a11 - a01 * a01
    s, q = (a11 * b0 - a01 * b1) / det, (-a01 * b0 + a00 * b1) / det
    e = ((s[:, None, None] * p + q[:, None, None] - t).abs() * w).sum((1, 2)) / a11
    return s, q, e


@pytest.mark.parametrize("name", ["A", "B", "C"])
def test_align_scale_shift_matches_oracle(name):
    import common

    g = synth.with_mono_prior(small_graphs()[name] if name == "A" else synth.make_config(name))
    w = g.valid_depth_mask.astype(np.float32)
    s, q, e = oracle.align_scale_and_shift(g.mono_disps, g.disps, w, precision="f64")
    tp, tt, tw = (torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (g.mono_disps, g.disps, w))
    gs, gq, ge = common.align_scale_and_shift(tp, tt, tw)
    np.testing.assert_allclose(gs.cpu().numpy(), s, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(gq.cpu().numpy(), q, rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(ge.cpu().numpy(), e, rtol=1e-4, atol=1e-6)
    # bool weights and the 2-D form of the reference signature (depth_video.py:308), against torch fp64
    gs1, gq1, ge1 = common.align_scale_and_shift(tp[1], tt[1], tw[1] > 0)
    rs, rq, re_ = _torch_align(tp[1:2], tt[1:2], tw[1:2])
    assert abs(gs1.item() - rs.item()) < 1e-4 * abs(rs.item()) + 1e-5
    assert abs(gq1.item() - rq.item()) < 1e-4 * abs(rq.item()) + 1e-5
    assert abs(ge1.item() - re_.item()) < 1e-4 * abs(re_.item()) + 1e-6
    # weights = None means all ones
    gs2, _, _ = common.align_scale_and_shift(tp, tt, None)
    rs2, _, _ = _torch_align(tp, tt, torch.ones_like(tp))
    np.testing.assert_allclose(gs2.cpu().numpy(), rs2.cpu().numpy(), rtol=1e-4, atol=1e-5)


def test_moba_mirror_matches_oracle():
    import geom_ba

    g = small_graphs()["R"]
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  1, 1e-4, 0.1, motion_only=True, mode=oracle.MODE_GEOM_PY, precision="f64")
    d = to_dev(g)
    target = d.targets.permute(0, 2, 3, 1).contiguous()[None]
    weight = d.weights.permute(0, 2, 3, 1).contiguous()[None]
    poses = geom_ba.MoBA(target, weight, None, d.poses[None], d.disps[None], d.intr[None, None].repeat(1, g.n_frames, 1),
                         d.ii, d.jj, fixedp=g.t0)
    assert rel(sign_normalise(poses[0].cpu().numpy()), sign_normalise(o.poses)) < STATE_RTOL


# ------------------------------------------------------------------------------------------
# THE parity gate of the default semantics: the reference's own CUDA droid_backends.ba.  The fixtures
# were written on a B200 by the unmodified reference extension (tests/golden/make_golden_cuda.py):
# configs A, B, C and E of BASELINE.json, the ragged / tiny graphs, pixels on every side of the
# MIN_DEPTH test, stereo self edges, rotation updates around the 1e-4 thresholds of the hand-rolled
# exponential, motion_only, depth_only and the sensor-depth prior.
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", helpers.DROID_CASES)
def test_droid_mode_matches_reference_cuda_golden(name):
    g, kw, stride, z = helpers.load_droid_golden(name)
    r = run_ba(g, **kw)
    assert r.info.chol_failures == 0 and r.info.watchdog_trips == 0
    assert rel(sign_normalise(r.poses), sign_normalise(z["poses"])) < STATE_RTOL
    assert rel(r.disps[::stride], z["disps"]) < STATE_RTOL
    # the returned updates of the last iteration (small numbers: looser bound, informative only)
    tol = 2e-2 if name == "step4" else 2e-3
    assert rel(r.dx, z["dx"]) < tol
    if "dz" in z.files:
        assert r.dz.shape == z["dz"].shape and rel(r.dz, z["dz"]) < tol
    if kw.get("motion_only"):
        np.testing.assert_array_equal(r.disps, g.disps)
    if kw.get("depth_only"):
        np.testing.assert_array_equal(r.poses, g.poses)


def test_config_c_matches_fp64_oracle():
    """config C (200 keyframes / 3000 edges / 48 x 64, BASELINE.json configs[2]) against the fp64 oracle"""
    g = synth.make_config("C")
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  g.iters, g.lm, g.ep, precision="f64")
    r = run_ba(g)
    assert rel(sign_normalise(r.poses), sign_normalise(o.poses)) < STATE_RTOL
    assert rel(r.disps, o.disps) < STATE_RTOL


def test_dspo_clamps_every_row_like_the_reference():
    """golden = the reference's BA_with_scale_shift on negative disparities, some on a keyframe that has no
    out-edge: ba.py:214 clamps the whole tensor"""
    import geom_ba

    g, z, _ = load_golden("dspo_neg")
    d = to_dev(g)
    tg = d.targets.permute(0, 2, 3, 1).contiguous()[None]
    wg = d.weights.permute(0, 2, 3, 1).contiguous()[None]
    disps, poses = d.disps[None], d.poses[None]
    intr = d.intr[None, None].repeat(1, g.n_frames, 1)
    mono = torch.from_numpy(g.mono_disps).cuda()[None]
    mask = torch.from_numpy(g.valid_depth_mask).cuda()[None]
    scales, shifts = torch.from_numpy(g.scales).cuda()[None], torch.from_numpy(g.shifts).cuda()[None]
    for _ in range(2):
        poses, disps, wqs = geom_ba.BA_with_scale_shift(tg, wg, d.eta, poses, disps, intr, d.ii, d.jj, mono, scales,
                                                        shifts, mask, 0, g.lm, g.ep, alpha=g.alpha)
        scales, shifts = wqs[..., 0], wqs[..., 1]
    out = disps[0].cpu().numpy()
    assert out.min() == 0.0 and (out[4] == 0).any()
    assert rel(out, z["dspo_disps_f64"]) < STATE_RTOL
    assert rel(wqs[0].cpu().numpy(), z["dspo_wqs_f64"]) < STATE_RTOL


def test_geom_mode_clamps_every_row():
    """src/geom/ba.py:119: disps.clamp(min=0) acts on every keyframe, also one outside unique(ii)"""
    g = synth.make_graph(7, 16, 2, iters=1, seed=41, ht=16, wd=24)
    keep = g.ii != 4
    g.ii, g.jj, g.targets, g.weights = g.ii[keep], g.jj[keep], g.targets[keep], g.weights[keep]
    g.eta = g.eta[: len(np.unique(g.ii))]
    g.disps[4, :3] = -0.2
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  1, g.lm, g.ep, mode=oracle.MODE_GEOM_PY, precision="f64")
    r = run_ba(g, mode="geom_py", iters=1)
    assert (o.disps[4, :3] == 0).all() and (r.disps[4, :3] == 0).all()
    assert rel(r.disps, o.disps) < STATE_RTOL


def test_deterministic_switch(monkeypatch):
    """GBA_DETERMINISTIC=1 removes the fp32 atomics (pixel splits of the Gram kernels): what is left are unordered
    fp64 sums, far below fp32 resolution -- two runs then give the same fp32 state to 1e-7 (bitwise in practice),
    and the same state as the default path within the noise allowance the other tests use (1e-5)"""
    g = small_graphs()["B"]  # 20 keyframes: the default path splits every keyframe's pixels over several CTAs
    monkeypatch.setenv("GBA_DETERMINISTIC", "1")
    a, b = run_ba(g), run_ba(g)
    monkeypatch.delenv("GBA_DETERMINISTIC")
    c = run_ba(g)
    assert rel(a.poses, b.poses) < 1e-7 and rel(a.disps, b.disps) < 1e-7
    assert rel(c.poses, a.poses) < 1e-5 and rel(c.disps, a.disps) < 1e-5
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.t0, g.t1,
                  g.iters, g.lm, g.ep, precision="f64")
    assert rel(sign_normalise(a.poses), sign_normalise(o.poses)) < STATE_RTOL and rel(a.disps, o.disps) < STATE_RTOL
    print("deterministic runs bitwise equal:", np.array_equal(a.poses, b.poses) and np.array_equal(a.disps, b.disps))


def test_spine_placement_variants(monkeypatch):
    """k_chain picks its critical-path CTAs by SM when the grid fills the machine (default) or by block index
    (GBA_CHAIN_SPINE_BY_BLOCK=1, also what smaller grids use): same factorisation, so the same state up to the
    order of the fp64 atomics"""
    g = synth.make_config("C")
    a = run_ba(g)
    monkeypatch.setenv("GBA_CHAIN_SPINE_BY_BLOCK", "1")
    b = run_ba(g)
    monkeypatch.delenv("GBA_CHAIN_SPINE_BY_BLOCK")
    assert a.info.chol_failures == 0 and b.info.chol_failures == 0
    assert rel(a.poses, b.poses) < 1e-6 and rel(a.disps, b.disps) < 1e-6


# ------------------------------------------------------------------------------------------
# boundary hardening
# ------------------------------------------------------------------------------------------
def test_misaligned_pointers_are_refused():
    """every per-pixel array is read with 16-byte vector loads: a view that starts 4 bytes into an allocation
    must come back as GBA_EINVAL, not as a misaligned-address fault"""
    import droid_backends

    g = small_graphs()["tiny"]
    d = to_dev(g)
    args = [d.poses, d.disps, d.intr, d.sens, d.targets, d.weights, d.eta, d.ii, d.jj, g.t0, g.t1, 1, 1e-4, 0.1, False]
    for idx in (1, 4, 5, 6):
        t = args[idx]
        shifted = torch.zeros(t.numel() + 1, dtype=t.dtype, device=t.device)[1:].view(t.shape)
        shifted.copy_(t)
        assert shifted.is_contiguous() and shifted.data_ptr() % 16 == 4
        bad = list(args)
        bad[idx] = shifted
        with pytest.raises(RuntimeError, match="aligned"):
            droid_backends.ba(*bad)
    droid_backends.ba(*args)  # and the context still works
    torch.cuda.synchronize()
    L = _gba.lib()
    out = torch.zeros(g.n_edges, 32, device="cuda")
    off = torch.zeros(d.targets.numel() + 1, device="cuda")[1:].view(d.targets.shape)
    rc = L.gba_edge_blocks_static(d.poses.data_ptr(), d.disps.data_ptr(), d.intr.data_ptr(), off.data_ptr(),
                                  d.weights.data_ptr(), d.ii.data_ptr(), d.jj.data_ptr(), g.n_frames, g.n_edges,
                                  g.disps.shape[1], g.disps.shape[2], 0, out.data_ptr(), None)
    assert rc == -1 and b"aligned" in L.gba_last_error(None)


def test_sharded_begin_rejects_a_local_edge_missing_from_the_whole_graph():
    """gba_ba_begin derives the tile structure from ii_all / jj_all; a local edge that is not part of it would
    add into a tile that does not exist.  It must be refused (GBA_EINVAL), not written somewhere."""
    g = synth.make_config("C")  # large enough that a far edge needs a tile of its own
    L = _gba.lib()
    N, ht, wd = g.disps.shape
    kx = np.unique(np.concatenate([np.arange(g.t0, g.t1), g.ii]))
    far = (np.abs(g.ii - g.jj) > 60)
    assert far.any()
    keep = ~far  # the "whole graph" the caller declares lacks the long-range edges ...
    d = to_dev(g)  # ... but the local edge list still has them
    h = ctypes.c_void_p()
    assert L.gba_create(ctypes.byref(h), 0, N, g.n_edges, ht, wd) == 0
    ii_all, jj_all = torch.from_numpy(g.ii[keep]).cuda(), torch.from_numpy(g.jj[keep]).cuda()
    pad = g.n_edges - int(keep.sum())
    ii_all = torch.cat([ii_all, ii_all[:pad]]).contiguous()  # n_all >= n_edges, yet not a superset
    jj_all = torch.cat([jj_all, jj_all[:pad]]).contiguous()
    kx_t, own = torch.from_numpy(kx).cuda(), torch.ones(len(kx), dtype=torch.uint8, device="cuda")
    rc = L.gba_ba_begin(h, d.poses.data_ptr(), d.disps.data_ptr(), d.intr.data_ptr(), d.sens.data_ptr(),
                        d.targets.data_ptr(), d.weights.data_ptr(), d.eta.data_ptr(), d.ii.data_ptr(), d.jj.data_ptr(),
                        N, g.n_edges, g.eta.shape[0], g.t0, g.t1, g.lm, g.ep, 0, 0, 0, kx_t.data_ptr(), len(kx),
                        own.data_ptr(), ii_all.data_ptr(), jj_all.data_ptr(), int(ii_all.shape[0]), None)
    assert rc == -1 and b"ii_all" in L.gba_last_error(h)
    # an out-of-range index in the whole-graph list is still reported although the CSR is rebuilt afterwards
    bad = torch.from_numpy(g.ii).cuda().clone()
    bad[7] = N + 5
    rc = L.gba_ba_begin(h, d.poses.data_ptr(), d.disps.data_ptr(), d.intr.data_ptr(), d.sens.data_ptr(),
                        d.targets.data_ptr(), d.weights.data_ptr(), d.eta.data_ptr(), d.ii.data_ptr(), d.jj.data_ptr(),
                        N, g.n_edges, g.eta.shape[0], g.t0, g.t1, g.lm, g.ep, 0, 0, 0, kx_t.data_ptr(), len(kx),
                        own.data_ptr(), bad.data_ptr(), d.jj.data_ptr(), g.n_edges, None)
    assert rc == -1 and b"outside" in L.gba_last_error(h)
    L.gba_destroy(h)


# ------------------------------------------------------------------------------------------
# the call site: DepthVideo.ba / .dspo (src/depth_video.py:198-298) driving the drop-ins
# ------------------------------------------------------------------------------------------
def test_dispatcher_pose_depth_branch():
    """depth_video.py:214-220: view/permute/contiguous, the 16 positional arguments, clamp(min=1e-5)"""
    from stub_depth_video import StubDepthVideo

    g, poses, disps, mono = helpers.dispatcher_case(0)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    v = StubDepthVideo(t(poses), t(disps), t(g.intrinsics), t(mono), g.n_frames, ba_type="DBA")
    target = t(g.targets).permute(0, 2, 3, 1).contiguous()[None]  # [1,E,ht,wd,2] as the update operator emits it
    weight = t(g.weights).permute(0, 2, 3, 1).contiguous()[None]
    v.ba(target, weight, t(g.eta), t(g.ii), t(g.jj), t0=g.t0, t1=None, iters=2, lm=g.lm, ep=g.ep)
    assert v.calls == ["pose_depth"]
    o = oracle.ba(poses, disps, g.intrinsics, np.zeros_like(disps), g.targets, g.weights, g.eta, g.ii, g.jj, g.t0,
                  g.n_frames, 2, g.lm, g.ep, precision="f64")
    assert rel(sign_normalise(v.poses.cpu().numpy()), sign_normalise(o.poses)) < STATE_RTOL
    assert rel(v.disps.cpu().numpy(), np.maximum(o.disps, 1e-5)) < STATE_RTOL
    assert v.disps.min().item() >= float(np.float32(1e-5))  # rows beyond the keyframes in use were 0 and are clamped too


def test_dispatcher_depth_scale_branch():
    """depth_video.py:221-288: depth_filter -> valid mask -> align_scale_and_shift -> mono_thres edge filter ->
    itrs x BA_with_scale_shift -> rebind, against the same pipeline on the CPU oracle"""
    from stub_depth_video import StubDepthVideo

    g, poses, disps, mono = helpers.dispatcher_case(1)
    n = g.n_frames
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    v = StubDepthVideo(t(poses), t(disps), t(g.intrinsics), t(mono), n, mono_thres=0.5)
    target = t(g.targets).permute(0, 2, 3, 1).contiguous()[None]
    weight = t(g.weights).permute(0, 2, 3, 1).contiguous()[None]
    kx_all = np.unique(g.ii)
    eta = g.eta[: len(kx_all)]
    v.ba(target, weight, t(eta), t(g.ii), t(g.jj), t0=g.t0, t1=None, iters=2, lm=g.lm, ep=g.ep, opt_type="depth_scale")
    assert v.calls == ["depth_scale"]
    # the device versions of the two Python loops of this branch (video_ops) against what the torch statements gave
    import video_ops

    m_dev, cnt_dev, mean_dev = video_ops.valid_depth_mask(t(poses), t(disps), t(g.intrinsics), torch.arange(n).cuda(),
                                                          v.mv_thresh, v.visible_num)
    assert torch.equal(m_dev, v.valid_depth_mask_small[:n])
    assert torch.equal(cnt_dev.long(), v.valid_depth_mask_small[:n].sum(dim=[1, 2]))
    # ---- the same steps on the CPU oracle ----
    depths = 1.0 / disps[:n].astype(np.float64)
    thresh = (v.mv_thresh * depths.mean(axis=(1, 2))).astype(np.float32)
    cnt = oracle.depth_filter(poses, disps, g.intrinsics, np.arange(n), thresh)
    dd = np.where(cnt >= v.visible_num, (1.0 / disps[:n]).astype(np.float32), np.nan)
    med = np.array([np.sort(r[~np.isnan(r)])[(np.sum(~np.isnan(r)) - 1) // 2] if (~np.isnan(r)).any() else np.nan
                    for r in dd.reshape(n, -1)], dtype=np.float32)  # torch.nanmedian takes the lower middle
    mask = dd < 3 * med[:, None, None]
    got_mask = v.valid_depth_mask_small[:n].cpu().numpy()
    assert (got_mask != mask).mean() < 2e-3  # depth_filter votes may flip on threshold-borderline pixels
    s, q, e = oracle.align_scale_and_shift(mono[:n], disps[:n], got_mask.astype(np.float32), precision="f64")
    avg = disps[:n].mean(axis=(1, 2))
    with np.errstate(invalid="ignore", divide="ignore"):
        invalid = (e / avg > v.mono_thres) | np.isnan(e) | (s < 0) | (got_mask.sum(axis=(1, 2)) < 0.5 * got_mask[0].size)
    assert invalid[5] and not invalid.all()
    drop = np.isin(g.ii, np.where(invalid)[0]) | np.isin(g.jj, np.where(invalid)[0])
    ii_t, jj_t = g.ii[~drop], g.jj[~drop]
    np.testing.assert_array_equal(v.last_filter[0].cpu().numpy(), ii_t)
    np.testing.assert_array_equal(v.last_filter[1].cpu().numpy(), jj_t)
    inv_dev, keep_dev, src_dev = video_ops.mono_filter(t(e.astype(np.float32)), mean_dev, t(s.astype(np.float32)), cnt_dev,
                                                       g.disps.shape[1], g.disps.shape[2], v.mono_thres, t(g.ii), t(g.jj),
                                                       poses.shape[0])
    np.testing.assert_array_equal(inv_dev.cpu().numpy(), invalid)
    np.testing.assert_array_equal(keep_dev.cpu().numpy(), ~drop)
    eta_t = eta[np.isin(kx_all, np.unique(ii_t))]
    assert v.last_filter[2] == eta_t.shape[0]
    tg = np.ascontiguousarray(g.targets.transpose(0, 2, 3, 1))[~drop]
    wg = np.ascontiguousarray(g.weights.transpose(0, 2, 3, 1))[~drop]
    buf = poses.shape[0]
    sc, sh = np.ones(buf), np.zeros(buf)
    sc[:n], sh[:n] = s, q
    maskbuf = np.zeros(disps.shape, dtype=bool)
    maskbuf[:n] = got_mask
    dz = disps
    for _ in range(2):
        o = oracle.dspo(poses, dz, g.intrinsics, tg, wg, eta_t, ii_t, jj_t, mono, sc, sh, maskbuf, 0, g.lm, g.ep, 0.01,
                        interleaved=True, precision="f64")
        dz, sc, sh = o.disps, o.scales, o.shifts
    assert rel(v.disps.cpu().numpy(), np.maximum(dz, 1e-5)) < STATE_RTOL
    ok = ~invalid
    assert rel(v.depth_scale.cpu().numpy()[:n][ok], sc[:n][ok]) < STATE_RTOL
    assert rel(v.depth_shift.cpu().numpy()[:n][ok], sh[:n][ok]) < 1e-3  # shifts are ~0.05: absolute scale of the scales
    np.testing.assert_array_equal(v.poses.cpu().numpy(), poses)  # stage 2 leaves the poses alone
