"""CPU: the C oracle (oracle/ba_oracle.c) against golden vectors produced by the reference
itself (tests/golden/make_golden.py imports reference src/geom/ba.py on CPU)."""
import numpy as np
import pytest

import synth
import helpers
from helpers import STATE_RTOL, load_golden, rel, sign_normalise
from oracle import oracle


# ------------------------------------------------------------------------------------------
# droid_cuda mode: fixtures written ON A B200 by the reference's own CUDA droid_backends.ba,
# frame_distance, depth_filter, iproj, projmap (tests/golden/make_golden_cuda.py).  Config E is
# left to the GPU suite: the dense fp64 oracle needs minutes for 5994 unknowns.
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", [n for n in helpers.DROID_CASES if n != "E"])
def test_ba_droid_mode_matches_reference_cuda(name):
    g, kw, stride, z = helpers.load_droid_golden(name)
    for prec in ("f64", "f32"):
        o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj,
                      g.t0, g.t1, g.iters, g.lm, g.ep, precision=prec, **kw)
        assert o.fail == 0
        # the reference is fp32 with fp32 block reductions: 2.6e-6 is the worst seen (config C)
        assert rel(sign_normalise(o.poses), sign_normalise(z["poses"])) < 1e-5, prec
        assert rel(o.disps[::stride], z["disps"]) < 1e-5, prec
        # updates of the last iteration: small numbers, hence the looser bound (step4: dx ~ 1e-6)
        assert rel(o.dx, z["dx"]) < (5e-3 if name == "step4" else 5e-4), prec
        if "dz" in z.files:
            assert rel(o.dz, z["dz"]) < (5e-3 if name == "step4" else 5e-4), prec
    if kw.get("motion_only"):
        np.testing.assert_array_equal(z["disps"], g.disps)  # the reference leaves disparities alone
    if kw.get("depth_only"):
        np.testing.assert_array_equal(z["poses"], g.poses)


@pytest.fixture(scope="module")
def geomfn():
    return np.load(helpers.os.path.join(helpers.GOLDEN, "geomfn.npz"))


@pytest.mark.parametrize("name", ["A", "R", "C"])
def test_frame_distance_matches_reference_cuda(name, geomfn):
    g, poses, ii, jj = helpers.frame_distance_case(name)
    np.testing.assert_allclose([poses.astype(np.float64).sum(), g.disps.astype(np.float64).sum()],
                               geomfn["fd_%s_checksum" % name], rtol=1e-9)
    for key, a, b in (("ij", ii, jj), ("ji", jj, ii)):
        ref = geomfn["fd_%s_%s" % (name, key)]
        got = oracle.frame_distance(poses, g.disps, g.intrinsics, a, b, 0.3, precision="f64")
        np.testing.assert_array_equal(got == 1000.0, ref == 1000.0)
        ok = ref != 1000.0
        np.testing.assert_allclose(got[ok], ref[ok], rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("case", helpers.DEPTH_FILTER_CASES)
def test_depth_filter_matches_reference_cuda(case, geomfn):
    poses, disps, intr, inds, th = helpers.depth_filter_case(*case)
    ref = geomfn["df_%d" % case[0]].astype(np.float32)
    got = oracle.depth_filter(poses, disps, intr, inds, th)
    diff = np.abs(got - ref)
    # thresholded decisions on fp32 geometry: a pixel exactly on a threshold may flip with the rounding
    assert diff.max() <= 1 and (diff > 0).mean() <= 2e-3


@pytest.mark.parametrize("name", ["A", "C"])
def test_iproj_projmap_match_reference_cuda(name, geomfn):
    g, poses, ks, es = helpers.reproj_case(name)
    got = oracle.iproj(poses, g.disps, g.intrinsics, precision="f64")[::ks]
    np.testing.assert_allclose(got, geomfn["iproj_%s" % name], rtol=2e-5, atol=2e-5)
    c, v = oracle.projmap(poses, g.disps, g.intrinsics, g.ii[::es], g.jj[::es], precision="f64")
    rc, rv = geomfn["projmap_%s_coords" % name], geomfn["projmap_%s_valid" % name].astype(np.float64)
    same = v == rv
    assert same.mean() > 0.9999 and 0 < (rv == 0).sum() < rv.size
    ok = np.broadcast_to(same & (np.abs(rc[..., :1]) < 1e4), rc.shape)
    np.testing.assert_allclose(c[ok], rc[ok], rtol=1e-4, atol=2e-3)


@pytest.mark.parametrize("name", ["A", "R", "near", "step6"])
def test_ba_geom_mode_matches_reference_fp64(name):
    g, z, kw = load_golden(name)
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj,
                  g.t0, g.t1, g.iters, g.lm, g.ep, mode=oracle.MODE_GEOM_PY, precision="f64")
    assert o.fail == 0
    # same algorithm in the same precision: far inside the 1e-4 bar
    assert rel(sign_normalise(o.poses), sign_normalise(z["poses_f64"])) < 1e-8
    assert rel(o.disps, z["disps_f64"].reshape(o.disps.shape)) < 1e-7


def test_edge_case_graphs_reach_the_branches():
    """near: pixels on every side of Z = 0.1 / 0.2 / 0.25 and behind the camera (projective_ops.py:44,112,
    droid_kernels.cu:302-306); step4 / step6: rotation updates on both sides of 1e-4 (droid_kernels.cu:119,162)
    and of lietorch's EPS 1e-6 (so3.h:153-170); stereo: self edges present"""
    Z = synth.reprojected_depth(synth.make_edge_case("near"))
    for lo, hi in ((-1e9, 0.0), (0.0, 0.1), (0.1, 0.2), (0.2, 0.25), (0.25, 1e9)):
        assert ((Z >= lo) & (Z < hi)).sum() >= 20, (lo, hi)
    for name, thr, mode in (("step4", 1e-4, oracle.MODE_DROID_CUDA), ("step6", 1e-6, oracle.MODE_GEOM_PY)):
        g = synth.make_edge_case(name)
        o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj,
                      g.t0, g.t1, 1, g.lm, g.ep, mode=mode, precision="f64")
        th = np.linalg.norm(o.dx[:, 3:], axis=1)
        assert th.min() < thr < th.max(), (name, th.min(), th.max())
    g = synth.make_edge_case("stereo")
    assert (g.ii == g.jj).sum() == 3


@pytest.mark.parametrize("name", ["A", "R", "near", "step6"])
def test_ba_geom_mode_fp32_within_bar(name):
    g, z, kw = load_golden(name)
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj,
                  g.t0, g.t1, g.iters, g.lm, g.ep, mode=oracle.MODE_GEOM_PY, precision="f32")
    for ref in ("f32", "f64"):
        assert rel(sign_normalise(o.poses), sign_normalise(z["poses_" + ref])) < STATE_RTOL
        assert rel(o.disps, z["disps_" + ref].reshape(o.disps.shape)) < STATE_RTOL


@pytest.mark.parametrize("name", ["A", "R"])
def test_dspo_matches_reference(name):
    g, z, kw = load_golden(name)
    synth.with_mono_prior(g, seed=kw["seed"] + 1)
    tg = np.ascontiguousarray(g.targets.transpose(0, 2, 3, 1))
    wg = np.ascontiguousarray(g.weights.transpose(0, 2, 3, 1))
    for prec, tol in (("f64", 1e-6), ("f32", STATE_RTOL)):
        disps, sc, sh = g.disps, g.scales, g.shifts
        for _ in range(2):
            o = oracle.dspo(g.poses, disps, g.intrinsics, tg, wg, g.eta, g.ii, g.jj, g.mono_disps, sc, sh,
                            g.valid_depth_mask, 0, g.lm, g.ep, g.alpha, interleaved=True, precision=prec)
            assert o.fail == 0
            disps, sc, sh = o.disps, o.scales, o.shifts
        wq = np.stack([sc, sh], -1)
        assert rel(disps, z["dspo_disps_f64"]) < tol
        assert rel(wq, z["dspo_wqs_f64"]) < tol


def test_dspo_clamps_the_whole_tensor_like_the_reference():
    """negative disparities, also on a keyframe outside unique(ii): ba.py:214 clamps everything"""
    g, z, _ = load_golden("dspo_neg")
    assert (g.disps[4] < 0).any() and 4 not in set(g.ii.tolist())
    tg = np.ascontiguousarray(g.targets.transpose(0, 2, 3, 1))
    wg = np.ascontiguousarray(g.weights.transpose(0, 2, 3, 1))
    disps, sc, sh = g.disps, g.scales, g.shifts
    for _ in range(2):
        o = oracle.dspo(g.poses, disps, g.intrinsics, tg, wg, g.eta, g.ii, g.jj, g.mono_disps, sc, sh,
                        g.valid_depth_mask, 0, g.lm, g.ep, g.alpha, interleaved=True, precision="f64")
        disps, sc, sh = o.disps, o.scales, o.shifts
    assert z["dspo_disps_f64"].min() == 0.0 and (z["dspo_disps_f64"][4] == 0).any()
    assert rel(disps, z["dspo_disps_f64"]) < 1e-6
    assert rel(np.stack([sc, sh], -1), z["dspo_wqs_f64"]) < 1e-6


def test_planar_and_interleaved_layouts_agree():
    g = synth.with_mono_prior(synth.make_graph(5, 8, 1, seed=2, ht=8, wd=12))
    tg = np.ascontiguousarray(g.targets.transpose(0, 2, 3, 1))
    wg = np.ascontiguousarray(g.weights.transpose(0, 2, 3, 1))
    a = oracle.dspo(g.poses, g.disps, g.intrinsics, tg, wg, g.eta, g.ii, g.jj, g.mono_disps, g.scales, g.shifts,
                    g.valid_depth_mask, 0, 1e-4, 0.1, 0.01, interleaved=True)
    b = oracle.dspo(g.poses, g.disps, g.intrinsics, g.targets, g.weights, g.eta, g.ii, g.jj, g.mono_disps,
                    g.scales, g.shifts, g.valid_depth_mask, 0, 1e-4, 0.1, 0.01, interleaved=False)
    np.testing.assert_array_equal(a.disps, b.disps)


def test_modes_differ_as_surveyed():
    """the two reference implementations do not agree to 1e-4 (D1, D2)."""
    g, z, kw = load_golden("A")
    a = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj,
                  g.t0, g.t1, 1, g.lm, g.ep, mode=oracle.MODE_DROID_CUDA, precision="f64")
    b = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj,
                  g.t0, g.t1, 1, g.lm, g.ep, mode=oracle.MODE_GEOM_PY, precision="f64")
    assert rel(a.dx, b.dx) > 1e-3
    assert rel(a.dz, b.dz) > 1e-3


def test_index_structures_and_flags():
    g = synth.make_graph(7, 12, 1, seed=5, ht=8, wd=12, t0=2)
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj,
                  g.t0, g.t1, 1, 1e-4, 0.1, mode=oracle.MODE_DROID_CUDA, debug=True)
    kx = np.unique(np.concatenate([np.arange(g.t0, g.t1), g.ii]))
    np.testing.assert_array_equal(o.kx, kx)
    assert o.dx.shape == (g.t1 - g.t0, 6) and o.dz.shape == (len(kx), g.hw)
    # symmetric up to fp32 rounding of (E_r Q) E_s^T vs (E_s Q) E_r^T, as in the reference
    np.testing.assert_allclose(o.M, o.M.T, rtol=1e-4, atol=1e-5)
    # motion_only leaves disparities alone; depth_only leaves poses alone
    m = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj,
                  g.t0, g.t1, 2, 1e-4, 0.1, motion_only=True)
    np.testing.assert_array_equal(m.disps, g.disps)
    assert rel(m.poses, g.poses) > 0
    d = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj,
                  g.t0, g.t1, 2, 1e-4, 0.1, depth_only=True)
    np.testing.assert_array_equal(d.poses, g.poses)
    assert rel(d.disps, g.disps) > 0
    # fixed poses never move
    np.testing.assert_array_equal(o.poses[: g.t0], g.poses[: g.t0])


def test_non_spd_gives_zero_update():
    """Reference contract: failed factorisation -> zero pose update (droid_kernels.cu:1207-1210)."""
    g = synth.make_graph(5, 8, 1, seed=7, ht=8, wd=12)
    o = oracle.ba(g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj,
                  g.t0, g.t1, 1, -2.0, 0.0, motion_only=True)  # lm = -2 flips the diagonal sign
    assert o.fail == 1
    np.testing.assert_array_equal(o.dx, 0)
    np.testing.assert_array_equal(o.poses, g.poses)


def test_frame_distance_oracle_basics():
    """droid_kernels.cu:518-657: zero on the diagonal, positive elsewhere, 1000 when the target looks away"""
    g = synth.make_graph(6, 10, 1, seed=4, ht=16, wd=24)
    ii, jj = np.meshgrid(np.arange(6), np.arange(6), indexing="ij")
    d = oracle.frame_distance(g.poses, g.disps, g.intrinsics, ii.reshape(-1), jj.reshape(-1), 0.3).reshape(6, 6)
    np.testing.assert_allclose(np.diag(d), 0.0, atol=1e-4)
    assert (d[~np.eye(6, dtype=bool)] > 0).all() and (d < 1000).all()
    poses = g.poses.copy()
    poses[3, 3:] = [0.0, 1.0, 0.0, 0.0]  # 180 degrees about y: everything ends up behind camera 3
    d2 = oracle.frame_distance(poses, g.disps, g.intrinsics, np.array([0]), np.array([3]), 0.3)
    assert d2[0] == 1000.0
    d64 = oracle.frame_distance(g.poses, g.disps, g.intrinsics, ii.reshape(-1), jj.reshape(-1), 0.3, precision="f64")
    assert rel(d.reshape(-1), d64) < 1e-5


def test_depth_filter_oracle_on_consistent_scene():
    """droid_kernels.cu:661-775: with exact multi-view-consistent depth every neighbour that sees the
    point agrees; a keyframe with wrong depth collects no votes and gives none."""
    from helpers import planar_scene

    n, ht, wd = 12, 24, 32
    poses, disps, intr = planar_scene(n, ht, wd)
    thresh = 0.02 * (1.0 / disps).mean(axis=(1, 2))
    cnt = oracle.depth_filter(poses, disps, intr, np.arange(n), thresh)
    assert cnt.shape == (n, ht, wd) and cnt.min() >= 0 and cnt.max() == 6
    # neighbours that exist: -1,-2,-3,+3,+4,+5
    exist = np.array([sum(0 <= k + o < n for o in (-1, -2, -3, 3, 4, 5)) for k in range(n)])
    assert (cnt.reshape(n, -1).max(axis=1) == exist).all()
    # interior pixels stay inside every neighbour's image: full count there
    assert (cnt[5, 8:16, 10:22] == 6).all()
    bad = disps.copy()
    bad[5] *= 2.0
    cnt2 = oracle.depth_filter(poses, bad, intr, np.arange(n), thresh)
    assert cnt2[5].max() == 0
    assert (cnt2[6] <= cnt[6]).all() and cnt2[6].sum() < cnt[6].sum()  # 6 lost neighbour 5 (= 6-1)
    np.testing.assert_array_equal(cnt2[9], cnt[9])  # 9 looks at 8, 7, 6, 12.., never at 5
    sub = oracle.depth_filter(poses, disps, intr, np.array([7, 2]), thresh[[7, 2]])
    np.testing.assert_array_equal(sub, cnt[[7, 2]])
    c64 = oracle.depth_filter(poses, disps, intr, np.arange(n), thresh, precision="f64")
    assert (c64 != cnt).mean() < 2e-3


def test_align_scale_shift_oracle_recovers_affine():
    """src/utils/common.py:402-438: target = 1.7 * prediction - 0.3 on the weighted pixels"""
    rng = np.random.default_rng(3)
    p = rng.uniform(0.2, 2.0, (5, 16, 24)).astype(np.float32)
    w = (rng.uniform(size=p.shape) > 0.3).astype(np.float32)
    t = (1.7 * p - 0.3).astype(np.float32)
    t[w == 0] = 99.0  # ignored pixels
    s, q, e = oracle.align_scale_and_shift(p, t, w)
    np.testing.assert_allclose(s, 1.7, rtol=1e-4)
    np.testing.assert_allclose(q, -0.3, atol=1e-4)
    assert e.max() < 1e-4
    t2 = t + (rng.normal(0, 0.05, p.shape) * w).astype(np.float32)
    s2, q2, e2 = oracle.align_scale_and_shift(p, t2, w, precision="f64")
    assert 0.02 < e2.mean() < 0.06 and abs(s2 - 1.7).max() < 0.05


def test_iproj_projmap_oracle_basics():
    """droid_kernels.cu:779-850 and :427-516: the identity pose gives (x, y, 1) / d; a keyframe maps onto
    itself; projmap agrees with the reprojection that generated the synthetic targets"""
    import synth

    g = synth.make_config("A")
    pts = oracle.iproj(g.poses, g.disps, g.intrinsics, precision="f64")
    fx, fy, cx, cy = g.intrinsics
    v, u = np.meshgrid(np.arange(g.disps.shape[1]), np.arange(g.disps.shape[2]), indexing="ij")
    np.testing.assert_allclose(pts[0, ..., 0], (u - cx) / fx / g.disps[0], rtol=1e-6)  # pose 0 is the identity
    np.testing.assert_allclose(pts[0, ..., 2], 1.0 / g.disps[0], rtol=1e-6)
    c, val = oracle.projmap(g.poses, g.disps, g.intrinsics, np.array([3]), np.array([3]), precision="f64")
    np.testing.assert_allclose(c[0, ..., 0], u, atol=1e-9)
    np.testing.assert_allclose(c[0, ..., 1], v, atol=1e-9)
    assert (val == 1).all() and (c[..., 2] == 0).all()
    c2, _ = oracle.projmap(g.poses_gt, g.disps_gt, g.intrinsics, g.ii, g.jj, precision="f64")
    tg = synth.project(g.poses_gt, g.disps_gt, g.intrinsics, g.ii, g.jj)  # [E,2,ht,wd]
    np.testing.assert_allclose(c2[..., 0], tg[:, 0], rtol=1e-4, atol=1e-3)
    np.testing.assert_allclose(c2[..., 1], tg[:, 1], rtol=1e-4, atol=1e-3)
