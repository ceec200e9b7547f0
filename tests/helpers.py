"""Shared helpers for the parity tests."""
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")

# The parity bar of BASELINE.json:north_star: state after a fixed iteration count within
# 1e-4 relative (fp32 arithmetic) of the reference.
STATE_RTOL = 1e-4


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def sign_normalise(poses):
    """Quaternion double cover: make qw >= 0 before comparing 7-vectors."""
    p = np.array(poses, dtype=np.float64, copy=True)
    s = np.where(p[:, 6:7] < 0, -1.0, 1.0)
    p[:, 3:] *= s
    return p


def checksum(g):
    return np.array([float(np.sum(a.astype(np.float64))) for a in
                     (g.poses, g.disps, g.targets, g.weights, g.eta, g.ii.astype(np.float64), g.jj.astype(np.float64))])


GOLDEN_CASES = {
    "A": dict(n=8, e=24, r=2, iters=4, lm=1e-4, ep=0.1, seed=0),
    "R": dict(n=6, e=14, r=2, iters=2, lm=1e-5, ep=1e-2, seed=3, ht=30, wd=40, t0=2),
}


def dspo_neg_graph():
    """DSPO input with negative disparities, also on a keyframe WITHOUT out-edges (ba.py:214 clamps the
    whole tensor); must stay identical to tests/golden/make_golden.py::dspo_neg_graph"""
    import synth

    g = synth.make_graph(7, 16, 2, iters=2, lm=1e-4, ep=0.1, seed=31, ht=16, wd=24)
    keep = g.ii != 4
    g.ii, g.jj, g.targets, g.weights = g.ii[keep], g.jj[keep], g.targets[keep], g.weights[keep]
    g.eta = g.eta[: len(np.unique(g.ii))]
    synth.with_mono_prior(g, seed=32)
    rng = np.random.default_rng(33)
    neg = rng.uniform(size=g.disps.shape) < 0.05
    g.disps[neg] = -g.disps[neg] * 0.3
    return g


def load_golden(name):
    import synth

    if name in ("near", "step6"):
        g, kw = synth.make_edge_case(name), {}
    elif name == "dspo_neg":
        g, kw = dspo_neg_graph(), {}
    else:
        kw = GOLDEN_CASES[name]
        g = synth.make_graph(**kw)
    z = np.load(os.path.join(GOLDEN, "geom_%s.npz" % name))
    np.testing.assert_allclose(checksum(g), z["checksum"], rtol=1e-9, err_msg="synthetic generator drifted")
    return g, z, kw


def small_graphs():
    import synth

    return {
        "A": synth.make_config("A"),
        "R": synth.make_graph(6, 14, 2, iters=2, lm=1e-5, ep=1e-2, seed=3, ht=30, wd=40, t0=2),
        "tiny": synth.make_graph(5, 8, 1, iters=2, seed=2, ht=8, wd=12),
        "B": synth.make_config("B"),
    }


# ---- fixtures produced by the reference's own CUDA extension (tests/golden/make_golden_cuda.py) ----
# name -> (graph builder, keyword arguments of the ba call, stride of the stored disparity keyframes)
def droid_case(name):
    import synth

    if name in ("A", "R", "tiny", "B"):
        return small_graphs()[name], {}, 1
    if name == "C":
        return synth.make_config("C"), {}, 4
    if name == "E":
        return synth.make_config("E"), {}, 20
    if name in synth.EDGE_CASES:
        return synth.make_edge_case(name), {}, 1
    if name == "R_motion":
        return small_graphs()["R"], dict(motion_only=True), 1
    if name == "R_depth":
        return small_graphs()["R"], dict(depth_only=True), 1
    if name == "tiny_sens":
        g = small_graphs()["tiny"]
        rng = np.random.default_rng(4)
        g.disps_sens = (g.disps_gt * (rng.uniform(size=g.disps.shape) < 0.5)).astype(np.float32)
        return g, {}, 1
    raise KeyError(name)


DROID_CASES = ("A", "R", "tiny", "B", "C", "E", "near", "stereo", "step4", "R_motion", "R_depth", "tiny_sens")


def checksum_full(g):
    return np.array([float(np.sum(np.asarray(a).astype(np.float64))) for a in
                     (g.poses, g.disps, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj, g.intrinsics)])


def load_droid_golden(name):
    """(graph, call kwargs, stride, fixture) of one reference-CUDA fixture; checks that the seeded
    generator still produces the inputs the fixture was computed from."""
    g, kw, stride = droid_case(name)
    z = np.load(os.path.join(GOLDEN, "droid_%s.npz" % name))
    np.testing.assert_allclose(checksum_full(g), z["checksum"], rtol=1e-9, err_msg="synthetic generator drifted")
    return g, kw, stride, z


def frame_distance_case(name):
    """inputs of the frame_distance fixtures: every ordered keyframe pair, one keyframe looking the other way"""
    import synth

    g = small_graphs()[name] if name != "C" else synth.make_config("C")
    n = g.n_frames
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    poses = g.poses.copy()
    poses[n // 2, 3:] = [0.0, 1.0, 0.0, 0.0]
    return g, poses, ii.reshape(-1).astype(np.int64), jj.reshape(-1).astype(np.int64)


DEPTH_FILTER_CASES = [(12, 24, 32, 0.0), (40, 48, 64, 0.01), (9, 21, 30, 0.02)]


def depth_filter_case(n, ht, wd, noise):
    poses, disps, intr = planar_scene(n, ht, wd, seed=n, noise=noise)
    thresh = (0.02 * (1.0 / disps).mean(axis=(1, 2))).astype(np.float32)
    inds = np.concatenate([np.arange(n), [3, 0]]).astype(np.int64)  # repeated / unordered entries are legal
    th = np.concatenate([thresh, thresh[[3, 0]] * 0.1]).astype(np.float32)
    return poses, disps, intr, inds, th


def reproj_case(name):
    """inputs of the iproj / projmap fixtures; returns graph, poses (keyframe 1 looks the other way),
    keyframe stride of the stored iproj rows, edge stride of the stored projmap rows"""
    import synth

    g = small_graphs()["A"] if name == "A" else synth.make_config("C")
    poses = g.poses.copy()
    poses[1, 3:] = [0.0, 1.0, 0.0, 0.0]
    return g, poses, (1 if name == "A" else 8), (1 if name == "A" else 100)


def planar_scene(n_frames, ht, wd, seed=0, noise=0.0):
    """Keyframes moving sideways in front of a tilted plane, with the exact disparity each one sees:
    multi-view-consistent depth for the depth_filter tests.  Returns poses [n,7] (world->camera,
    tx ty tz qx qy qz qw), disps [n,ht,wd], intrinsics [4] as float32."""
    rng = np.random.default_rng(seed)
    intr = np.array([0.9 * wd, 0.9 * wd, wd / 2.0 - 0.5, ht / 2.0 - 0.5], dtype=np.float64)
    nrm, c = np.array([0.15, -0.1, 1.0]), 4.0  # plane nrm . X = c in world coordinates
    poses = np.zeros((n_frames, 7))
    disps = np.zeros((n_frames, ht, wd))
    v, u = np.meshgrid(np.arange(ht), np.arange(wd), indexing="ij")
    rays = np.stack([(u - intr[2]) / intr[0], (v - intr[3]) / intr[1], np.ones_like(u, dtype=np.float64)], -1)
    for k in range(n_frames):
        w = rng.normal(0, 0.02, 3)
        th = np.linalg.norm(w)
        q = np.concatenate([np.sin(th / 2) * w / th, [np.cos(th / 2)]])
        x, y, z, s = q
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - s * z), 2 * (x * z + s * y)],
                      [2 * (x * y + s * z), 1 - 2 * (x * x + z * z), 2 * (y * z - s * x)],
                      [2 * (x * z - s * y), 2 * (y * z + s * x), 1 - 2 * (x * x + y * y)]])
        t = np.array([-0.12 * k, 0.0, 0.0]) + rng.normal(0, 0.02, 3)
        poses[k, :3], poses[k, 3:] = t, q
        # camera point p = depth * ray,  world X = R^T (p - t),  nrm . X = c
        depth = (c + nrm @ (R.T @ t)) / (rays @ (R @ nrm))
        disps[k] = 1.0 / depth
    if noise > 0:
        disps *= 1.0 + rng.normal(0, noise, disps.shape)
    return poses.astype(np.float32), disps.astype(np.float32), intr.astype(np.float32)


def dispatcher_case(seed=0, n=24, e=140, r=4, ht=24, wd=32, buf=32):
    """Inputs for the call-site tests: a multi-view-consistent scene (planar_scene), so that depth_filter lets
    most pixels through, a mono prior with one keyframe that has none, in a video buffer longer than the
    keyframes in use.  Returns (graph, poses [buf,7], disps [buf,ht,wd], mono_disps [buf,ht,wd])."""
    import synth

    rng = np.random.default_rng(600 + seed)
    pgt, dgt, intr = planar_scene(n, ht, wd, seed=70 + seed)
    ii, jj = synth.make_edges(n, e, r, rng)
    tg = synth.project(pgt, dgt, intr, ii, jj) + rng.normal(0.0, 0.3, size=(len(ii), 2, ht, wd))
    poses = pgt.astype(np.float64).copy()
    poses[1:] = synth.se3_retr(rng.normal(0.0, 0.003, size=(n - 1, 6)), poses[1:])
    disps = dgt * rng.uniform(0.98, 1.02, size=dgt.shape)
    kx = np.unique(np.concatenate([np.arange(1, n), ii]))
    eta = (0.2 * rng.uniform(0.0, 0.01, size=(len(kx), ht, wd)) + 1e-7).astype(np.float32)
    g = synth.Graph(poses=poses.astype(np.float32), disps=disps.astype(np.float32), intrinsics=intr,
                    disps_sens=np.zeros_like(dgt), targets=tg.astype(np.float32),
                    weights=rng.uniform(0.0, 1.0, size=tg.shape).astype(np.float32), eta=eta, ii=ii, jj=jj, t0=1, t1=n,
                    iters=2, lm=1e-4, ep=0.1, poses_gt=pgt, disps_gt=dgt)
    synth.with_mono_prior(g, seed=601 + seed)
    # a clean affine prior (no dropped pixels: the scale fit of depth_video.py:233 does not mask them and would be
    # ill-conditioned on this nearly fronto-parallel scene) ...
    sc, sh = rng.uniform(0.7, 1.3, size=(n, 1, 1)), rng.normal(0.0, 0.02, size=(n, 1, 1))
    g.mono_disps = ((dgt.astype(np.float64) - sh) / sc).astype(np.float32)
    g.mono_disps[5] = 0.0  # ... except keyframe 5, which has none: the mono_thres filter drops its edges
    pad = lambda a: np.concatenate([a, np.zeros((buf - a.shape[0],) + a.shape[1:], a.dtype)], 0)
    pbuf = pad(g.poses)
    pbuf[n:, 6] = 1.0
    return g, pbuf, pad(g.disps), pad(g.mono_disps)


# ---- proximity edge selection (reference src/factor_graph.py:386-462) -------------------------------------
# name: (keyframes, (ht, wd), seed, t_start, t_start_loop, nms, radius, thresh, max_factors, loop)
PROX_CASES = {
    "global": (60, (16, 24), 1, 0, None, 2, 2, 12.0, 400, False),
    "capped": (60, (16, 24), 2, 0, None, 1, 3, 30.0, 520, False),
    "window": (48, (16, 24), 3, 4, 20, 2, 2, 14.0, 10 ** 6, False),
    "loop": (64, (16, 24), 4, 0, 30, 2, 2, 14.0, 10 ** 6, True),
}


def prox_distances(n, hw, seed, t_start, t_loop, loop=False):
    """a trajectory that revisits places (a slow circle) so that far-apart keyframes are close again; returns
    poses, disps, intrinsics and the bidirectional frame distance of every (i in [t_loop, n), j in [t_start, n))
    pair as the reference's DepthVideo.distance computes it (float32)"""
    from oracle import oracle

    ht, wd = hw
    rng = np.random.default_rng(900 + seed)
    intr = np.array([0.9 * wd, 0.9 * wd, wd / 2.0 - 0.5, ht / 2.0 - 0.5], dtype=np.float32)
    poses = np.zeros((n, 7), dtype=np.float64)
    ang = 2 * np.pi * np.arange(n) / (0.45 * n)  # a bit more than two laps
    poses[:, 0] = 0.6 * np.cos(ang) + rng.normal(0, 0.01, n)
    poses[:, 1] = 0.6 * np.sin(ang) + rng.normal(0, 0.01, n)
    poses[:, 2] = rng.normal(0, 0.01, n)
    poses[:, 6] = 1.0
    disps = rng.uniform(0.2, 0.3, size=(n, ht, wd)).astype(np.float32)
    t_loop = t_start if (t_loop is None or not loop) else t_loop  # factor_graph.py:387-388
    ii, jj = np.meshgrid(np.arange(t_loop, n), np.arange(t_start, n), indexing="ij")
    ii, jj = ii.reshape(-1), jj.reshape(-1)
    p32 = poses.astype(np.float32)
    d1 = oracle.frame_distance(p32, disps, intr, ii, jj, 0.25, precision="f32")
    d2 = oracle.frame_distance(p32, disps, intr, jj, ii, 0.25, precision="f32")
    return p32, disps, intr, (np.float32(0.5) * (d1 + d2)).astype(np.float32)


def mask_case(seed=0, n=14, ht=24, wd=32):
    """inputs of the valid-depth mask / mono_thres filter fixtures: disparities with far outliers, vote counts
    0..6, one keyframe without any surviving pixel, one with an even and one with an odd number of survivors"""
    rng = np.random.default_rng(700 + seed)
    disps = rng.uniform(0.2, 1.5, size=(n, ht, wd)).astype(np.float32)
    disps[rng.uniform(size=disps.shape) < 0.05] *= 0.02      # very far points: depth > 3 median
    count = rng.integers(0, 7, size=(n, ht, wd)).astype(np.float32)
    count[3] = 0.0                                            # nothing survives: nanmedian of all-NaN
    count[4].reshape(-1)[:] = 0.0
    count[4].reshape(-1)[:10] = 6.0                           # even number of survivors
    count[5].reshape(-1)[:] = 0.0
    count[5].reshape(-1)[:11] = 6.0                           # odd number
    count[6][: ht // 2 + 4] = 0.0                             # fewer than half of the pixels valid
    error = rng.uniform(0.0, 0.2, size=n).astype(np.float32)
    error[7] = np.nan
    scale = rng.uniform(0.5, 1.5, size=n).astype(np.float32)
    scale[8] = -0.3
    e = 60
    ii = rng.integers(0, n + 2, size=e).astype(np.int64)      # some edges touch keyframes beyond the first n
    jj = rng.integers(0, n + 2, size=e).astype(np.int64)
    return disps, count, error, scale, ii, jj
