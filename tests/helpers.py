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


def load_golden(name):
    import synth

    kw = GOLDEN_CASES[name]
    g = synth.make_graph(**kw)
    z = np.load(os.path.join(GOLDEN, "geom_%s.npz" % name))
    np.testing.assert_allclose(checksum(g), z["checksum"], rtol=1e-9, err_msg="synthetic generator drifted")
    return g, z, kw


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
