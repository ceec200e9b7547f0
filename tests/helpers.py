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
