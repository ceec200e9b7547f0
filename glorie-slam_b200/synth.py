"""Seeded synthetic factor graphs for the dense-BA / DSPO hot path.

Host-side helper (numpy only) shared by the parity tests, the golden-fixture
generator and bench.py.  The recipe is SURVEY.md section 8(d): a smooth
camera trajectory, disparities around 1 (as after `DepthVideo.normalize`,
reference src/depth_video.py:146-153), temporal edges (i, i+-d) for d = 1..r
in both directions plus seeded long-range pairs, targets = reprojection at
ground truth + pixel noise, then perturbed initial poses / disparities so
the solve has something to do.  Weights follow the sigmoid range of the
DROID update head (reference src/modules/droid_net/droid_net.py:93), eta the
0.2 * eta + 1e-7 damping of reference src/factor_graph.py:248.

Pose storage is the reference's: [tx, ty, tz, qx, qy, qz, qw], world->camera.
"""
from __future__ import annotations

import dataclasses
from typing import Optional

import numpy as np

# name -> (keyframes, edges, temporal radius, iterations, lm, ep)
CONFIGS = {
    "A": dict(n=8, e=24, r=2, iters=4, lm=1e-4, ep=0.1),       # CPU-runnable reference case
    "B": dict(n=20, e=250, r=7, iters=2, lm=1e-4, ep=0.1),     # frontend local window
    "C": dict(n=200, e=3000, r=7, iters=2, lm=1e-5, ep=1e-2),  # backend global BA
    "D": dict(n=200, e=3000, r=7, iters=2, lm=1e-5, ep=1e-2),  # C + mono prior (DSPO)
    "E": dict(n=1000, e=20000, r=10, iters=2, lm=1e-5, ep=1e-2),  # ScanNet-scale global BA
}

HT, WD = 48, 64
INTRINSICS = (65.0, 65.0, 32.0, 24.0)


# ----------------------------------------------------------------------------
# minimal SE3 helpers (float64, quaternion xyzw) -- generator only
# ----------------------------------------------------------------------------
def quat_mul(a, b):
    ax, ay, az, aw = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    bx, by, bz, bw = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack(
        [
            aw * bx + ax * bw + ay * bz - az * by,
            aw * by - ax * bz + ay * bw + az * bx,
            aw * bz + ax * by - ay * bx + az * bw,
            aw * bw - ax * bx - ay * by - az * bz,
        ],
        axis=-1,
    )


def quat_rot(q, x):
    qv = q[..., :3]
    uv = 2.0 * np.cross(qv, x)
    return x + q[..., 3:4] * uv + np.cross(qv, uv)


def se3_exp(xi):
    """xi = (tau, phi) -> (t, q)."""
    xi = np.asarray(xi, dtype=np.float64)
    tau, phi = xi[..., :3], xi[..., 3:]
    th2 = np.sum(phi * phi, axis=-1, keepdims=True)
    th = np.sqrt(th2)
    small = th < 1e-6
    ths = np.where(small, 1.0, th)
    imag = np.where(small, 0.5 - th2 / 48.0, np.sin(0.5 * ths) / ths)
    real = np.where(small, 1.0 - th2 / 8.0, np.cos(0.5 * ths))
    q = np.concatenate([imag * phi, real], axis=-1)
    a = np.where(small, 0.5 - th2 / 24.0, (1.0 - np.cos(ths)) / (ths * ths))
    b = np.where(small, 1.0 / 6.0 - th2 / 120.0, (ths - np.sin(ths)) / (ths * ths * ths))
    c1 = np.cross(phi, tau)
    c2 = np.cross(phi, c1)
    t = tau + a * c1 + b * c2
    return t, q


def se3_mul(ta, qa, tb, qb):
    return quat_rot(qa, tb) + ta, quat_mul(qa, qb)


def se3_inv(t, q):
    qi = q * np.array([-1.0, -1.0, -1.0, 1.0])
    return -quat_rot(qi, t), qi


def se3_retr(xi, pose):
    """Exp(xi) * pose for [N,7] poses."""
    dt, dq = se3_exp(xi)
    t, q = se3_mul(dt, dq, pose[..., :3], pose[..., 3:])
    return np.concatenate([t, q], axis=-1)


# ----------------------------------------------------------------------------
@dataclasses.dataclass
class Graph:
    """One synthetic BA problem, arrays laid out as `droid_backends.ba` takes them."""

    poses: np.ndarray        # [N,7] f32  (perturbed start)
    disps: np.ndarray        # [N,ht,wd] f32 (perturbed start)
    intrinsics: np.ndarray   # [4] f32
    disps_sens: np.ndarray   # [N,ht,wd] f32 (zeros)
    targets: np.ndarray      # [E,2,ht,wd] f32 (planar u,v)
    weights: np.ndarray      # [E,2,ht,wd] f32
    eta: np.ndarray          # [K,ht,wd] f32
    ii: np.ndarray           # [E] i64
    jj: np.ndarray           # [E] i64
    t0: int
    t1: int
    iters: int
    lm: float
    ep: float
    poses_gt: np.ndarray
    disps_gt: np.ndarray
    # DSPO extras (filled by with_mono_prior)
    mono_disps: Optional[np.ndarray] = None
    scales: Optional[np.ndarray] = None
    shifts: Optional[np.ndarray] = None
    valid_depth_mask: Optional[np.ndarray] = None
    alpha: float = 0.01

    @property
    def n_frames(self):
        return self.poses.shape[0]

    @property
    def n_edges(self):
        return self.ii.shape[0]

    @property
    def hw(self):
        return self.disps.shape[1] * self.disps.shape[2]


def make_edges(n, e, r, rng):
    """Temporal edges by increasing |i-j| (both directions) up to the quota, then
    seeded random long-range pairs; no self edges, no duplicates."""
    pairs = []
    seen = set()

    def add(i, j):
        if i == j or (i, j) in seen or len(pairs) >= e:
            return
        seen.add((i, j))
        pairs.append((i, j))

    for d in range(1, r + 1):
        for i in range(n):
            if i + d < n:
                add(i, i + d)
                add(i + d, i)
    guard = 0
    while len(pairs) < e and guard < 100 * e + 1000:
        guard += 1
        i = int(rng.integers(0, n))
        j = int(rng.integers(0, n))
        if abs(i - j) <= r:
            continue
        add(i, j)
        add(j, i)
    # fall back to any remaining free pair (tiny graphs)
    if len(pairs) < e:
        for d in range(r + 1, n):
            for i in range(n - d):
                add(i, i + d)
                add(i + d, i)
    arr = np.asarray(pairs, dtype=np.int64).reshape(-1, 2)
    return arr[:, 0].copy(), arr[:, 1].copy()


def project(poses, disps, intr, ii, jj, chunk=512):
    """pi(G_ij X_i) for every edge pixel, float64: returns [E,2,ht,wd]."""
    fx, fy, cx, cy = [float(v) for v in intr]
    n, ht, wd = disps.shape
    v, u = np.meshgrid(np.arange(ht, dtype=np.float64), np.arange(wd, dtype=np.float64), indexing="ij")
    x0 = ((u - cx) / fx).reshape(-1)
    y0 = ((v - cy) / fy).reshape(-1)
    out = np.empty((ii.shape[0], 2, ht * wd), dtype=np.float32)
    t = poses[:, :3].astype(np.float64)
    q = poses[:, 3:].astype(np.float64)
    for s in range(0, ii.shape[0], chunk):
        i = ii[s : s + chunk]
        j = jj[s : s + chunk]
        tinv, qinv = se3_inv(t[i], q[i])
        tij, qij = se3_mul(t[j], q[j], tinv, qinv)
        z = disps[i].reshape(len(i), -1).astype(np.float64)
        pts = np.stack([np.broadcast_to(x0, z.shape), np.broadcast_to(y0, z.shape), np.ones_like(z)], axis=-1)
        xj = quat_rot(qij[:, None, :], pts) + tij[:, None, :] * z[..., None]
        d = 1.0 / xj[..., 2]
        out[s : s + chunk, 0] = fx * xj[..., 0] * d + cx
        out[s : s + chunk, 1] = fy * xj[..., 1] * d + cy
    return out.reshape(ii.shape[0], 2, ht, wd)


def make_graph(n, e, r=2, iters=2, lm=1e-4, ep=0.1, ht=HT, wd=WD, seed=0,
               pose_noise=0.01, disp_noise=0.1, pixel_noise=0.5, t0=1, t1=None) -> Graph:
    rng = np.random.default_rng(seed)
    intr = np.asarray(INTRINSICS, dtype=np.float32)
    if (ht, wd) != (HT, WD):
        intr = np.asarray([1.02 * wd, 1.02 * wd, wd / 2.0, ht / 2.0], dtype=np.float32)

    # smooth trajectory T_k = Exp(xi_k) T_{k-1}, T_0 = I
    poses_gt = np.zeros((n, 7), dtype=np.float64)
    poses_gt[:, 6] = 1.0
    xis = rng.normal(0.0, 0.02, size=(n, 6))
    for k in range(1, n):
        poses_gt[k] = se3_retr(xis[k], poses_gt[k - 1])

    disps_gt = rng.uniform(0.5, 1.5, size=(n, ht, wd))
    ii, jj = make_edges(n, e, r, rng)

    targets = project(poses_gt, disps_gt, intr, ii, jj)
    targets = targets + rng.normal(0.0, pixel_noise, size=targets.shape).astype(np.float32)
    weights = rng.uniform(0.0, 1.0, size=targets.shape).astype(np.float32)

    poses = poses_gt.copy()
    poses[1:] = se3_retr(rng.normal(0.0, pose_noise, size=(n - 1, 6)), poses_gt[1:])
    disps = disps_gt * rng.uniform(1.0 - disp_noise, 1.0 + disp_noise, size=disps_gt.shape)

    t1 = n if t1 is None else t1
    kx = np.unique(np.concatenate([np.arange(t0, t1), ii]))
    eta = (0.2 * rng.uniform(0.0, 0.01, size=(kx.shape[0], ht, wd)) + 1e-7).astype(np.float32)

    return Graph(
        poses=poses.astype(np.float32), disps=disps.astype(np.float32), intrinsics=intr,
        disps_sens=np.zeros((n, ht, wd), dtype=np.float32), targets=targets.astype(np.float32),
        weights=weights, eta=eta, ii=ii, jj=jj, t0=t0, t1=t1, iters=iters, lm=lm, ep=ep,
        poses_gt=poses_gt.astype(np.float32), disps_gt=disps_gt.astype(np.float32),
    )


def with_mono_prior(g: Graph, seed=1, alpha=0.01) -> Graph:
    """Config D: add the synthetic mono-depth prior inputs of BA_with_scale_shift."""
    rng = np.random.default_rng(seed)
    n = g.n_frames
    s = rng.uniform(0.7, 1.3, size=(n, 1, 1))
    q = rng.normal(0.0, 0.05, size=(n, 1, 1))
    mono = (g.disps_gt.astype(np.float64) - q) / s
    drop = rng.uniform(size=mono.shape) < 0.05
    mono[drop] = 0.0
    g.mono_disps = mono.astype(np.float32)
    g.valid_depth_mask = rng.uniform(size=mono.shape) < 0.7
    g.scales = np.ones((n,), dtype=np.float32)
    g.shifts = np.zeros((n,), dtype=np.float32)
    g.alpha = alpha
    return g


def make_config(name: str, seed=0, **over) -> Graph:
    c = dict(CONFIGS[name])
    c.update(over)
    g = make_graph(c["n"], c["e"], c["r"], c["iters"], c["lm"], c["ep"], seed=seed,
                   **{k: v for k, v in c.items() if k in ("ht", "wd", "t0", "t1")})
    if name == "D":
        with_mono_prior(g, seed=seed + 1)
    return g


# ----------------------------------------------------------------------------
# Edge cases the reference has code for
# ----------------------------------------------------------------------------
EDGE_CASES = ("near", "stereo", "step4", "step6")


def reprojected_depth(g: Graph):
    """Z_j of every edge pixel at the graph's START state, float64 [E,ht,wd]: which side of the
    MIN_DEPTH tests (droid_kernels.cu:302-306, projective_ops.py:44,112) a pixel falls on."""
    n, ht, wd = g.disps.shape
    fx, fy, cx, cy = [float(v) for v in g.intrinsics]
    v, u = np.meshgrid(np.arange(ht, dtype=np.float64), np.arange(wd, dtype=np.float64), indexing="ij")
    x0, y0 = ((u - cx) / fx).reshape(-1), ((v - cy) / fy).reshape(-1)
    t, q = g.poses[:, :3].astype(np.float64), g.poses[:, 3:].astype(np.float64)
    tinv, qinv = se3_inv(t[g.ii], q[g.ii])
    tij, qij = se3_mul(t[g.jj], q[g.jj], tinv, qinv)
    z = g.disps[g.ii].reshape(len(g.ii), -1).astype(np.float64)
    pts = np.stack([np.broadcast_to(x0, z.shape), np.broadcast_to(y0, z.shape), np.ones_like(z)], axis=-1)
    xj = quat_rot(qij[:, None, :], pts) + tij[:, None, :] * z[..., None]
    return xj[..., 2].reshape(len(g.ii), ht, wd)


def make_edge_case(name: str) -> Graph:
    """Small graphs (16 x 24 pixels) that drive the branches ordinary synthetic graphs never reach.

    near    4 % of the pixels are very close points (disparity 4..40): their reprojected depth Z_j
            lands below 0.1, in 0.1..0.2, in 0.2..0.25 and above, and behind the camera
    stereo  three stereo self edges ii == jj among ordinary ones (droid_kernels.cu:219-229,323,356)
    step4   nearly converged: the rotation update |phi| straddles 1e-4, i.e. theta^2 < 1e-8 and
            theta > 1e-4 both occur in the hand-rolled exponential (droid_kernels.cu:119,162)
    step6   converged further: |phi| straddles lietorch's EPS = 1e-6 (so3.h:153-170)
    """
    ht, wd = 16, 24
    if name == "near":
        g = make_graph(8, 24, 2, iters=2, seed=21, ht=ht, wd=wd)
        rng = np.random.default_rng(121)
        # the camera advances 0.07 per keyframe along its optical axis, so that close points
        # (a quarter of the pixels, disparity 2..16) come within MIN_DEPTH of the neighbouring cameras
        pgt = g.poses_gt.astype(np.float64)
        pgt[:, 2] -= 0.07 * np.arange(g.n_frames)
        mask = rng.uniform(size=g.disps.shape) < 0.25
        big = rng.uniform(2.0, 16.0, size=g.disps.shape)
        dgt = g.disps_gt.astype(np.float64)
        dgt[mask] = big[mask]
        tg = np.clip(project(pgt, dgt, g.intrinsics, g.ii, g.jj).astype(np.float64), -1e3, 1e3)
        g.targets = (tg + rng.normal(0.0, 0.5, size=tg.shape)).astype(np.float32)
        g.poses_gt, g.disps_gt = pgt.astype(np.float32), dgt.astype(np.float32)
        poses = pgt.copy()
        poses[1:] = se3_retr(rng.normal(0.0, 0.01, size=(g.n_frames - 1, 6)), pgt[1:])
        g.poses = poses.astype(np.float32)
        g.disps = (dgt * rng.uniform(0.9, 1.1, size=dgt.shape)).astype(np.float32)
        return g
    if name == "stereo":
        g = make_graph(8, 24, 2, iters=2, seed=22, ht=ht, wd=wd)
        rng = np.random.default_rng(122)
        selfs = np.array([2, 5, 6], dtype=np.int64)
        fx, cx = float(g.intrinsics[0]), float(g.intrinsics[2])
        v, u = np.meshgrid(np.arange(ht, dtype=np.float64), np.arange(wd, dtype=np.float64), indexing="ij")
        tg = np.zeros((len(selfs), 2, ht, wd))
        for a, k in enumerate(selfs):  # G_ij = (t = (-0.1, 0, 0), q = identity): u shifts by -0.1 fx z
            tg[a, 0] = u - 0.1 * fx * g.disps_gt[k]
            tg[a, 1] = v
        tg += rng.normal(0.0, 0.5, size=tg.shape)
        where = [3, 11, 20]  # positions inside the edge list
        ii, jj, T, W = list(g.ii), list(g.jj), list(g.targets), list(g.weights)
        for a, pos in enumerate(where):
            ii.insert(pos, int(selfs[a])); jj.insert(pos, int(selfs[a]))
            T.insert(pos, tg[a].astype(np.float32))
            W.insert(pos, rng.uniform(0.0, 1.0, size=tg[a].shape).astype(np.float32))
        g.ii, g.jj = np.asarray(ii, dtype=np.int64), np.asarray(jj, dtype=np.int64)
        g.targets, g.weights = np.stack(T).astype(np.float32), np.stack(W).astype(np.float32)
        return g
    if name in ("step4", "step6"):
        s = 1e-4 if name == "step4" else 1e-6
        return make_graph(8, 24, 2, iters=2, seed=23 if name == "step4" else 24, ht=ht, wd=wd, pose_noise=s,
                          disp_noise=s, pixel_noise=0.0)
    raise KeyError(name)
