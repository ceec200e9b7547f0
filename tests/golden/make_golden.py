"""Generates tests/golden/geom_*.npz by RUNNING THE REFERENCE ITSELF on CPU.

Run in the build container only (needs /root/reference and oracle/_ref/lietorch_backends.so,
see oracle/build_ref.py):   python tests/golden/make_golden.py

Each fixture stores the outputs of the unmodified reference functions
  src/geom/ba.py::BA                  (geom_py semantics, 4 or 2 iterations)
  src/geom/ba.py::BA_with_scale_shift (DSPO stage 2, 2 iterations)
in float32 and float64, plus checksums of the seeded synthetic inputs
(glorie-slam_b200/synth.py) so that generator drift is detected instead of
silently comparing different problems.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "glorie-slam_b200"))
import synth  # noqa: E402
from oracle import ref_harness  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def checksum(g):
    return np.array([float(np.sum(a.astype(np.float64))) for a in
                     (g.poses, g.disps, g.targets, g.weights, g.eta, g.ii.astype(np.float64), g.jj.astype(np.float64))])


CASES = {
    # name: (make_graph kwargs)
    "A": dict(n=8, e=24, r=2, iters=4, lm=1e-4, ep=0.1, seed=0),
    # ragged: ScanNet-size grid (30x40), two fixed poses, backend damping
    "R": dict(n=6, e=14, r=2, iters=2, lm=1e-5, ep=1e-2, seed=3, ht=30, wd=40, t0=2),
}


def dspo_neg_graph():
    """DSPO input with negative disparities, also on a keyframe WITHOUT out-edges (not in unique(ii)):
    ba.py:214 clamps the whole tensor, not only the rows it solved for."""
    g = synth.make_graph(7, 16, 2, iters=2, lm=1e-4, ep=0.1, seed=31, ht=16, wd=24)
    keep = g.ii != 4  # keyframe 4 keeps only in-edges
    g.ii, g.jj, g.targets, g.weights = g.ii[keep], g.jj[keep], g.targets[keep], g.weights[keep]
    g.eta = g.eta[: len(np.unique(g.ii))]
    synth.with_mono_prior(g, seed=32)
    rng = np.random.default_rng(33)
    neg = rng.uniform(size=g.disps.shape) < 0.05
    g.disps[neg] = -g.disps[neg] * 0.3
    return g


EXTRA = {
    # name: builder -- branches ordinary graphs never reach (synth.make_edge_case)
    "near": lambda: synth.make_edge_case("near"),
    "step6": lambda: synth.make_edge_case("step6"),
}


def main():
    import sys as _sys

    only = _sys.argv[1:]
    for name, build in EXTRA.items():
        if only and name not in only:
            continue
        g = build()
        out = dict(checksum=checksum(g))
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            p, d = ref_harness.run_BA(g, dtype=dt)
            out["poses_" + tag] = p
            out["disps_" + tag] = d
        np.savez_compressed(os.path.join(HERE, "geom_%s.npz" % name), **out)
        print("wrote geom_%s.npz" % name, {k: v.shape for k, v in out.items()})
    if not only or "dspo_neg" in only:
        g = dspo_neg_graph()
        out = dict(checksum=checksum(g))
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            d, wq = ref_harness.run_DSPO(g, iters=2, dtype=dt)
            out["dspo_disps_" + tag] = d
            out["dspo_wqs_" + tag] = wq
        np.savez_compressed(os.path.join(HERE, "geom_dspo_neg.npz"), **out)
        print("wrote geom_dspo_neg.npz", {k: v.shape for k, v in out.items()})
    for name, kw in CASES.items():
        if only and name not in only:
            continue
        g = synth.make_graph(**kw)
        out = dict(checksum=checksum(g))
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            p, d = ref_harness.run_BA(g, dtype=dt)
            out["poses_" + tag] = p
            out["disps_" + tag] = d
        synth.with_mono_prior(g, seed=kw["seed"] + 1)
        for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
            d, wq = ref_harness.run_DSPO(g, iters=2, dtype=dt)
            out["dspo_disps_" + tag] = d
            out["dspo_wqs_" + tag] = wq
        np.savez_compressed(os.path.join(HERE, "geom_%s.npz" % name), **out)
        print("wrote geom_%s.npz" % name, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
