"""Generates tests/golden/droid_*.npz and geomfn_*.npz by RUNNING THE REFERENCE'S OWN CUDA EXTENSION.

Run on a B200 box (needs a GPU and oracle/_ref/ref_droid_backends.so, the reference's
src/lib/{droid.cpp,droid_kernels.cu,...} compiled unmodified for sm_100 by oracle/build_ref.py):

    python tests/golden/make_golden_cuda.py [--out DIR] [--only NAME ...]

Each droid_<case>.npz stores what the unmodified `droid_backends.ba` (src/lib/droid_kernels.cu:
1314-1437) leaves in `poses` / `disps` and returns as `[dx, dz]`, plus checksums of the seeded
synthetic inputs (tests/helpers.py::droid_case) so that generator drift is detected instead of
silently comparing different problems.  geomfn_*.npz hold the outputs of the reference's
frame_distance (:518-657), depth_filter (:661-775), iproj (:779-850) and projmap (:427-516).
Large cases store every `stride`-th disparity keyframe / output row (tests/helpers.py).
"""
import argparse
import importlib.util
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "glorie-slam_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import helpers  # noqa: E402


def load_ref():
    so = os.path.join(ROOT, "oracle", "_ref", "ref_droid_backends.so")
    spec = importlib.util.spec_from_file_location("ref_droid_backends", so)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    return ref


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.dirname(os.path.abspath(__file__)))
    ap.add_argument("--only", nargs="*", default=None)
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    ref = load_ref()
    want = lambda n: args.only is None or n in args.only

    for name in helpers.DROID_CASES:
        if not want(name):
            continue
        g, kw, stride = helpers.droid_case(name)
        poses, disps = dev(g.poses), dev(g.disps)
        dx, dz = ref.ba(poses, disps, dev(g.intrinsics), dev(g.disps_sens), dev(g.targets), dev(g.weights), dev(g.eta),
                        dev(g.ii), dev(g.jj), g.t0, g.t1, g.iters, g.lm, g.ep, bool(kw.get("motion_only", False)),
                        bool(kw.get("depth_only", False)))
        torch.cuda.synchronize()
        out = dict(checksum=helpers.checksum_full(g), poses=poses.cpu().numpy(), disps=disps.cpu().numpy()[::stride],
                   dx=dx.cpu().numpy(), stride=np.array(stride))
        if not kw.get("motion_only") and stride == 1:
            out["dz"] = dz.cpu().numpy()
        np.savez_compressed(os.path.join(args.out, "droid_%s.npz" % name), **out)
        print("wrote droid_%s.npz" % name, {k: v.shape for k, v in out.items()}, flush=True)

    if want("geomfn"):
        out = {}
        for name in ("A", "R", "C"):
            g, poses, ii, jj = helpers.frame_distance_case(name)
            tp, td, ti = dev(poses), dev(g.disps), dev(g.intrinsics)
            out["fd_%s_ij" % name] = ref.frame_distance(tp, td, ti, dev(ii), dev(jj), 0.3).cpu().numpy()
            out["fd_%s_ji" % name] = ref.frame_distance(tp, td, ti, dev(jj), dev(ii), 0.3).cpu().numpy()
            out["fd_%s_checksum" % name] = np.array([poses.astype(np.float64).sum(), g.disps.astype(np.float64).sum()])
        for c in helpers.DEPTH_FILTER_CASES:
            poses, disps, intr, inds, th = helpers.depth_filter_case(*c)
            r = ref.depth_filter(dev(poses), dev(disps), dev(intr), dev(inds), dev(th)).cpu().numpy()
            out["df_%d" % c[0]] = r.astype(np.int8)
            out["df_%d_checksum" % c[0]] = np.array([poses.astype(np.float64).sum(), disps.astype(np.float64).sum()])
        for name in ("A", "C"):
            g, poses, ks, es = helpers.reproj_case(name)
            tp, td, ti = dev(poses), dev(g.disps), dev(g.intrinsics)
            out["iproj_%s" % name] = ref.iproj(tp, td, ti).cpu().numpy()[::ks]
            ii, jj = g.ii[::es].copy(), g.jj[::es].copy()
            c, v = ref.projmap(tp, td, ti, dev(ii), dev(jj))
            out["projmap_%s_coords" % name] = c.cpu().numpy()
            out["projmap_%s_valid" % name] = v.cpu().numpy().astype(np.int8)
            out["reproj_%s_checksum" % name] = np.array([poses.astype(np.float64).sum(), g.disps.astype(np.float64).sum()])
        torch.cuda.synchronize()
        np.savez_compressed(os.path.join(args.out, "geomfn.npz"), **out)
        print("wrote geomfn.npz", {k: v.shape for k, v in out.items()}, flush=True)


if __name__ == "__main__":
    main()
