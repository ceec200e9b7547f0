"""Generates tests/golden/prox.npz by RUNNING THE REFERENCE'S OWN edge selection on CPU.

Run in the build container only (needs /root/reference):   python tests/golden/make_golden_prox.py

The function under test is `FactorGraph.add_backend_proximity_factors` (reference src/factor_graph.py:386-462),
called unbound on a stand-in `self` that supplies the three things it touches: `video.distance` (returns the
stored distance vector), `device`, and `add_factors` (records the edge list).  src/factor_graph.py imports the
DROID network module for names this function never uses; an empty stand-in module satisfies the import.
Also runs the valid-depth mask (src/depth_video.py:349-361) and the mono_thres edge filter (:245-261) as the
reference states them, on CPU tensors.
Stored per case: the inputs (distance vector d as float32, parameters) and the reference's edge list.
"""
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "glorie-slam_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import helpers  # noqa: E402
import synth  # noqa: E402
from oracle import oracle, ref_harness  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def load_factor_graph():
    ref_harness.load()  # src, src.geom, src.geom.projective_ops
    mods = types.ModuleType("src.modules")
    mods.__path__ = []
    dn = types.ModuleType("src.modules.droid_net")
    dn.CorrBlock = dn.AltCorrBlock = object
    sys.modules["src.modules"] = mods
    sys.modules["src.modules.droid_net"] = dn
    path = os.path.join(REF, "src", "factor_graph.py")
    fg = types.ModuleType("src.factor_graph")
    fg.__file__ = path
    exec(compile(open(path).read(), path, "exec"), fg.__dict__)
    return fg.FactorGraph


def main():
    FG = load_factor_graph()
    out = {}
    for name, (n, hw, seed, t_start, t_loop, nms, radius, thresh, maxf, loop) in helpers.PROX_CASES.items():
        poses, disps, intr, d = helpers.prox_distances(n, hw, seed, t_start, t_loop, loop)
        rec = {}

        class Self:
            device = "cpu"
            video = types.SimpleNamespace(distance=lambda ii, jj, beta: torch.from_numpy(d.copy()))
            ii = []

            def add_factors(self, ii, jj, remove=False):
                rec["ii"], rec["jj"] = ii.numpy().copy(), jj.numpy().copy()

        ret = FG.add_backend_proximity_factors(Self(), t_start, n, nms, radius, thresh, maxf, 0.25,
                                               t_start_loop=t_loop, loop=loop)
        out[name + "_d"] = d
        out[name + "_ii"] = rec.get("ii", np.zeros(0, dtype=np.int64))
        out[name + "_jj"] = rec.get("jj", np.zeros(0, dtype=np.int64))
        out[name + "_ret0"] = np.array(0 if ret == 0 else 1)
        print(name, "edges", len(out[name + "_ii"]), "finite candidates", int(np.isfinite(d).sum()), "ret", ret)
    # valid-depth mask (src/depth_video.py:349-361) and mono_thres filter (:245-261): the reference's statements,
    # executed with torch on CPU (src/depth_video.py itself cannot be imported without the whole SLAM stack)
    disps, count, error, scale, ii, jj = (torch.from_numpy(a) for a in helpers.mask_case())
    visible_num, mono_thres = 2, 0.1
    depths = 1.0 / disps
    multiview_masks = (count >= visible_num)
    depths[~multiview_masks] = torch.nan
    depths_median = depths.view(depths.shape[0], -1).nanmedian(dim=1).values
    masks = depths < 3 * depths_median[:, None, None]
    out["mask"] = masks.numpy()
    out["mask_median"] = depths_median.numpy()
    valid_d, est_d = masks, disps
    avg_disps = est_d.mean(dim=[1, 2])
    invalid_mono_mask = (error / avg_disps > mono_thres) | (error.isnan()) | (scale < 0) | \
                        (valid_d.sum(dim=[1, 2]) < valid_d.shape[1] * valid_d.shape[2] * 0.5)
    invalid_mono_index, = torch.where(invalid_mono_mask.clone())
    invalid_ii_mask = (ii < 0)
    idx_in_ii = torch.unique(ii)
    for idx in invalid_mono_index:
        invalid_ii_mask = invalid_ii_mask | (ii == idx) | (jj == idx)
    ii_t = ii[~invalid_ii_mask]
    idx_in_ii_t = torch.unique(ii_t)
    valid_eta_mask = torch.tensor([idx in idx_in_ii_t for idx in idx_in_ii])
    out["filter_invalid"] = invalid_mono_mask.numpy()
    out["filter_keep"] = (~invalid_ii_mask).numpy()
    out["filter_eta"] = valid_eta_mask.numpy()
    out["filter_avg"] = avg_disps.numpy()
    print("mask: valid fraction per keyframe", masks.float().mean(dim=[1, 2]).numpy().round(2), "invalid", invalid_mono_mask.numpy().astype(int))
    np.savez_compressed(os.path.join(HERE, "prox.npz"), **out)


if __name__ == "__main__":
    main()
