"""One gba_dspo call on config D (for ncu): python profiles/prof_dspo.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "glorie-slam_b200"))
import numpy as np, torch
import geom_ba, synth
g = synth.make_config("D")
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
N = g.n_frames
M = len(np.unique(g.ii))
args = (t(g.targets.transpose(0, 2, 3, 1))[None], t(g.weights.transpose(0, 2, 3, 1))[None], t(g.eta[:M]), t(g.poses)[None],
        t(g.disps)[None], t(g.intrinsics)[None, None].repeat(1, N, 1), t(g.ii), t(g.jj), t(g.mono_disps)[None],
        t(g.scales)[None], t(g.shifts)[None], t(g.valid_depth_mask)[None], 0, g.lm, g.ep)
for _ in range(2):
    geom_ba.BA_with_scale_shift(*args, alpha=g.alpha)
torch.cuda.synchronize()
print("done")
