"""One droid_backends.ba call on a synthetic config (for ncu): python profiles/prof_one_call.py E 2"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "glorie-slam_b200"))
import numpy as np, torch
import droid_backends, synth
name = sys.argv[1] if len(sys.argv) > 1 else "E"
calls = int(sys.argv[2]) if len(sys.argv) > 2 else 2
g = synth.make_config(name)
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
d = [t(x) for x in (g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj)]
droid_backends.RETURN_UPDATES = False
for _ in range(calls):
    d[0].copy_(t(g.poses)); d[1].copy_(t(g.disps))
    droid_backends.ba(*d, g.t0, g.t1, g.iters, g.lm, g.ep, False, False)
torch.cuda.synchronize()
print("done")
