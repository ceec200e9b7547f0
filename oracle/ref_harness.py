"""Imports the reference's own PyTorch BA (src/geom/ba.py) on CPU, unmodified.

TEST INFRASTRUCTURE ONLY, and usable only where /root/reference is mounted
(the build container): it is how the golden fixtures under tests/golden/ are
produced and how the C oracle is pinned.  Nothing on the GPU box imports it.

Three shims, none of which edits a reference file:
  1. `lietorch_backends` comes from oracle/_ref/ (oracle/build_ref.py).
  2. `torch_scatter` is not vendored (pinned pytorch-scatter 2.1.0,
     reference glorie_env.yaml:159); `scatter_sum` is restated as index_add_.
  3. src/geom/projective_ops.py:105 hard-codes device="cuda" and float32 for
     the stereo constant; the module source is patched IN MEMORY (that one
     line) so the CPU / float64 runs work.  Nothing is copied to disk.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

REF = os.environ.get("GLORIE_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
PYREF = os.path.join(HERE, "_ref", "pyref")  # staged copy for the GPU box (oracle/build_ref.py stage_python)
_loaded = None


def _roots():
    """(directory holding src/geom, directory holding the lietorch package): the reference tree where it
    is mounted, else the staged copy under the git-ignored oracle/_ref/pyref/"""
    if os.path.isdir(os.path.join(REF, "src", "geom")):
        return REF, os.path.join(REF, "thirdparty", "lietorch")
    if os.path.isdir(os.path.join(PYREF, "src", "geom")):
        return PYREF, PYREF
    return None, None


def available():
    return _roots()[0] is not None and os.path.exists(os.path.join(HERE, "_ref", "lietorch_backends.so"))


def _scatter_sum(src, index, dim=-1, out=None, dim_size=None):
    import torch

    if dim < 0:
        dim += src.dim()
    size = list(src.shape)
    size[dim] = int(dim_size) if dim_size is not None else (int(index.max()) + 1 if index.numel() else 0)
    res = torch.zeros(size, dtype=src.dtype, device=src.device)
    return res.index_add_(dim, index.to(src.device).long(), src)


def load():
    """Returns namespace(ba=<src.geom.ba>, lietorch=<lietorch>, pops=<projective_ops>)."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError("reference tree or oracle/_ref/lietorch_backends.so missing; run oracle/build_ref.py")
    import torch  # noqa: F401

    root, lt_root = _roots()
    sys.path.insert(0, os.path.join(HERE, "_ref"))
    sys.path.insert(0, lt_root)
    ts = types.ModuleType("torch_scatter")
    ts.scatter_sum = _scatter_sum
    sys.modules.setdefault("torch_scatter", ts)
    lietorch = importlib.import_module("lietorch")

    pkg = types.ModuleType("src")
    pkg.__path__ = [os.path.join(root, "src")]
    geom = types.ModuleType("src.geom")
    geom.__path__ = [os.path.join(root, "src", "geom")]
    sys.modules["src"] = pkg
    sys.modules["src.geom"] = geom
    pkg.geom = geom

    path = os.path.join(root, "src", "geom", "projective_ops.py")
    code = open(path).read()
    old = 'torch.as_tensor([-0.1, 0.0, 0.0, 0.0, 0.0, 0.0, 1.0], device="cuda")'
    new = "torch.as_tensor([-0.1, 0.0, 0.0, 0.0, 0.0, 0.0, 1.0], device=Gij.data.device, dtype=Gij.data.dtype)"
    assert old in code, "reference projective_ops.py:105 changed"
    pops = types.ModuleType("src.geom.projective_ops")
    pops.__file__ = path
    sys.modules["src.geom.projective_ops"] = pops
    exec(compile(code.replace(old, new), path, "exec"), pops.__dict__)
    geom.projective_ops = pops

    ba = importlib.import_module("src.geom.ba")
    _loaded = types.SimpleNamespace(ba=ba, lietorch=lietorch, pops=pops)
    return _loaded


def run_BA(g, iters=None, dtype=None, fixedp=None):
    """Reference src/geom/ba.py::BA on a synth.Graph for `iters` iterations (CPU).
    Tensors are truncated to t1 as the optimised-window convention D4 requires."""
    import torch

    m = load()
    dtype = dtype or torch.float32
    t1 = g.t1
    iters = g.iters if iters is None else iters
    fixedp = g.t0 if fixedp is None else fixedp
    E, _, ht, wd = g.targets.shape
    target = torch.from_numpy(g.targets[None]).permute(0, 1, 3, 4, 2).contiguous().to(dtype)
    weight = torch.from_numpy(g.weights[None]).permute(0, 1, 3, 4, 2).contiguous().to(dtype)
    eta = torch.from_numpy(g.eta).to(dtype)
    poses = m.lietorch.SE3(torch.from_numpy(g.poses[None, :t1]).to(dtype))
    disps = torch.from_numpy(g.disps[None, :t1]).to(dtype)
    intr = torch.from_numpy(g.intrinsics)[None, None].repeat(1, t1, 1).to(dtype)
    ii = torch.from_numpy(g.ii)
    jj = torch.from_numpy(g.jj)
    for _ in range(iters):
        poses, disps = m.ba.BA(target, weight, eta, poses, disps, intr, ii, jj,
                               lm=g.lm, ep=g.ep, fixedp=fixedp)
    return poses.data[0].numpy(), disps[0].numpy()


def run_DSPO(g, iters=1, dtype=None, ignore_frames=0):
    """Reference src/geom/ba.py::BA_with_scale_shift for `iters` iterations (CPU)."""
    import torch

    m = load()
    dtype = dtype or torch.float32
    n = g.n_frames
    target = torch.from_numpy(g.targets[None]).permute(0, 1, 3, 4, 2).contiguous().to(dtype)
    weight = torch.from_numpy(g.weights[None]).permute(0, 1, 3, 4, 2).contiguous().to(dtype)
    ii = torch.from_numpy(g.ii)
    jj = torch.from_numpy(g.jj)
    kx = torch.unique(ii)
    eta = torch.from_numpy(g.eta[: len(kx)]).to(dtype)
    poses = m.lietorch.SE3(torch.from_numpy(g.poses[None]).to(dtype))
    disps = torch.from_numpy(g.disps[None]).to(dtype)
    intr = torch.from_numpy(g.intrinsics)[None, None].repeat(1, n, 1).to(dtype)
    mono = torch.from_numpy(g.mono_disps[None]).to(dtype)
    scales = torch.from_numpy(g.scales[None]).to(dtype)
    shifts = torch.from_numpy(g.shifts[None]).to(dtype)
    mask = torch.from_numpy(g.valid_depth_mask[None])
    for _ in range(iters):
        poses, disps, wqs = m.ba.BA_with_scale_shift(
            target, weight, eta, poses, disps, intr, ii, jj, mono, scales, shifts, mask,
            ignore_frames, g.lm, g.ep, alpha=g.alpha)
        scales, shifts = wqs[..., 0], wqs[..., 1]
    return disps[0].numpy(), wqs[0].numpy()
