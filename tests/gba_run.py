"""GPU-side helpers: run libgba (through the Python mirror or the raw C ABI) on a synth.Graph."""
import ctypes
from types import SimpleNamespace

import numpy as np
import torch

import _gba


def to_dev(g, dev="cuda:0"):
    t = lambda a, dt=None: torch.from_numpy(np.ascontiguousarray(a)).to(dev) if dt is None else torch.from_numpy(
        np.ascontiguousarray(a)).to(dev).to(dt)
    return SimpleNamespace(
        poses=t(g.poses), disps=t(g.disps), intr=t(g.intrinsics), sens=t(g.disps_sens), targets=t(g.targets),
        weights=t(g.weights), eta=t(g.eta), ii=t(g.ii), jj=t(g.jj))


def run_ba(g, mode="droid_cuda", iters=None, motion_only=False, depth_only=False, lm=None, ep=None, eta=None):
    import droid_backends

    d = to_dev(g)
    if eta is not None:
        d.eta = torch.from_numpy(np.ascontiguousarray(eta)).cuda()
    droid_backends.SEMANTICS = mode
    try:
        dx, dz = droid_backends.ba(d.poses, d.disps, d.intr, d.sens, d.targets, d.weights, d.eta, d.ii, d.jj, g.t0,
                                   g.t1, g.iters if iters is None else iters, g.lm if lm is None else lm,
                                   g.ep if ep is None else ep, motion_only, depth_only)
    finally:
        droid_backends.SEMANTICS = "droid_cuda"
    torch.cuda.synchronize()
    ctx = _gba.context(0, g.n_frames, g.n_edges, g.disps.shape[1], g.disps.shape[2])
    info = _gba.query(ctx)
    return SimpleNamespace(poses=d.poses.cpu().numpy(), disps=d.disps.cpu().numpy(), dx=dx.cpu().numpy(),
                           dz=None if dz is None else dz.cpu().numpy(), info=info, ctx=ctx, dev=d)


def get_array(ctx, what, shape, dtype):
    out = np.zeros(shape, dtype=dtype)
    _gba.check(ctx, _gba.lib().gba_get(ctx, what, out.ctypes.data_as(ctypes.c_void_p), out.nbytes))
    return out


def split_phase(g, mode=0, owned_masks=None, iters=1):
    """begin / build / (sum systems) / solve through the C ABI.  With owned_masks = [m0, m1, ...]
    emulates edge-sharded ranks on ONE GPU: one context per rank, each given only the edges whose
    source keyframe it owns; the packed systems are summed (the all-reduce) before every solve.
    Returns the final state of rank 0 merged with the owned disparities of the others."""
    L = _gba.lib()
    N, ht, wd = g.disps.shape
    kx = np.unique(np.concatenate([np.arange(g.t0, g.t1), g.ii])) if mode == 0 else np.unique(g.ii)
    masks = owned_masks or [np.ones(len(kx), dtype=np.uint8)]
    ranks = []
    for m in masks:
        sel = np.isin(g.ii, kx[m.astype(bool)])
        h = ctypes.c_void_p()
        rc = L.gba_create(ctypes.byref(h), 0, N, g.n_edges, ht, wd)
        assert rc == 0, L.gba_last_error(None)
        d = SimpleNamespace(
            h=h, sel=sel, mask=torch.from_numpy(m.astype(np.uint8)).cuda(),
            poses=torch.from_numpy(g.poses).cuda(), disps=torch.from_numpy(g.disps).cuda(),
            intr=torch.from_numpy(g.intrinsics).cuda(), sens=torch.from_numpy(g.disps_sens).cuda(),
            targets=torch.from_numpy(np.ascontiguousarray(g.targets[sel])).cuda(),
            weights=torch.from_numpy(np.ascontiguousarray(g.weights[sel])).cuda(),
            eta=torch.from_numpy(g.eta).cuda(), ii=torch.from_numpy(np.ascontiguousarray(g.ii[sel])).cuda(),
            jj=torch.from_numpy(np.ascontiguousarray(g.jj[sel])).cuda(), kx=torch.from_numpy(kx).cuda(),
            ii_all=torch.from_numpy(g.ii).cuda(), jj_all=torch.from_numpy(g.jj).cuda())
        _gba.check(h, L.gba_ba_begin(h, d.poses.data_ptr(), d.disps.data_ptr(), d.intr.data_ptr(), d.sens.data_ptr(),
                                     d.targets.data_ptr(), d.weights.data_ptr(), d.eta.data_ptr(), d.ii.data_ptr(),
                                     d.jj.data_ptr(), N, int(sel.sum()), g.eta.shape[0], g.t0, g.t1, g.lm, g.ep, 0, 0,
                                     mode, d.kx.data_ptr(), len(kx), d.mask.data_ptr(), d.ii_all.data_ptr(),
                                     d.jj_all.data_ptr(), g.n_edges, None))
        ranks.append(d)
    systems = None
    for _ in range(iters):
        views = []
        for d in ranks:
            _gba.check(d.h, L.gba_ba_build(d.h))
            views.append(_gba.system_view(d.h, "cuda:0"))
        assert len(set(v.numel() for v in views)) == 1, "ranks disagree on the packed system size"
        # the all-reduce: every rank pads the diagonal of the last tile with 1, the sum stays SPD
        total = torch.stack(views).sum(dim=0)
        for v in views:
            v.copy_(total)
        torch.cuda.synchronize()
        if systems is None:
            n = 6 * (g.t1 - g.t0)
            sysm = np.zeros(n * n + n)
            _gba.check(ranks[0].h, L.gba_get(ranks[0].h, _gba.GET_SYSTEM, sysm.ctypes.data_as(ctypes.c_void_p),
                                             sysm.nbytes))
            systems = (sysm[: n * n].reshape(n, n), sysm[n * n:])
        for d in ranks:
            _gba.check(d.h, L.gba_ba_solve(d.h))
    torch.cuda.synchronize()
    poses = ranks[0].poses.cpu().numpy()
    disps = ranks[0].disps.cpu().numpy()
    for d, m in zip(ranks[1:], masks[1:]):
        own = kx[m.astype(bool)]
        disps[own] = d.disps.cpu().numpy()[own]
    infos = [_gba.query(d.h) for d in ranks]
    for d in ranks:
        L.gba_destroy(d.h)
    return SimpleNamespace(poses=poses, disps=disps, M=systems[0], rhs=systems[1], infos=infos, kx=kx)


def segmented(g, world, iters=2):
    """The SEGMENTED sharded solve, `world` ranks emulated on ONE GPU through the C ABI: every rank builds and
    eliminates its own band segments, the closure block + rhs ranges are summed (the all-reduce), every rank
    finishes.  Returns None when the graph's ordering offers no plan for `world` ranks, else the merged state."""
    import sharded

    L = _gba.lib()
    N, ht, wd = g.disps.shape
    kx, plan = sharded.plan(g.ii, g.jj, g.t0, g.t1, N, ht, wd, world, "cuda:0")
    if plan is None:
        return None
    ranks = []
    for r in range(world):
        m = (plan.owner == r)
        sel = sharded.local_edges(g.ii, kx, plan.owner, r)
        h = ctypes.c_void_p()
        assert L.gba_create(ctypes.byref(h), 0, N, g.n_edges, ht, wd) == 0, L.gba_last_error(None)
        t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
        d = SimpleNamespace(h=h, own=m, poses=t(g.poses), disps=t(g.disps), intr=t(g.intrinsics), sens=t(g.disps_sens),
                            targets=t(g.targets[sel]), weights=t(g.weights[sel]), eta=t(g.eta), ii=t(g.ii[sel]),
                            jj=t(g.jj[sel]), kx=t(kx), mask=t(m.astype(np.uint8)), ii_all=t(g.ii), jj_all=t(g.jj))
        _gba.check(h, L.gba_ba_begin(h, d.poses.data_ptr(), d.disps.data_ptr(), d.intr.data_ptr(), d.sens.data_ptr(),
                                     d.targets.data_ptr(), d.weights.data_ptr(), d.eta.data_ptr(), d.ii.data_ptr(),
                                     d.jj.data_ptr(), N, int(sel.sum()), g.eta.shape[0], g.t0, g.t1, g.lm, g.ep, 0, 0,
                                     0, d.kx.data_ptr(), len(kx), d.mask.data_ptr(), d.ii_all.data_ptr(),
                                     d.jj_all.data_ptr(), g.n_edges, None))
        lo, n = plan.seg_of_rank[r]
        _gba.check(h, L.gba_ba_set_shard(h, lo, n, int(r == 0)))
        ranks.append(d)
    shared_bytes = None
    for _ in range(iters):
        views = []
        for d in ranks:
            _gba.check(d.h, L.gba_ba_build(d.h))
            _gba.check(d.h, L.gba_ba_eliminate(d.h))
            views.append(_gba.system_view(d.h, "cuda:0", 1))
        assert len(set(v.numel() for v in views)) == 1
        total = torch.stack(views).sum(dim=0)
        for v in views:
            v.copy_(total)
        shared_bytes = (views[0].numel() * 8, _gba.system_view(ranks[0].h, "cuda:0", 0).numel() * 8)
        for d in ranks:
            _gba.check(d.h, L.gba_ba_finish(d.h))
    torch.cuda.synchronize()
    poses, disps = ranks[0].poses.cpu().numpy(), ranks[0].disps.cpu().numpy()
    for r, d in enumerate(ranks[1:], start=1):
        disps[kx[d.own]] = d.disps.cpu().numpy()[kx[d.own]]
        mine = np.where(plan.pose_owner == r)[0] + g.t0
        poses[mine] = d.poses.cpu().numpy()[mine]
    infos = [_gba.query(d.h) for d in ranks]
    for d in ranks:
        L.gba_destroy(d.h)
    return SimpleNamespace(poses=poses, disps=disps, infos=infos, plan=plan, shared_bytes=shared_bytes)
