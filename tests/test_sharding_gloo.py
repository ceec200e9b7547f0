"""CPU, world_size 2, gloo: the host-side logic of the edge-sharded path (partitioning, the
system all-reduce and the owned-row merge).  The CUDA kernels themselves are exercised by the
emulated two-rank test in test_gpu_parity.py."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import sharded
import synth


def test_partition_covers_and_balances():
    g = synth.make_graph(40, 400, 5, seed=1, ht=8, wd=12)
    kx = sharded.frame_list(g.ii, g.t0, g.t1)
    np.testing.assert_array_equal(kx, np.unique(np.concatenate([np.arange(g.t0, g.t1), g.ii])))
    for world in (1, 2, 4, 8):
        owner = sharded.partition(g.ii, kx, world)
        assert owner.shape == kx.shape and owner.min() >= 0 and owner.max() < world
        masks = [sharded.local_edges(g.ii, kx, owner, r) for r in range(world)]
        assert np.sum(masks, axis=0).tolist() == [1] * g.n_edges  # every edge on exactly one rank
        loads = np.array([m.sum() for m in masks])
        assert loads.max() - loads.min() <= np.bincount(np.searchsorted(kx, g.ii)).max()
        # all edges of a source keyframe are on its owner
        for r, m in enumerate(masks):
            assert set(owner[np.searchsorted(kx, g.ii[m])]) <= {r}


def test_partition_handles_edgeless_and_hub_frames():
    ii = np.array([3] * 50 + [1, 2, 2], dtype=np.int64)
    kx = sharded.frame_list(ii, 1, 8)
    owner = sharded.partition(ii, kx, 2)
    hub = owner[np.searchsorted(kx, 3)]
    assert owner[np.searchsorted(kx, 1)] != hub and owner[np.searchsorted(kx, 2)] != hub
    assert len(set(owner[np.searchsorted(kx, [4, 5, 6, 7])])) == 2  # edge-less frames are spread


def test_plan_assigns_keyframes_to_the_segment_they_touch():
    """host half of the segmented plan on a hand-made ordering: 4 segments of 32 pose positions (3 tiles each),
    separators and one long-range pair in the closure block"""
    P, t0 = 150, 1
    seg_pos, nseg, seg_tiles = 32, 4, 3
    perm = np.zeros(P, dtype=np.int64)
    closure = set(list(range(30, 36)) + list(range(66, 72)) + list(range(102, 108)) + [10, 140])  # separators + 2 long-range
    nxt, clo = [0, 0, 0, 0], nseg * seg_pos
    for pose in range(P):
        if pose in closure:
            perm[pose] = clo
            clo += 1
        else:
            sg = 0 if pose < 30 else 1 if pose < 66 else 2 if pose < 102 else 3
            perm[pose] = sg * seg_pos + nxt[sg]
            nxt[sg] += 1
    ii, jj = [], []
    for i in range(P + t0):
        for dlt in (1, 2, 3):
            for j in (i - dlt, i + dlt):
                if 0 <= j < P + t0:
                    ii.append(i); jj.append(j)
    ii += [11, 141]; jj += [141, 11]  # the long-range pair (frames 11 and 141 = poses 10 and 140)
    ii, jj = np.array(ii), np.array(jj)
    kx = sharded.frame_list(ii, t0, P + t0)
    for world in (2, 4):
        pl = sharded.plan_from_perm(ii, jj, kx, t0, P + t0, perm, nseg, seg_tiles, world)
        assert pl is not None and pl.n_segments == 4 and pl.seg_of_rank == [(r * (4 // world), 4 // world) for r in range(world)]
        per = nseg // world
        # every keyframe's own pose and all its destinations lie in its owner's segments or in the closure block
        for k, f in enumerate(kx):
            touched = np.concatenate([[f], jj[ii == f]]) - t0
            touched = touched[(touched >= 0) & (touched < P)]
            segs = perm[touched] // seg_pos
            segs = segs[perm[touched] < nseg * seg_pos]
            assert np.all(segs // per == pl.owner[k]), (f, segs, pl.owner[k])
        assert np.all(pl.pose_owner[sorted(closure)] == 0)
    assert sharded.plan_from_perm(ii, jj, kx, t0, P + t0, perm, nseg, seg_tiles, 3) is None  # 4 segments, 3 ranks
    # an edge across two segments that is not in the closure block: no plan
    bad_ii, bad_jj = np.append(ii, 20), np.append(jj, 50)
    assert sharded.plan_from_perm(bad_ii, bad_jj, kx, t0, P + t0, perm, nseg, seg_tiles, 2) is None


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = synth.make_graph(10, 30, 2, seed=4, ht=8, wd=12)
        kx = sharded.frame_list(g.ii, g.t0, g.t1)
        owner = sharded.partition(g.ii, kx, world)
        owned = torch.from_numpy((owner == rank).astype(np.uint8))
        kx_t = torch.from_numpy(kx)
        # each rank "updates" only its rows; rows outside kx stay as they are
        disps = torch.from_numpy(g.disps.copy())
        disps[kx_t[owned.bool()]] += float(rank + 1)
        sharded.merge_owned_rows(disps, kx_t, owned)
        expect = g.disps.copy()
        for r in range(world):
            expect[kx[owner == r]] += float(r + 1)
        ok_merge = bool(np.allclose(disps.numpy(), expect))
        # system all-reduce: sum of per-rank contributions, identical on every rank
        view = torch.full((1000,), float(rank + 1), dtype=torch.float64)
        sharded.allreduce_system(view)
        ok_sum = bool(torch.all(view == sum(range(1, world + 1))))
        # pose merge of the segmented solve: every rank moved its own poses, rank 0 also the shared ones
        poses = torch.from_numpy(g.poses.copy())
        pose_owner = torch.from_numpy((np.arange(g.t1 - g.t0) % world).astype(np.int32))
        poses[g.t0:g.t1][pose_owner == rank] += float(rank + 1)
        sharded.merge_poses(poses, g.t0, g.t1, pose_owner, rank)
        expect_p = g.poses.copy()
        for r in range(world):
            expect_p[g.t0:g.t1][pose_owner.numpy() == r] += float(r + 1)
        ok_sum = ok_sum and bool(np.allclose(poses.numpy(), expect_p))
        out[rank] = (ok_merge, ok_sum)
    finally:
        dist.destroy_process_group()


def test_gloo_world2_merge_and_allreduce():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert dict(out) == {0: (True, True), 1: (True, True)}
