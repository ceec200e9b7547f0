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
        out[rank] = (ok_merge, ok_sum)
    finally:
        dist.destroy_process_group()


def test_gloo_world2_merge_and_allreduce():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert dict(out) == {0: (True, True), 1: (True, True)}
