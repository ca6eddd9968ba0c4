"""CPU test of the multi-rank path (world_size 2, gloo): views are sharded round-robin, every rank
accumulates its views' volume/TF gradients privately, one all-reduce sums them. The per-rank compute
stands in for the CUDA path with the CPU oracle (same C-ABI, host pointers) -- what is tested here
is the host logic in diffdvr_b200/multigpu.py."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_loader as ol
import scenes
from diffdvr_b200 import capi, multigpu

N_VIEWS = 5


def view_scene(v):
    sc = scenes.make_scene(n=12, W=9, H=7, B=N_VIEWS, tf_mode=capi.TF_TEXTURE, R=4, seed=3, opacity=3.0)
    sc.camera_matrix = sc.camera_matrix[v:v + 1].copy()
    return sc


def local_gradients(views):
    o = ol.port(32)
    sc0 = view_scene(0)
    adj_volume = np.zeros(sc0.volume.shape, np.float32)
    adj_tf = np.zeros(sc0.tf.shape, np.float32)
    for v in views:
        sc = view_scene(v)
        c, t = scenes.np_forward(o, sc)
        adj_color = np.sign(c - 0.3).astype(np.float32)
        a = scenes.np_adjoint(o, sc, c, t, adj_color, volume=True, tf=True)
        adj_volume += a["adj_volume"]
        adj_tf += a["adj_tf"]
    return adj_volume, adj_tf


def worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        views = multigpu.shard_views(N_VIEWS, rank, world)
        adj_volume, adj_tf = (torch.from_numpy(a) for a in local_gradients(views))
        # host-sharded parameter: every rank contributes its slice of the volume, all ranks end with all of it
        full = torch.from_numpy(view_scene(0).volume.copy())
        b, e = multigpu.shard_bounds(full.numel(), rank, world)
        gathered = torch.full_like(full, float("nan"))
        multigpu.allgather_parameter(gathered, full.view(-1)[b:e].clone())
        # ... and reads back only its slice of the summed gradient
        mine = multigpu.reduce_scatter_gradient(adj_volume.clone()).clone()
        multigpu.allreduce_gradients([adj_volume, None, adj_tf])
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), adj_volume=adj_volume.numpy(), adj_tf=adj_tf.numpy(),
                 views=np.array(views), gathered=gathered.numpy(), grad_slice=mine.numpy(), bounds=np.array([b, e]))
    finally:
        dist.destroy_process_group()


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_views_partition():
    for n, w in [(64, 8), (5, 2), (3, 4), (0, 2)]:
        shards = [multigpu.shard_views(n, r, w) for r in range(w)]
        assert sorted(v for s in shards for v in s) == list(range(n))
        assert max(len(s) for s in shards) - min(len(s) for s in shards) <= 1
    with pytest.raises(ValueError):
        multigpu.shard_views(4, 2, 2)
    assert list(multigpu.batches([0, 2, 4, 6, 8], 2)) == [[0, 2], [4, 6], [8]]
    with pytest.raises(ValueError):
        list(multigpu.batches([1], 0))


def test_allreduce_is_a_noop_without_process_group():
    t = torch.ones(3)
    multigpu.allreduce_gradients([t])
    assert torch.equal(t, torch.ones(3))


def test_two_ranks_sum_to_the_single_process_gradient(tmp_path):
    world = 2
    mp.spawn(worker, args=(world, free_port(), str(tmp_path)), nprocs=world, join=True)
    want_volume, want_tf = local_gradients(range(N_VIEWS))
    seen = []
    for r in range(world):
        got = np.load(tmp_path / f"rank{r}.npz")
        seen += list(got["views"])
        # every rank holds the full sum; fp32 summation order differs -> gradient tolerance
        assert scenes.rel_l2(got["adj_volume"], want_volume) <= 1e-6
        assert scenes.rel_l2(got["adj_tf"], want_tf) <= 1e-6
        assert np.array_equal(got["gathered"], view_scene(0).volume)
        b, e = got["bounds"]
        assert np.array_equal(got["grad_slice"], got["adj_volume"].reshape(-1)[b:e])  # same collective result, own slice
    assert sorted(seen) == list(range(N_VIEWS))


def test_shard_bounds_cover_the_parameter_once():
    for numel, w in [(256 ** 3, 8), (12 ** 3, 2), (10, 4), (3, 8)]:
        bounds = [multigpu.shard_bounds(numel, r, w) for r in range(w)]
        assert bounds[0][0] == 0 and bounds[-1][1] == numel
        assert all(bounds[r][1] == bounds[r + 1][0] for r in range(w - 1))
        assert all(0 <= e - b <= -(-numel // w) for b, e in bounds)
    # single process: the slice is the whole parameter
    t = torch.arange(6.0)
    full = torch.zeros(6)
    multigpu.allgather_parameter(full, t)
    assert torch.equal(full, t) and multigpu.reduce_scatter_gradient(t) is not None
