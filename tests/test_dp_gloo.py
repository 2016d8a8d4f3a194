"""CPU, world_size 2, gloo: the view-parallel plumbing of b200gs.dp -- view sharding, the flat
gradient arena and its all-reduce -- checked against a single-process sum.  Per-view gradients
come from the CPU oracle (tests may use it); the GPU kernels are not involved here."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import scenes
from b200gs.dp import GradArena, shard_views


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _view_grads(view, n_views, P):
    from oracle import surfel_oracle as so
    torch.set_num_threads(1)
    sc, _ = scenes.build("tiny", P=P)
    cam = scenes.ring_camera(view, n_views, 48, 32, fovx=0.9)
    g = scenes.upstream_grads(48, 32, seed=view)
    _, gr = so.forward_backward(sc, cam, g)
    return gr


def _worker(rank, world, port, n_views, P, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    arena = GradArena(P, 16, "cpu")
    for v in shard_views(n_views, rank, world):
        gr = _view_grads(v, n_views, P)
        for k, dst in arena.views().items():
            dst += torch.from_numpy(gr[k]).reshape(dst.shape)
    arena.all_reduce()
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), arena.flat.numpy())
    dist.destroy_process_group()


def test_shard_views_partition():
    for world in (1, 2, 4, 8):
        seen = sorted(v for r in range(world) for v in shard_views(32, r, world))
        assert seen == list(range(32))
        assert max(len(shard_views(32, r, world)) for r in range(world)) == 32 // world


def test_arena_layout():
    a = GradArena(10, 16, "cpu")
    assert a.floats_per_gaussian == 59 and a.flat.numel() == 590
    v = a.views()
    assert v["sh"].shape == (10, 16, 3) and v["means3D"].shape == (10, 3) and v["refl"].shape == (10, 1)
    v["rotations"].fill_(1.0)
    assert float(a.flat.sum()) == 40.0 and a.flat.data_ptr() == v["means3D"].data_ptr()


@pytest.mark.timeout(600)
def test_two_rank_allreduce_matches_single_process(tmp_path):
    n_views, P, world = 4, 300, 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n_views, P, str(tmp_path)), nprocs=world, join=True)
    r0, r1 = np.load(tmp_path / "rank0.npy"), np.load(tmp_path / "rank1.npy")
    assert np.array_equal(r0, r1)                      # every rank holds the same reduced gradients
    ref = GradArena(P, 16, "cpu")
    for v in range(n_views):
        gr = _view_grads(v, n_views, P)
        for k, dst in ref.views().items():
            dst += torch.from_numpy(gr[k]).reshape(dst.shape)
    assert np.allclose(r0, ref.flat.numpy(), rtol=1e-5, atol=1e-6)
    assert np.abs(r0).sum() > 0


def _stats_worker(rank, world, port, P, out_dir):
    from b200gs.update import all_reduce_stats
    from oracle import update_oracle as uo
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    st = _local_stats(uo, shard_views(6, rank, world), P)
    flat = torch.from_numpy(np.stack([st[k] for k in ("xyz_gradient_accum", "denom", "accum_w", "denom_w", "max_radii2D")]))
    all_reduce_stats(flat)
    np.save(os.path.join(out_dir, f"stats{rank}.npy"), flat.numpy())
    dist.destroy_process_group()


def _local_stats(uo, views, P):
    st = {k: np.zeros(P, np.float32) for k in ("xyz_gradient_accum", "denom", "accum_w", "denom_w", "max_radii2D")}
    for v in views:
        rng = np.random.default_rng(900 + v)
        radii = (rng.integers(0, 50, P) * (rng.random(P) > 0.4)).astype(np.int32)
        uo.densify_stats(st, rng.standard_normal((P, 3)).astype(np.float32), radii,
                         (rng.random(P) * (rng.random(P) > 0.5)).astype(np.float32))
    return st


@pytest.mark.timeout(600)
def test_two_rank_densify_stats_allreduce(tmp_path):
    """rows 0-3 of the [5, P] statistics buffer are summed over ranks, row 4 (max_radii2D) maximised:
    the result equals one process that saw every view."""
    from oracle import update_oracle as uo
    P, world = 500, 2
    mp.spawn(_stats_worker, args=(world, _free_port(), P, str(tmp_path)), nprocs=world, join=True)
    r0, r1 = np.load(tmp_path / "stats0.npy"), np.load(tmp_path / "stats1.npy")
    assert np.array_equal(r0, r1)
    one = _local_stats(uo, range(6), P)
    assert np.array_equal(r0[1], one["denom"]) and np.array_equal(r0[3], one["denom_w"])
    assert np.array_equal(r0[4], one["max_radii2D"])
    assert np.allclose(r0[0], one["xyz_gradient_accum"], rtol=1e-6) and np.allclose(r0[2], one["accum_w"], rtol=1e-6)


def test_arena_lands_in_model_gradients():
    """The two ways an all-reduced arena reaches the optimizer (b200gs.dp.GradArena): bind_grads -- the leaves'
    .grad ARE the arena blocks -- and backward_through -- the chain rule through the model's activations
    (sigmoid / exp / normalize, scene/gaussian_model.py:124-147) applied once to the batch sum."""
    P = 7
    a = GradArena(P, 16, "cpu", accumulate=True)
    leaves = {"means3D": torch.zeros(P, 3, requires_grad=True), "sh": torch.zeros(P, 16, 3, requires_grad=True),
              "opacity": torch.zeros(P, 1, requires_grad=True)}
    a.bind_grads(leaves)
    a.view("opacity").add_(2.0)                       # what a backward kernel does
    assert leaves["opacity"].grad.data_ptr() == a.view("opacity").data_ptr()
    assert float(leaves["opacity"].grad.sum()) == 2.0 * P
    a.zero_()
    assert float(leaves["opacity"].grad.abs().sum()) == 0.0
    with pytest.raises(ValueError):
        a.bind_grads({"scales": torch.zeros(P, 3, requires_grad=True)})
    # activations in between: d/d raw = Jacobian^T x (gradient with respect to the activated value)
    raw_o = torch.randn(P, 1, requires_grad=True)
    raw_s = torch.randn(P, 2, requires_grad=True)
    raw_r = torch.randn(P, 4, requires_grad=True)
    act = {"opacity": torch.sigmoid(raw_o), "scales": torch.exp(raw_s), "rotations": torch.nn.functional.normalize(raw_r)}
    g = {k: torch.randn_like(v) for k, v in act.items()}
    for k in act:
        a.view(k).copy_(g[k])
    a.backward_through(act)
    so = torch.sigmoid(raw_o.detach())
    assert torch.allclose(raw_o.grad, g["opacity"] * so * (1 - so), atol=1e-6)
    assert torch.allclose(raw_s.grad, g["scales"] * torch.exp(raw_s.detach()), atol=1e-6)
    want = torch.autograd.grad(torch.nn.functional.normalize(raw_r), raw_r, g["rotations"])[0]
    assert torch.allclose(raw_r.grad, want, atol=1e-6)
