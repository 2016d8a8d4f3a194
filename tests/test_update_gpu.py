"""GPU parity of csrc/update.cu through the C ABI:
  * b200gs_adam_step vs the CPU oracle (oracle/update_oracle.py, pinned to torch.optim.Adam by
    tests/test_update_oracle_cpu.py) and vs torch.optim.Adam itself on the same device -- the call the
    reference makes (train.py:306) -- over several steps with the reference's seven groups, the
    f_dc / f_rest column split of the [P, M, 3] dL_dsh block, zero-gradient (culled) rows, an lr change;
  * b200gs_densify_stats vs the oracle and vs the reference's lines run as torch ops on the device.
Tolerances: moments 1e-5 relative (absolute floor 1e-6 of the largest entry), parameters to 1e-5 of the
distance one step can move them; counters and max_radii2D exact."""
import numpy as np
import pytest
import torch

from test_update_oracle_cpu import LRS, SHAPES, make_grads, make_params

pytestmark = pytest.mark.gpu


def _close_moment(got, want, what):
    want = np.asarray(want)
    np.testing.assert_allclose(np.asarray(got), want, rtol=1e-5, atol=1e-6 * float(np.abs(want).max()) + 1e-37, err_msg=what)


@pytest.mark.parametrize("P", [1, 257, 70_001])
def test_adam_step_matches_oracle_and_torch(P):
    from b200gs.update import GaussianAdam, surfel_param_groups
    from oracle import update_oracle as uo
    params = make_params(P, 3)
    dev = {k: v.cuda().contiguous() for k, v in params.items()}
    opt = GaussianAdam(surfel_param_groups(**dev, lrs=LRS), lr=0.0, eps=1e-15)     # scene/gaussian_model.py:207
    tp = {k: torch.nn.Parameter(v.cuda().clone()) for k, v in params.items()}
    topt = torch.optim.Adam([{"params": [tp[k]], "lr": LRS[k], "name": k} for k in SHAPES], lr=0.0, eps=1e-15)
    o = {k: (v.numpy().copy(), np.zeros(v.shape, np.float32), np.zeros(v.shape, np.float32)) for k, v in params.items()}
    for it in range(1, 6):
        gr = make_grads(P, 200 + it)
        lr_xyz = LRS["xyz"] * (0.25 if it >= 3 else 1.0)
        for grp in opt.param_groups + topt.param_groups:          # update_learning_rate (gaussian_model.py:216-222)
            if grp["name"] == "xyz":
                grp["lr"] = lr_xyz
        dc, rest = uo.split_sh_grad(gr["sh"].numpy())
        gnp = dict(xyz=gr["xyz"].numpy(), f_dc=dc, f_rest=rest, opacity=gr["opacity"].numpy(), scaling=gr["scaling"].numpy(),
                   rotation=gr["rotation"].numpy(), refl=gr["refl"].numpy())
        for k in SHAPES:
            o[k] = uo.adam_step(o[k][0], gnp[k], o[k][1], o[k][2], lr_xyz if k == "xyz" else LRS[k], it)
            tp[k].grad = torch.from_numpy(gnp[k]).cuda()
        topt.step()
        # the product reads dL_dsh as one [P, M, 3] block, like the backward writes it
        opt.step({k: v.cuda().contiguous() for k, v in gr.items()})
        torch.cuda.synchronize()
        for k in SHAPES:
            st = opt.state[dev[k]]              # keyed by the parameter tensor, like torch's
            _close_moment(st["exp_avg"].cpu().numpy(), o[k][1], f"{k} exp_avg it{it} (oracle)")
            _close_moment(st["exp_avg_sq"].cpu().numpy(), o[k][2], f"{k} exp_avg_sq it{it} (oracle)")
            _close_moment(st["exp_avg"].cpu().numpy(), topt.state[tp[k]]["exp_avg"].cpu().numpy(), f"{k} exp_avg it{it} (torch)")
            _close_moment(st["exp_avg_sq"].cpu().numpy(), topt.state[tp[k]]["exp_avg_sq"].cpu().numpy(), f"{k} exp_avg_sq it{it} (torch)")
            lr = lr_xyz if k == "xyz" else LRS[k]
            got = dev[k].cpu().numpy()
            for want, who in ((o[k][0], "oracle"), (tp[k].detach().cpu().numpy(), "torch")):
                assert np.abs(got - want).max() <= 1e-5 * max(LRS[k], lr) * it + 2e-7 * np.abs(want).max(), (k, it, who)
    assert float(opt.state[dev["xyz"]]["step"]) == 5.0 and set(opt.state[dev["xyz"]]) == set(topt.state[tp["xyz"]])


def test_adam_reads_arena_blocks_and_rejects_bad_input():
    import b200gs
    from b200gs.dp import GradArena
    from b200gs.update import GaussianAdam, arena_grads, surfel_param_groups
    P = 1000
    dev = {k: v.cuda().contiguous() for k, v in make_params(P, 9).items()}
    before = {k: v.clone() for k, v in dev.items()}
    arena = GradArena(P, 16, "cuda")
    arena.flat.normal_()
    opt = GaussianAdam(surfel_param_groups(**dev, lrs=LRS), lr=0.0, eps=1e-15)
    opt.step(arena_grads(arena))
    torch.cuda.synchronize()
    # first Adam step: every parameter moves by lr * sign(g) (m / sqrt(v) = +-1)
    g = arena.views()
    for k, gk, sl in (("xyz", "means3D", None), ("f_dc", "sh", slice(0, 1)), ("f_rest", "sh", slice(1, 16)), ("opacity", "opacity", None),
                      ("scaling", "scales", None), ("rotation", "rotations", None), ("refl", "refl", None)):
        gg = g[gk] if sl is None else g[gk][:, sl, :]
        want = before[k] - LRS[k] * torch.sign(gg).reshape(before[k].shape)
        assert torch.allclose(dev[k], want, rtol=0, atol=2e-6 * LRS[k] + 1e-7), k
    with pytest.raises(ValueError):
        opt.step({"xyz": torch.zeros(P, 3)})                    # host tensor: no CPU path
    with pytest.raises(ValueError):
        opt.step({"xyz": torch.zeros((P, 2), device="cuda")})   # wrong size
    bad = b200gs.Adam()
    bad.groups, bad.step, bad.beta1, bad.beta2 = 1, 0, 0.9, 0.999
    assert b200gs.adam_step(bad, None) != 0 and b"step" in b200gs.last_error()


@pytest.mark.parametrize("P", [1, 1000, 100_003])
def test_densify_stats_match_oracle_and_reference_lines(P):
    from b200gs.update import DensifyStats
    from oracle import update_oracle as uo
    g = torch.Generator().manual_seed(11)
    ds = DensifyStats(P)
    st = {k: np.zeros(P, np.float32) for k in DensifyStats.ROWS}
    max_radii2D = torch.zeros(P, device="cuda")
    acc, den = torch.zeros((P, 1), device="cuda"), torch.zeros((P, 1), device="cuda")
    acc_w, den_w = torch.zeros((P, 1), device="cuda"), torch.zeros((P, 1), device="cuda")
    for view in range(4):
        radii = (torch.randint(0, 60, (P,), generator=g) * (torch.rand(P, generator=g) > 0.3)).int()
        vgrad = torch.randn((P, 3), generator=g) * 1e-3
        rw = (torch.rand(P, generator=g) * (torch.rand(P, generator=g) > 0.5)).float()
        # view 2 comes without weights: only the first three statistics move
        uo.densify_stats(st, vgrad.numpy(), radii.numpy(), rw.numpy() if view != 2 else None)
        r_d, v_d, w_d = radii.cuda(), vgrad.cuda(), rw.cuda()
        ds.add(v_d, r_d, w_d if view != 2 else None)
        vis = r_d > 0                       # the reference's lines on the device (train.py:242-243, gaussian_model.py:579-584)
        max_radii2D[vis] = torch.max(max_radii2D[vis], r_d[vis])
        acc[vis] += torch.norm(v_d[vis], dim=-1, keepdim=True)
        den[vis] += 1
        if view != 2:
            mask = w_d > 0.0
            acc_w[mask] += w_d[mask].unsqueeze(-1)
            den_w[mask] += 1
    torch.cuda.synchronize()
    for k, ref in (("max_radii2D", max_radii2D), ("denom", den), ("denom_w", den_w)):
        assert np.array_equal(getattr(ds, k).cpu().numpy(), st[k]), k
        assert torch.equal(getattr(ds, k), ref.reshape(-1)), k
    for k, ref in (("xyz_gradient_accum", acc), ("accum_w", acc_w)):
        np.testing.assert_allclose(getattr(ds, k).cpu().numpy(), st[k], rtol=2e-6, atol=1e-12, err_msg=k)
        assert torch.allclose(getattr(ds, k), ref.reshape(-1), rtol=2e-6, atol=1e-12), k


def test_training_iteration_end_to_end():
    """render -> backward -> statistics -> Adam on the tiny scene: the whole per-iteration device work of
    train.py:134-306 that touches per-Gaussian data, through the drop-in modules."""
    import harness as hz
    import scenes
    from b200gs.update import DensifyStats, GaussianAdam
    sc_np, cam_np = scenes.build("tiny")
    sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
    P = sc["means3D"].shape[0]
    grads = hz.to_dev(scenes.upstream_grads(cam_np["width"], cam_np["height"]))
    C = hz.new_surfel_C()
    shs = sc["shs"]
    f_dc, f_rest = shs[:, :1, :].contiguous(), shs[:, 1:, :].contiguous()
    groups = [dict(params=[sc["means3D"]], lr=LRS["xyz"], name="xyz"), dict(params=[f_dc], lr=LRS["f_dc"], name="f_dc"),
              dict(params=[f_rest], lr=LRS["f_rest"], name="f_rest"), dict(params=[sc["opacities"]], lr=LRS["opacity"], name="opacity")]
    opt, ds = GaussianAdam(groups, lr=0.0, eps=1e-15), DensifyStats(P)
    x0 = sc["means3D"].clone()
    f = hz.surfel_forward(C, sc, cam)
    b = hz.surfel_backward(C, sc, cam, f, grads)
    ds.add(b["means2D"].contiguous(), f["radii"], f["weights"].reshape(-1).contiguous())
    opt.step({"xyz": b["means3D"], "sh": b["sh"], "opacity": b["opacity"]})
    torch.cuda.synchronize()
    vis = f["radii"] > 0
    assert int(ds.denom.sum()) == int(vis.sum()) > 0
    assert torch.equal(ds.max_radii2D, f["radii"].clamp(min=0).float())
    moved = (sc["means3D"] - x0).abs().amax(1) > 0
    assert bool(moved.any()) and not bool((moved & ~(b["means3D"].abs().amax(1) > 0)).any())
