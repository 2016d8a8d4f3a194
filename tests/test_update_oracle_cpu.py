"""Pins oracle/update_oracle.py (CPU): against torch.optim.Adam built exactly as the reference builds
it (scene/gaussian_model.py:196-207) and against the reference's statistics lines run as literal
torch (train.py:242-243, scene/gaussian_model.py:579-584)."""
import numpy as np
import torch

from oracle import update_oracle as uo

LRS = dict(xyz=1.6e-4, f_dc=2.5e-3, f_rest=2.5e-3 / 20.0, opacity=0.05, scaling=0.005, rotation=0.001, refl=0.005)
SHAPES = dict(xyz=(3,), f_dc=(1, 3), f_rest=(15, 3), opacity=(1,), scaling=(2,), rotation=(4,), refl=(1,))


def make_params(P, seed):
    g = torch.Generator().manual_seed(seed)
    return {k: torch.randn((P,) + s, generator=g) for k, s in SHAPES.items()}


def make_grads(P, seed, sparse=0.4):
    """Gradients as the backward leaves them: dL_dsh for f_dc+f_rest together, zero rows for culled Gaussians."""
    g = torch.Generator().manual_seed(seed)
    vis = (torch.rand(P, generator=g) > sparse).float()
    out = {}
    for k, s in dict(xyz=(3,), sh=(16, 3), opacity=(1,), scaling=(2,), rotation=(4,), refl=(1,)).items():
        t = torch.randn((P,) + s, generator=g) * 10.0 ** float(torch.randint(-6, 1, (1,), generator=g))
        out[k] = t * vis.view((P,) + (1,) * len(s))
    return out


def test_adam_matches_torch_optim_adam():
    P = 257
    params = make_params(P, 1)
    tp = {k: torch.nn.Parameter(v.clone()) for k, v in params.items()}
    opt = torch.optim.Adam([{"params": [tp[k]], "lr": LRS[k], "name": k} for k in SHAPES], lr=0.0, eps=1e-15)
    o = {k: (v.numpy().copy(), np.zeros(v.shape, np.float32), np.zeros(v.shape, np.float32)) for k, v in params.items()}
    for it in range(1, 8):
        gr = make_grads(P, 100 + it)
        dc, rest = uo.split_sh_grad(gr["sh"].numpy())
        gnp = dict(xyz=gr["xyz"].numpy(), f_dc=dc, f_rest=rest, opacity=gr["opacity"].numpy(), scaling=gr["scaling"].numpy(),
                   rotation=gr["rotation"].numpy(), refl=gr["refl"].numpy())
        if it == 4:   # the xyz lr is rescheduled every iteration in the reference (update_learning_rate)
            for g_ in opt.param_groups:
                if g_["name"] == "xyz":
                    g_["lr"] = LRS["xyz"] * 0.5
        for k in SHAPES:
            tp[k].grad = torch.from_numpy(gnp[k]).clone()
            lr = LRS[k] * (0.5 if (k == "xyz" and it >= 4) else 1.0)
            o[k] = uo.adam_step(o[k][0], gnp[k], o[k][1], o[k][2], lr, it)
        opt.step()
        for k in SHAPES:
            st = opt.state[tp[k]]
            # moments: relative, with an absolute floor for entries where g and m cancel
            for got, want in ((o[k][1], st["exp_avg"].numpy()), (o[k][2], st["exp_avg_sq"].numpy())):
                np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6 * float(np.abs(want).max()), err_msg=f"{k} it{it}")
            # one step moves a parameter by <= lr * O(1); compare the positions to a fraction of that
            assert np.abs(o[k][0] - tp[k].detach().numpy()).max() <= 1e-5 * LRS[k] * it + 2e-7 * np.abs(o[k][0]).max(), (k, it)


def test_adam_zero_gradient_rows_still_move_on_momentum():
    """Dense Adam: a Gaussian culled this view (zero gradient) keeps decaying its moments and moving."""
    p = np.ones((4, 3), np.float32)
    g = np.zeros((4, 3), np.float32); g[0] = 1.0
    p1, m1, v1 = uo.adam_step(p, g, np.zeros_like(p), np.zeros_like(p), 0.1, 1)
    assert np.all(p1[1:] == 1.0) and np.all(p1[0] < 1.0)
    p2, m2, v2 = uo.adam_step(p1, np.zeros_like(g), m1, v1, 0.1, 2)
    assert np.all(p2[0] < p1[0]) and np.all(m2[0] < m1[0]) and np.all(p2[1:] == 1.0)


def test_densify_stats_match_reference_lines():
    P = 1000
    g = torch.Generator().manual_seed(5)
    max_radii2D = torch.zeros(P)
    xyz_gradient_accum, denom = torch.zeros((P, 1)), torch.zeros((P, 1))
    accum_w, denom_w = torch.zeros((P, 1)), torch.zeros((P, 1))
    st = {k: np.zeros(P, np.float32) for k in ("xyz_gradient_accum", "denom", "accum_w", "denom_w", "max_radii2D")}
    for view in range(5):
        radii = (torch.randint(0, 40, (P,), generator=g) * (torch.rand(P, generator=g) > 0.3)).int()
        vgrad = torch.randn((P, 3), generator=g) * 1e-3
        rw = (torch.rand((P, 1), generator=g) * (torch.rand((P, 1), generator=g) > 0.5)).float()
        # --- the reference's lines, verbatim semantics (train.py:242-243, gaussian_model.py:579-584) ---
        visibility_filter = radii > 0
        max_radii2D[visibility_filter] = torch.max(max_radii2D[visibility_filter], radii[visibility_filter])
        xyz_gradient_accum[visibility_filter] += torch.norm(vgrad[visibility_filter], dim=-1, keepdim=True)
        denom[visibility_filter] += 1
        mask = (rw > 0.0).squeeze()
        accum_w[mask] += rw[mask].unsqueeze(-1).reshape(-1, 1)
        denom_w[mask] += 1
        # --- oracle ---
        uo.densify_stats(st, vgrad.numpy(), radii.numpy(), rw.numpy().reshape(-1))
    np.testing.assert_array_equal(st["max_radii2D"], max_radii2D.numpy())
    np.testing.assert_array_equal(st["denom"], denom.numpy().reshape(-1))
    np.testing.assert_array_equal(st["denom_w"], denom_w.numpy().reshape(-1))
    np.testing.assert_allclose(st["xyz_gradient_accum"], xyz_gradient_accum.numpy().reshape(-1), rtol=1e-6)
    np.testing.assert_allclose(st["accum_w"], accum_w.numpy().reshape(-1), rtol=1e-6)
