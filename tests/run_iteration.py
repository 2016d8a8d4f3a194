"""GPU harness: ONE TRAINING ITERATION of the reference's loop (train.py:134-306) on a synthetic scene, every
per-iteration device step through this repo's sm_100a kernels, next to the same iteration through the reference's own
pieces on the same GPU:

  step                          this repo                                  reference arm
  rasterize fwd + bwd           diff_surfel_rasterization (libb200gs)      the reference's own package on oracle/_ref/ref_surfel_C
  surface attributes            b200gs.shade.surface_attributes            the torch lines (oracle/shade_oracle.py on the device)
  reflection pass               b200gs.shade.reflect_shade                 torch lines + oracle/_ref/ref_cubemap_C
  L1 + SSIM                     fused_ssim (libb200gs)                     oracle/_ref/ref_ssim_C through the same autograd shape
  densification statistics      b200gs.update.DensifyStats                 train.py:242-243 + add_densification_stats in torch
  Adam                          b200gs.update.GaussianAdam (+ torch Adam   torch.optim.Adam over all groups
                                for the env map)

  python tests/run_iteration.py [c2 bench ...] > gpurun_out/iteration.jsonl

BASELINE.json configs[1] is `c2` (300k Gaussians, 800x800, reflection channels + cubemap lookup, one train iteration).
Per config: median ms per iteration over 20 timed iterations after 5 warm-ups (CUDA events around the whole iteration,
host work included), and the agreement of the two arms after the same number of iterations (loss, parameters)."""
import json
import math
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import harness as hz  # noqa: E402
import scenes  # noqa: E402
from run_configs import timed  # noqa: E402
from test_update_oracle_cpu import LRS  # noqa: E402

L_ENV, DEPTH_RATIO, LAMBDA_DSSIM, LAMBDA_NORMAL, LAMBDA_DIST = 128, 0.0, 0.2, 0.05, 100.0
LAMBDA_NORMAL = float(os.environ.get("ITER_LAMBDA_NORMAL", LAMBDA_NORMAL))      # diagnosis switches
LAMBDA_DIST = float(os.environ.get("ITER_LAMBDA_DIST", LAMBDA_DIST))
LAMBDA_DSSIM = float(os.environ.get("ITER_LAMBDA_DSSIM", LAMBDA_DSSIM))
NO_REFL = os.environ.get("ITER_NO_REFL", "0") == "1"


class View:
    """The camera attributes the renderer reads (scene/cameras.py)."""
    def __init__(self, cam_np):
        W, H = cam_np["width"], cam_np["height"]
        self.image_width, self.image_height = W, H
        self.world_view_transform = torch.from_numpy(cam_np["viewmatrix"]).reshape(4, 4).cuda()
        self.full_proj_transform = torch.from_numpy(cam_np["projmatrix"]).reshape(4, 4).cuda()
        self.camera_center = torch.from_numpy(cam_np["campos"]).cuda()
        self.tanfovx, self.tanfovy = cam_np["tanfovx"], cam_np["tanfovy"]
        w2c = np.asarray(cam_np["w2c"], np.float32)
        self.R, self.T = np.ascontiguousarray(w2c[:3, :3].T), np.ascontiguousarray(w2c[:3, 3])   # 3DGS stores R transposed
        K = np.array([[W / (2 * self.tanfovx), 0, W / 2], [0, H / (2 * self.tanfovy), H / 2], [0, 0, 1]], np.float32)
        self.HWK = (H, W, K)


def make_model(sc_np):
    sc = hz.to_dev(sc_np)
    shs = sc["shs"]
    p = dict(xyz=sc["means3D"].clone(), f_dc=shs[:, :1, :].contiguous(), f_rest=shs[:, 1:, :].contiguous(),
             opacity=sc["opacities"].clone(), scaling=sc["scales"].clone(), rotation=sc["rotations"].clone(), refl=sc["refl"].clone())
    g = torch.Generator().manual_seed(3)
    env_tex = (torch.rand((6, 3, L_ENV, L_ENV), generator=g) - 0.5).cuda()
    return {k: v.requires_grad_(True) for k, v in p.items()}, env_tex.requires_grad_(True), torch.zeros(3, device="cuda", requires_grad=True)


def build_arm(arm, sc_np, cam_np):
    from oracle import shade_oracle as so
    if arm == "new":
        import diff_surfel_rasterization as dsr
    else:      # the reference's own python package on its own extension (tests/refpy.py)
        import refpy
        ref_arm = refpy.load_arm("ref")
        if ref_arm is None:
            return None
        dsr = ref_arm.diff_surfel_rasterization
    view = View(cam_np)
    W, H = view.image_width, view.image_height
    P = sc_np["means3D"].shape[0]
    params, env_tex, env_fail = make_model(sc_np)
    gt = torch.rand((3, H, W), generator=torch.Generator().manual_seed(11)).cuda()
    bg = torch.zeros(3, device="cuda")
    mask = torch.ones((P, 3), dtype=torch.bool, device="cuda")
    rs = dsr.GaussianRasterizationSettings(
        image_height=H, image_width=W, tanfovx=view.tanfovx, tanfovy=view.tanfovy, bg=bg, scale_modifier=1.0,
        viewmatrix=view.world_view_transform, projmatrix=view.full_proj_transform, sh_degree=3, campos=view.camera_center,
        prefiltered=False, debug=False)
    new = arm == "new"
    C_new = hz.new_surfel_C()
    C_ref = None if new else hz.ref_module("ref_surfel_C")
    ref_cm = None if new else hz.ref_module("ref_cubemap_C")
    ref_ss = None if new else hz.ref_module("ref_ssim_C")
    if not new and (C_ref is None or ref_cm is None or ref_ss is None):
        return None

    if new:
        import fused_ssim as fs
        from b200gs.shade import camera_constants, reflect_shade, surface_attributes, surface_camera_constants
        from b200gs.update import DensifyStats, GaussianAdam, surfel_param_groups
        opt = GaussianAdam(surfel_param_groups(**params, lrs=LRS), lr=0.0, eps=1e-15)
        stats = DensifyStats(P)
        cam_s = surface_camera_constants(view.world_view_transform, view.full_proj_transform, W, H)
        cam_r = camera_constants(view.HWK, view.R, view.T)

        class Env:
            params = {"Cubemap_texture": env_tex, "Cubemap_failv": env_fail}
            interp_id, seamless = 1, 1
        env_opt = torch.optim.Adam([env_tex, env_fail], lr=0.01, eps=1e-15)
    else:
        C1, C2 = 0.01 ** 2, 0.03 ** 2

        class RefSSIM(torch.autograd.Function):        # submodules/fused-ssim/fused_ssim/__init__.py:8-36 on the reference backend
            @staticmethod
            def forward(ctx, a, b):
                m, d1, d2, d3 = ref_ss.fusedssim(C1, C2, a, b, True)
                ctx.save_for_backward(a.detach(), b, d1, d2, d3)
                return m

            @staticmethod
            def backward(ctx, g):
                a, b, d1, d2, d3 = ctx.saved_tensors
                return ref_ss.fusedssim_backward(C1, C2, a, b, g.contiguous(), d1, d2, d3), None

        class RefLookup(torch.autograd.Function):      # cubemapencoder/cubemap_encoder.py:19-64 on the reference backend
            @staticmethod
            def forward(ctx, inputs, emb, failv):
                inputs, emb = inputs.contiguous(), emb.contiguous()
                B = inputs.shape[0]
                o = torch.empty((3, B), device="cuda")
                ref_cm.cubemap_encode_forward(inputs, emb, failv, o, 1, 1, B, 3, L_ENV)
                ctx.save_for_backward(inputs, emb)
                return o.permute(1, 0)

            @staticmethod
            def backward(ctx, g):
                inputs, emb = ctx.saved_tensors
                ge, gi, gf = torch.zeros_like(emb), torch.zeros_like(inputs), torch.zeros(3, device="cuda")
                ref_cm.cubemap_encode_backward(g.permute(1, 0).contiguous(), inputs, emb, ge, gi, gf, 1, 1, inputs.shape[0], 3, L_ENV)
                return gi, ge, gf

        opt = torch.optim.Adam([{"params": [params[k]], "lr": LRS[k], "name": k} for k in params] +
                               [{"params": [env_tex, env_fail], "lr": 0.01, "name": "env"}], lr=0.0, eps=1e-15)
        max_radii2D = torch.zeros(P, device="cuda")
        acc, den, aw, dw = (torch.zeros((P, 1), device="cuda") for _ in range(4))
        Rs, Ts = torch.from_numpy(view.R).cuda(), torch.from_numpy(view.T).cuda()
        i, j = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32), indexing="xy")
        pixel_camera = torch.tensor(np.dot(np.stack([i, j, np.ones_like(i)], axis=2), np.linalg.inv(view.HWK[2]).T)).cuda()

    state = {"loss": None}

    def iteration(do_step=True):
        try:
            m2d = torch.zeros_like(params["xyz"], requires_grad=True)
            shs = torch.cat((params["f_dc"], params["f_rest"]), dim=1)                    # get_features
            base, radii, allmap, refl_map, weights = dsr.GaussianRasterizer(rs)(
                means3D=params["xyz"], means2D=m2d, opacities=params["opacity"], shs=shs, refl_strengths=params["refl"],
                scales=params["scaling"], rotations=params["rotation"], env_scope_mask=mask)
            if new:
                sa = surface_attributes(allmap, view, DEPTH_RATIO, cam=cam_s)
                n_hw3, surf_normal, rend_dist = sa["render_normal_hw3"], sa["surf_normal"], sa["rend_dist"]
                image, _ = reflect_shade(Env, view.HWK, view.R, view.T, n_hw3, base, refl_map, cam=cam_r)
                ssim_value = fs.fused_ssim(image.unsqueeze(0), gt.unsqueeze(0))
            else:
                torch.set_default_device("cuda")
                try:
                    n_hw3, _, surf_normal = so.surface_attributes(allmap, view.world_view_transform, view.full_proj_transform, DEPTH_RATIO)
                finally:
                    torch.set_default_device("cpu")
                rend_dist = allmap[6:7]
                Rt = Rs.T
                rays_o = (-Rt.T @ Ts.unsqueeze(-1)).flatten()
                rays_d = (pixel_camera - Ts[None, None]).reshape(-1, 3) @ Rt - rays_o[None]
                rays_d = (rays_d / torch.norm(rays_d, dim=1, keepdim=True)).reshape(H, W, 3)
                refl = rays_d - 2 * n_hw3 * torch.sum(rays_d * n_hw3, dim=-1, keepdim=True)
                refl_color = torch.sigmoid(RefLookup.apply(refl.reshape(-1, 3), env_tex, env_fail)).reshape(H, W, 3).permute(2, 0, 1)
                image = (1 - refl_map) * base + refl_map * refl_color
                ssim_value = RefSSIM.apply(image.unsqueeze(0), gt.unsqueeze(0)).mean()
            if NO_REFL:
                image = base
            Ll1 = torch.abs(image - gt).mean()
            normal_loss = LAMBDA_NORMAL * (1 - (n_hw3.permute(2, 0, 1) * surf_normal).sum(dim=0)).mean()
            loss = (1.0 - LAMBDA_DSSIM) * Ll1 + LAMBDA_DSSIM * (1.0 - ssim_value) + normal_loss + LAMBDA_DIST * rend_dist.mean()
            loss.backward()
            if not do_step:          # gradients of the first iteration, for the parity gate
                g = {k: v.grad.detach().clone() for k, v in params.items()}
                g["means2D"], g["loss"] = m2d.grad.detach().clone(), loss.detach().clone()
                if env_tex.grad is not None:
                    g["env_texture"] = env_tex.grad.detach().clone()
                for v in list(params.values()) + [env_tex, env_fail]:
                    v.grad = None
                return g
            with torch.no_grad():
                if new:
                    stats.add(m2d.grad.contiguous(), radii, weights.reshape(-1).contiguous())
                    opt.step()
                    env_opt.step()
                    env_opt.zero_grad(set_to_none=True)
                    opt.zero_grad(set_to_none=True)
                else:
                    vis = radii > 0
                    max_radii2D[vis] = torch.max(max_radii2D[vis], radii[vis])
                    acc[vis] += torch.norm(m2d.grad[vis], dim=-1, keepdim=True)
                    den[vis] += 1
                    msk = (weights > 0.0).squeeze()
                    aw[msk] += weights[msk].reshape(-1, 1)
                    dw[msk] += 1
                    opt.step()
                    opt.zero_grad(set_to_none=True)
            state["loss"] = loss.detach()
        finally:
            pass
    return iteration, params, state


def run(name):
    sc_np, cam_np = scenes.build(name)
    out = {"config": name, "P": int(sc_np["means3D"].shape[0]), "W": cam_np["width"], "H": cam_np["height"], "env_L": L_ENV}
    res, first = {}, {}
    for arm in ("new", "reference"):
        built = build_arm(arm, sc_np, cam_np)
        if built is None:
            out["reference"] = "oracle/_ref extensions not built"
            continue
        it, params, state = built
        first[arm] = it(do_step=False)
        out[arm + "_ms"] = timed(it, warmup=5, iters=20)           # 25 iterations on both arms
        torch.cuda.synchronize()
        res[arm] = (float(state["loss"]), {k: v.detach().clone() for k, v in params.items()})
        del it, params, state, built
        torch.cuda.empty_cache()
    if "reference" in res:
        out["speedup"] = out["reference_ms"] / out["new_ms"]
        out["first_iteration"] = {"loss": {a: float(first[a]["loss"]) for a in first},
                                  "grad_rel_l2": {k: hz.rel_l2(first["new"][k], first["reference"][k]) for k in first["new"] if k != "loss"}}
        out["loss_after_25"] = {"new": res["new"][0], "reference": res["reference"][0]}
        # Adam moves a parameter by ~lr per step whatever the gradient's size, so noise-level gradients (atomic order) can
        # flip single steps: the mean is the meaningful figure, the max is bounded by 2 x 25 steps
        out["param_mean_abs_diff_over_lr_after_25"] = {k: float((res["new"][1][k] - res["reference"][1][k]).abs().mean()) / LRS[k] for k in LRS}
        out["param_max_abs_diff_over_lr_after_25"] = {k: float((res["new"][1][k] - res["reference"][1][k]).abs().max()) / LRS[k] for k in LRS}
    return out


if __name__ == "__main__":
    for n in (sys.argv[1:] or ["c2", "bench"]):
        print(json.dumps(run(n)), flush=True)
        torch.cuda.empty_cache()
