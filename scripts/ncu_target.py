"""One warmed-up step of a chosen piece of the path between cudaProfilerStart / Stop, for
`ncu --profile-from-start off --set full ... python scripts/ncu_target.py <target>` (scripts/gpu_ncu_round.sh).

  surfel   bench workload, view 0: forward + backward of the surfel rasterizer (default culled, depth-sliced mode)
  exact    the same in exact-lists mode
  gauss    c3 through the 3DGS (EWA) rasterizer, forward + backward
  cubemap  c2's environment lookup: 640k rays, L = 128, C = 3, seamless bilinear, forward + backward
  glue     surface attributes + reflection pass at 1920x1080, forward + backward (b200gs.shade)
  ssim     fused SSIM forward + backward, 3 x 1920x1080
  adam     fused Adam over the 59 floats of 3M Gaussians + densification statistics
"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import harness as hz  # noqa: E402
import scenes  # noqa: E402


def profiled(fn, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    fn()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()


def main(target):
    if target in ("surfel", "exact"):
        sc_np, cam_np = scenes.build("bench")
        sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
        grads = hz.to_dev(scenes.upstream_grads(cam_np["width"], cam_np["height"]))
        C = hz.new_surfel_C()

        def step():
            f = hz.surfel_forward(C, sc, cam, exact=(target == "exact"))
            hz.surfel_backward(C, sc, cam, f, grads)
        profiled(step)
    elif target == "gauss":
        sc_np, cam_np = scenes.build("c3")
        sc, cam = hz.gauss_scene(hz.to_dev(sc_np)), hz.to_dev(cam_np)
        grads = hz.to_dev(hz.gauss_grads(cam_np["width"], cam_np["height"]))
        C = hz.new_gauss_C()

        def step():
            f = hz.gauss_forward(C, sc, cam)
            hz.gauss_backward(C, sc, cam, f, grads)
        profiled(step)
    elif target == "cubemap":
        import _cubemapencoder as cm
        from run_configs import cubemap_step
        step, _, _ = cubemap_step(cm)
        profiled(step)
    elif target == "glue":
        from b200gs.shade import camera_constants, reflect_shade, surface_attributes, surface_camera_constants
        from test_shade_gpu import _Env, _camera, _inputs
        H, W, L = 1080, 1920, 128
        HWK, R, T = _camera(H, W, 5)
        n, base, r, tex, fail = _inputs(H, W, L, 6)
        ln = [t.cuda().requires_grad_(True) for t in (n, base, r, tex, fail)]
        env, cam = _Env(ln[3], ln[4]), camera_constants(HWK, R, T)
        sc_np, cam_np = scenes.build("c4")
        f = hz.surfel_forward(hz.new_surfel_C(), hz.to_dev(sc_np), hz.to_dev(cam_np))
        allmap = f["others"].detach().clone().requires_grad_(True)
        view = torch.from_numpy(cam_np["viewmatrix"]).reshape(4, 4)
        proj = torch.from_numpy(cam_np["projmatrix"]).reshape(4, 4)
        scam = surface_camera_constants(view, proj, W, H)

        def step():
            for t in ln + [allmap]:
                t.grad = None
            sa = surface_attributes(allmap, None, 0.0, cam=scam)
            final, _ = reflect_shade(env, HWK, R, T, sa["render_normal_hw3"], ln[1], ln[2], cam=cam)
            (final.sum() + sa["surf_normal"].sum() + sa["surf_depth"].sum()).backward()
        profiled(step)
    elif target == "ssim":
        import fused_ssim as fs
        g = torch.Generator().manual_seed(1)
        a = torch.rand((1, 3, 1080, 1920), generator=g).cuda().requires_grad_(True)
        b = torch.rand((1, 3, 1080, 1920), generator=g).cuda()

        def step():
            a.grad = None
            fs.fused_ssim(a, b).backward()
        profiled(step)
    elif target == "adam":
        from b200gs.update import DensifyStats, GaussianAdam, surfel_param_groups
        from test_update_oracle_cpu import LRS, SHAPES
        P = 3_000_000
        g = torch.Generator(device="cuda").manual_seed(1)
        params = {k: torch.randn((P,) + s, device="cuda", generator=g) for k, s in SHAPES.items()}
        grads = {k: torch.randn((P,) + s, device="cuda", generator=g) * 1e-3
                 for k, s in dict(xyz=(3,), sh=(16, 3), opacity=(1,), scaling=(2,), rotation=(4,), refl=(1,)).items()}
        opt = GaussianAdam(surfel_param_groups(**params, lrs=LRS), lr=0.0, eps=1e-15)
        ds = DensifyStats(P)
        radii = torch.randint(0, 60, (P,), device="cuda", generator=g).int()
        vg = torch.randn((P, 3), device="cuda", generator=g)
        w = torch.rand(P, device="cuda", generator=g)

        def step():
            opt.step(grads)
            ds.add(vg, radii, w)
        profiled(step)
    else:
        raise SystemExit(__doc__)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "surfel")
