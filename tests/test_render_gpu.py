"""The reference's UNMODIFIED Python on top of this repo's packages (north_star: "so gaussian_renderer/__init__.py,
train.py and render.py call it unchanged"), compared with the same Python on top of the reference's own extensions.

tests/refpy.py imports the byte-for-byte copies of gaussian_renderer/__init__.py, scene/gaussian_model.py,
scene/cameras.py, utils/*.py (oracle/_ref/py, made by oracle/build_ref.py) twice -- arm "ref" with
oracle/_ref/ref_surfel_C.so / ref_cubemap_C.so behind `diff_surfel_rasterization._C` / `_cubemapencoder`, arm "new"
with gaussian-splatting-reflection_b200/python first on sys.path -- and this file calls the reference's
render() (gaussian_renderer/__init__.py:42-219) and render_fast() (:221-325) on both with identical parameters:

  * every key of the returned dict: radii / visibility exact, images <= 1e-5 abs;
  * the training loss of train.py:167-190 (L1 + SSIM with the fork's own utils/loss_utils.py + normal consistency),
    backward through the reference's glue, and ALL parameter gradients (the raw, pre-activation parameters of
    GaussianModel + the cube-map texture + viewspace_points) <= 1e-4 relative (or 4x the reference's own
    run-to-run difference);
  * the call variants of render(): initial_stage, env_scope_radius > 0 (a (P,) mask), pipe.compute_cov3D_python
    (surfel transMat_precomp), depth_ratio = 1, and 2DGS's distortion loss on rend_dist;
  * render_fast() under torch.no_grad() (eval_fps.py:25-59);
  * this repo's fused glue kernels (b200gs.shade: surface_attributes, reflect_shade) against what the reference's
    render() computed -- which pins their oracle with reference-executed values;
  * the reference's optimizer surgery (densification_postfix, prune_points, reset_opacity;
    scene/gaussian_model.py:383-478) on b200gs.update.GaussianAdam against torch.optim.Adam.
"""
import numpy as np
import pytest
import torch

import harness as hz
import refpy
import scenes

pytestmark = pytest.mark.gpu

IMG_ATOL = 1e-5
GRAD_RTOL = 1e-4
L_ENV = 128


@pytest.fixture(scope="module")
def arms():
    if not refpy.available():
        pytest.skip("oracle/_ref/py not present (python oracle/build_ref.py needs /root/reference)")
    ref, new = refpy.load_arm("ref"), refpy.load_arm("new")
    if ref is None or new is None:
        pytest.skip("reference extensions not built")
    return ref, new


def _inputs(name, seed=3):
    sc_np, cam_np = scenes.build(name)
    g = torch.Generator().manual_seed(seed)
    env_tex = torch.rand((6, 3, L_ENV, L_ENV), generator=g) - 0.5          # cubemap_encoder.py:94 init
    gt = torch.rand((3, cam_np["height"], cam_np["width"]), generator=g)
    return sc_np, cam_np, env_tex, gt


def _loss(arm, pkg, gt, opt, lambda_dist=0.0):
    """train.py:167-190 with the fork's own torch ssim (utils/loss_utils.py:63-101) in both arms."""
    image = pkg["render"]
    Ll1 = arm.loss_utils.l1_loss(image, gt)
    loss = (1.0 - opt.lambda_dssim) * Ll1 + opt.lambda_dssim * (1.0 - arm.loss_utils.ssim(image, gt))
    normal_error = (1 - (pkg["rend_normal"] * pkg["surf_normal"]).sum(dim=0))[None]
    loss = loss + opt.lambda_normal * normal_error.mean()
    if lambda_dist:
        loss = loss + lambda_dist * pkg["rend_dist"].mean()
    return loss


def _params(gm):
    p = dict(xyz=gm._xyz, f_dc=gm._features_dc, f_rest=gm._features_rest, scaling=gm._scaling, rotation=gm._rotation,
             opacity=gm._opacity, refl=gm._refl_strength)
    p["env_texture"] = gm.env_map.params["Cubemap_texture"]
    p["env_failv"] = gm.env_map.params["Cubemap_failv"]
    return p


def _run_arm(arm, inputs, initial_stage=False, lambda_dist=0.0, render_kw=None, pipe_kw=None, bg=(1.0, 1.0, 1.0)):
    sc_np, cam_np, env_tex, gt = inputs
    gm = refpy.make_model(arm, sc_np, env_tex, cubemap_resol=L_ENV)
    cam = refpy.make_camera(arm, cam_np, image=gt)
    pipe, opt = arm.pipe(), arm.opt()
    for k, v in (pipe_kw or {}).items():
        setattr(pipe, k, v)
    bgt = torch.tensor(bg, dtype=torch.float32, device="cuda")
    arm.general_utils.pixel_camera = None                                   # the fork caches the ray grid per process
    pkg = arm.render(cam, gm, pipe, bgt, initial_stage=initial_stage, **(render_kw or {}))
    loss = _loss(arm, pkg, cam.original_image, opt, lambda_dist)
    loss.backward()
    grads = {k: (v.grad.detach().clone() if v.grad is not None else None) for k, v in _params(gm).items()}
    grads["viewspace_points"] = pkg["viewspace_points"].grad.detach().clone()
    out = {k: (v.detach() if torch.is_tensor(v) else v) for k, v in pkg.items()}
    return out, grads, float(loss), (gm, cam, pipe, bgt)


def _compare(rn, rr, gn, gr, gr2, what):
    assert set(rn) == set(rr), what
    for k in rr:
        if k == "viewspace_points":
            continue
        a, b = rn[k], rr[k]
        assert a.shape == b.shape and a.dtype == b.dtype, (what, k)
        if a.dtype in (torch.bool, torch.int32):
            assert torch.equal(a, b), (what, k)
        elif k == "gaussian_weights":
            assert int((b > a).sum()) == 0, (what, k)                        # the reference's max is racy: ref <= exact max
        else:
            assert hz.max_abs(a, b) <= IMG_ATOL, (what, k, hz.max_abs(a, b))
    report = {}
    for k in gr:
        if gr[k] is None:
            assert gn[k] is None, (what, k)
            continue
        noise, err = hz.rel_l2(gr2[k], gr[k]), hz.rel_l2(gn[k], gr[k])
        report[k] = (f"{err:.2e}", f"noise {noise:.2e}")
        assert torch.isfinite(gn[k]).all(), (what, k)
        assert err <= max(GRAD_RTOL, 4 * noise), (what, k, err, noise)
    print(what, report)


CASES = [
    ("small", dict()),
    ("small", dict(initial_stage=True)),
    ("small", dict(render_kw=dict(env_scope_center=[0.1, 0.0, 0.2], env_scope_radius=0.9))),       # (P,) bool mask
    ("small", dict(pipe_kw=dict(compute_cov3D_python=True))),                                       # transMat_precomp
    ("small", dict(pipe_kw=dict(depth_ratio=1.0))),
    ("small", dict(lambda_dist=100.0)),
    ("c2", dict()),
    ("c2", dict(lambda_dist=100.0)),
]


@pytest.mark.parametrize("name,kw", CASES, ids=[f"{n}-{'-'.join(sorted(k)) or 'default'}" for n, k in CASES])
def test_reference_render_unmodified(name, kw, arms):
    ref, new = arms
    inputs = _inputs(name)
    rr, gr, lr, _ = _run_arm(ref, inputs, **kw)
    _, gr2, _, _ = _run_arm(ref, inputs, **kw)
    rn, gn, ln, _ = _run_arm(new, inputs, **kw)
    assert abs(ln - lr) <= 2e-6 * max(1.0, abs(lr)), (ln, lr)
    _compare(rn, rr, gn, gr, gr2, f"render() {name} {kw}")


@pytest.mark.parametrize("initial_stage", [False, True])
def test_reference_render_fast_no_grad(initial_stage, arms):
    """eval_fps.py:25-59: forward only under torch.no_grad()."""
    ref, new = arms
    inputs = _inputs("c2")
    outs = []
    for arm in (ref, new):
        sc_np, cam_np, env_tex, gt = inputs
        gm = refpy.make_model(arm, sc_np, env_tex, cubemap_resol=L_ENV)
        cam = refpy.make_camera(arm, cam_np, image=gt)
        arm.general_utils.pixel_camera = None
        with torch.no_grad():
            outs.append(arm.render_fast(cam, gm, arm.pipe(), torch.zeros(3, device="cuda"), initial_stage=initial_stage))
    rr, rn = outs
    assert set(rr) == set(rn)
    for k in rr:
        assert not rn[k].requires_grad
        assert hz.max_abs(rn[k], rr[k]) <= IMG_ATOL, (k, hz.max_abs(rn[k], rr[k]))


def test_fused_glue_kernels_against_reference_render(arms):
    """b200gs.shade.surface_attributes / reflect_shade fuse gaussian_renderer/__init__.py:143-201.
    Their oracle (oracle/shade_oracle.py) restates those lines; here the lines THEMSELVES run (the reference's
    render() on the reference's extensions) and the kernels are compared with what they returned, forward and
    backward, on c2."""
    from b200gs import shade
    ref, _ = arms
    inputs = _inputs("c2")
    sc_np, cam_np, env_tex, gt = inputs
    gm = refpy.make_model(ref, sc_np, env_tex, cubemap_resol=L_ENV)
    cam = refpy.make_camera(ref, cam_np, image=gt)
    pipe = ref.pipe()
    ref.general_utils.pixel_camera = None
    # capture what render() receives from the rasterizer by wrapping the rasterizer module's forward
    captured = {}
    GR = ref.diff_surfel_rasterization.GaussianRasterizer
    orig_forward = GR.forward

    def spy(self, *a, **k):
        out = orig_forward(self, *a, **k)
        captured["raster"] = out
        return out
    GR.forward = spy
    try:
        pkg = ref.render(cam, gm, pipe, torch.ones(3, device="cuda"))
    finally:
        GR.forward = orig_forward
    base, radii, allmap, refl_map, weights = captured["raster"]
    up = {k: torch.randn_like(pkg[k]) for k in ("render", "rend_normal", "surf_depth", "surf_normal", "refl_color_map")}
    ref_scalar = sum((pkg[k] * up[k]).sum() for k in up)
    g_ref = torch.autograd.grad(ref_scalar, [allmap, base, refl_map, gm.env_map.params["Cubemap_texture"]], retain_graph=False)

    allmap_n = allmap.detach().clone().requires_grad_(True)
    base_n = base.detach().clone().requires_grad_(True)
    refl_n = refl_map.detach().clone().requires_grad_(True)
    new_arm = refpy.load_arm("new")
    env_new = new_arm.cubemapencoder.CubemapEncoder(output_dim=3, resolution=L_ENV).cuda()
    with torch.no_grad():
        env_new.params["Cubemap_texture"].copy_(env_tex.cuda())
    sa = shade.surface_attributes(allmap_n, cam, pipe.depth_ratio)
    final, refl_color = shade.reflect_shade(env_new, cam.HWK, cam.R, cam.T, sa["render_normal_hw3"], base_n, refl_n)
    got = {"render": final, "rend_normal": sa["rend_normal"], "surf_depth": sa["surf_depth"], "surf_normal": sa["surf_normal"],
           "refl_color_map": refl_color}
    for k in up:
        assert got[k].shape == pkg[k].shape, k
        d = hz.max_abs(got[k], pkg[k].detach())
        # surf_normal normalises cross products of finite differences of neighbouring surface points: the differences
        # cancel, so a last-place change in a point moves a component of the unit normal by up to ~1e-4
        tol = {"surf_depth": 2e-5 * float(pkg[k].abs().max()), "surf_normal": 1e-4}.get(k, 2e-5)
        assert d <= tol, (k, d)
    new_scalar = sum((got[k] * up[k]).sum() for k in up)
    g_new = torch.autograd.grad(new_scalar, [allmap_n, base_n, refl_n, env_new.params["Cubemap_texture"]])
    for name, a, b in zip(("allmap", "base", "refl_map", "env_texture"), g_new, g_ref):
        # depth_to_normal normalises cross products of finite differences: a few border / degenerate pixels carry
        # huge gradients that swamp a global norm, so compare the bulk (all but the 1e-4 largest reference entries)
        lim = torch.quantile(b.abs().flatten()[:: max(1, b.numel() // 4_000_000)].float(), 0.9999)
        sel = b.abs() <= lim
        err = float(((a - b)[sel]).double().norm() / b[sel].double().norm())
        assert err <= 1e-3, (name, err)


def test_reference_optimizer_surgery_on_gaussian_adam(arms):
    """GaussianAdam must survive the reference's densify / prune / reset-opacity surgery, which
    looks optimizer state up BY PARAMETER TENSOR (scene/gaussian_model.py:399,:414,:467).  Two reference
    GaussianModels with identical parameters, one on torch.optim.Adam (training_setup as is), one with the optimizer
    swapped for GaussianAdam built from the same group list; identical gradient streams; then the reference's own
    densification_postfix / prune_points / reset_opacity on both and more steps.  Parameters must stay equal (to the
    Adam kernel's parity with torch: bit-identical in practice) and the state shapes must follow the parameters."""
    from b200gs.update import GaussianAdam
    _, new = arms
    sc_np, _ = scenes.build("tiny")
    P0 = sc_np["means3D"].shape[0]
    models = []
    env_tex = torch.rand((6, 3, 16, 16), generator=torch.Generator().manual_seed(1)) - 0.5
    for use_new in (False, True):
        gm = refpy.make_model(new, sc_np, env_tex, cubemap_resol=16)
        opt_args = new.opt()
        gm.spatial_lr_scale = 1.0
        gm.training_setup(opt_args)                              # builds torch.optim.Adam(l, lr=0.0, eps=1e-15)
        if use_new:
            groups = [{k: v for k, v in g.items() if k in ("params", "lr", "name")} for g in gm.optimizer.param_groups]
            gm.optimizer = GaussianAdam(groups, lr=0.0, eps=1e-15)
        models.append(gm)
    g = torch.Generator(device="cuda").manual_seed(9)

    def step_both(n):
        for _ in range(n):
            P = models[0]._xyz.shape[0]
            grads = {k: torch.randn(v.shape, device="cuda", generator=g) * 1e-3 for k, v in _params(models[0]).items()}
            for gm in models:
                for k, v in _params(gm).items():
                    v.grad = grads[k].clone()
                gm.optimizer.step()
                gm.optimizer.zero_grad(set_to_none=True)
            assert models[1]._xyz.shape[0] == P

    def check(tag):
        for (k, a), (_, b) in zip(_params(models[0]).items(), _params(models[1]).items()):
            assert a.shape == b.shape, (tag, k)
            assert torch.allclose(a, b, rtol=0, atol=1e-7), (tag, k, float((a.detach() - b.detach()).abs().max()))
        for grp in models[1].optimizer.param_groups:
            for p in grp["params"]:
                st = models[1].optimizer.state.get(p)
                if st:
                    assert st["exp_avg"].shape == p.shape and st["exp_avg_sq"].shape == p.shape, (tag, grp["name"])

    step_both(3)
    check("after 3 steps")
    # densification_postfix: append 100 clones (scene/gaussian_model.py:480-503)
    sel = torch.arange(100, device="cuda")
    for gm in models:
        gm.densification_postfix(gm._xyz[sel], gm._features_dc[sel], gm._features_rest[sel], gm._opacity[sel],
                                 gm._scaling[sel], gm._rotation[sel], gm._refl_strength[sel])
    assert models[1]._xyz.shape[0] == P0 + 100
    step_both(2)
    check("after densification_postfix")
    mask = torch.zeros(P0 + 100, dtype=torch.bool, device="cuda")
    mask[::7] = True
    for gm in models:
        gm.prune_points(mask)
    step_both(2)
    check("after prune_points")
    for gm in models:
        gm.reset_opacity()
    step_both(2)
    check("after reset_opacity")
    # capture / restore round trip of the optimizer state (scene/gaussian_model.py:55-94)
    sd = models[1].optimizer.state_dict()
    torch_sd = models[0].optimizer.state_dict()
    assert set(sd["state"][0]) == set(torch_sd["state"][0])
    models[1].optimizer.load_state_dict(torch_sd)               # interchangeable with torch's
    models[1].optimizer.load_state_dict(sd)
    step_both(1)
    check("after state_dict round trip")
