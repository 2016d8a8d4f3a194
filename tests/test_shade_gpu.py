"""GPU parity of the fused reflection pass (csrc/cubemap.cu reflect_shade_*) through the
C ABI / b200gs.shade: against the CPU oracle (oracle/shade_oracle.py, float64) on small images, and on the device
against the reference's torch lines (gaussian_renderer/__init__.py:22-35, :201, utils/general_utils.py:177-197 restated)
driving the UNMODIFIED reference cubemap extension (oracle/_ref/ref_cubemap_C) at 800x800.  Tolerances: images 2e-5 abs
(fast-math lookup, like the reference's -use_fast_math build), gradients 1e-3 of the largest entry vs the fp64 oracle
and vs the reference composition (direction gradients divide by small face coordinates)."""
import numpy as np
import pytest
import torch

import harness as hz

pytestmark = pytest.mark.gpu


def _camera(H, W, seed=0):
    rng = np.random.default_rng(seed)
    f = 0.9 * W
    K = np.array([[f, 0, W / 2], [0, f, H / 2], [0, 0, 1]], np.float32)
    a = rng.normal(size=(3, 3))
    Q, _ = np.linalg.qr(a)
    if np.linalg.det(Q) < 0:
        Q[:, 0] = -Q[:, 0]
    return (H, W, K), Q.astype(np.float32), rng.normal(size=3).astype(np.float32)


def _inputs(H, W, L, seed=0):
    g = torch.Generator().manual_seed(seed)
    n = torch.randn((H, W, 3), generator=g)
    n = n / n.norm(dim=-1, keepdim=True)
    n[0, 0] = 0.0                                   # an unnormalised / empty normal (background pixel)
    base = torch.rand((3, H, W), generator=g)
    r = torch.rand((1, H, W), generator=g)
    tex = torch.rand((6, 3, L, L), generator=g) - 0.5
    fail = torch.randn(3, generator=g)
    return n, base, r, tex, fail


class _Env:
    """Stand-in for CubemapEncoder: the attributes reflect_shade reads."""
    def __init__(self, tex, fail, interp_id=1, seamless=1):
        self.params = {"Cubemap_texture": tex, "Cubemap_failv": fail}
        self.interp_id, self.seamless = interp_id, seamless


@pytest.mark.parametrize("H,W,L,mode", [(24, 33, 8, (1, 1)), (17, 40, 16, (1, 0)), (9, 9, 4, (0, 1))])
def test_against_cpu_oracle(H, W, L, mode):
    from b200gs.shade import reflect_shade
    from oracle import shade_oracle as so
    HWK, R, T = _camera(H, W, 1)
    n, base, r, tex, fail = _inputs(H, W, L, 2)
    up_f = torch.randn((3, H, W), generator=torch.Generator().manual_seed(3))
    up_e = torch.randn((3, H, W), generator=torch.Generator().manual_seed(4))
    # oracle, float64
    lo = [t.double().requires_grad_(True) for t in (n, base, r, tex, fail)]
    fo, eo = so.reflect_shade(lo[3], lo[4], HWK, R, T, lo[0], lo[1], lo[2], *mode)
    ((fo * up_f.double()).sum() + (eo * up_e.double()).sum()).backward()
    # kernel
    ld = [t.cuda().requires_grad_(True) for t in (n, base, r, tex, fail)]
    fn, en = reflect_shade(_Env(ld[3], ld[4], *mode), HWK, R, T, ld[0], ld[1], ld[2])
    ((fn * up_f.cuda()).sum() + (en * up_e.cuda()).sum()).backward()
    assert float((fn.cpu().double() - fo).abs().max()) <= 2e-5 and float((en.cpu().double() - eo).abs().max()) <= 2e-5
    for name, a, b in zip(("normal", "base", "refl", "texture", "failv"), ld, lo):
        ref = b.grad
        err = float((a.grad.cpu().double() - ref).abs().max())
        assert err <= 1e-3 * float(ref.abs().max()) + 1e-6, (name, err, float(ref.abs().max()))


def test_against_reference_lines_on_reference_cubemap_extension():
    refC = hz.ref_module("ref_cubemap_C")
    if refC is None:
        pytest.skip("oracle/_ref/ref_cubemap_C not built")
    from b200gs.shade import reflect_shade
    H = W = 800
    L = 128
    HWK, R, T = _camera(H, W, 5)
    n, base, r, tex, fail = _inputs(H, W, L, 6)
    up = torch.randn((3, H, W), generator=torch.Generator().manual_seed(7)).cuda()

    class RefLookup(torch.autograd.Function):       # cubemapencoder/cubemap_encoder.py:19-64 on the reference's backend
        @staticmethod
        def forward(ctx, inputs, emb, failv):
            inputs, emb = inputs.contiguous(), emb.contiguous()
            B = inputs.shape[0]
            out = torch.empty((3, B), device="cuda")
            refC.cubemap_encode_forward(inputs, emb, failv, out, 1, 1, B, 3, L)
            ctx.save_for_backward(inputs, emb)
            return out.permute(1, 0)

        @staticmethod
        def backward(ctx, g):
            inputs, emb = ctx.saved_tensors
            go = g.permute(1, 0).contiguous()
            ge, gi, gf = torch.zeros_like(emb), torch.zeros_like(inputs), torch.zeros(3, device="cuda")
            refC.cubemap_encode_backward(go, inputs, emb, ge, gi, gf, 1, 1, inputs.shape[0], 3, L)
            return gi, ge, gf

    lr = [t.cuda().requires_grad_(True) for t in (n, base, r, tex, fail)]
    Rt, Tt = torch.from_numpy(R).cuda().T, torch.from_numpy(T).cuda()
    i, j = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32), indexing="xy")
    pixel_camera = torch.tensor(np.dot(np.stack([i, j, np.ones_like(i)], axis=2), np.linalg.inv(HWK[2]).T)).cuda()
    rays_o = (-Rt.T @ Tt.unsqueeze(-1)).flatten()
    rays_d = (pixel_camera - Tt[None, None]).reshape(-1, 3) @ Rt - rays_o[None]
    rays_d = (rays_d / torch.norm(rays_d, dim=1, keepdim=True)).reshape(H, W, 3)
    refl = rays_d - 2 * lr[0] * torch.sum(rays_d * lr[0], dim=-1, keepdim=True)
    refl_color = torch.sigmoid(RefLookup.apply(refl.reshape(-1, 3), lr[3], lr[4])).reshape(H, W, 3).permute(2, 0, 1)
    final_ref = (1 - lr[2]) * lr[1] + lr[2] * refl_color
    (final_ref * up).sum().backward()

    ln = [t.cuda().requires_grad_(True) for t in (n, base, r, tex, fail)]
    fn, en = reflect_shade(_Env(ln[3], ln[4]), HWK, R, T, ln[0], ln[1], ln[2])
    (fn * up).sum().backward()
    assert float((fn - final_ref).abs().max()) <= 2e-5 and float((en - refl_color).abs().max()) <= 2e-5
    for name, a, b in zip(("normal", "base", "refl", "texture", "failv"), ln, lr):
        # a handful of pixels sit on texel / face boundaries where a 1e-7 change of the direction picks other taps
        d = (a.grad - b.grad).abs()
        tol = 1e-3 * float(b.grad.abs().max()) + 1e-6
        assert float((d > tol).float().mean()) <= 1e-4, (name, float(d.max()), tol)


def test_errors_and_empty():
    import b200gs
    from b200gs.shade import reflect_shade
    HWK, R, T = _camera(8, 8, 0)
    n, base, r, tex, fail = _inputs(8, 8, 4, 0)
    with pytest.raises(RuntimeError):
        reflect_shade(_Env(tex, fail), HWK, R, T, n, base, r)                       # host tensors: no CPU path
    with pytest.raises(RuntimeError):
        reflect_shade(_Env(tex.cuda()[:, :2], fail.cuda()), HWK, R, T, n.cuda(), base.cuda(), r.cuda())
    assert b200gs.reflect_shade_forward(4, 4, None, None, None, None, None, None, None, None, 4, 1, 1, None, None, None) != 0


# ---------------------------------------------------------------------------------------------------------------
# surface attributes (second half of row 2)
def _view(H, W, seed=0):
    """world_view_transform / full_proj_transform in the transposed storage of scene/cameras.py."""
    import scenes
    cam = scenes.make_camera((2.5 + 0.1 * seed, 1.0, 1.5), W, H, fovx=0.9)
    return torch.from_numpy(cam["viewmatrix"]).reshape(4, 4).clone(), torch.from_numpy(cam["projmatrix"]).reshape(4, 4).clone()


def _allmap(H, W, seed):
    g = torch.Generator().manual_seed(seed)
    yy, xx = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    alpha = 0.2 + 0.8 * torch.rand((H, W), generator=g)
    depth = 3.0 + 0.5 * torch.sin(xx / 7.0) + 0.4 * torch.cos(yy / 5.0) + 0.05 * torch.rand((H, W), generator=g)
    a = torch.zeros((8, H, W))
    a[0], a[1] = depth * alpha, alpha
    a[2:5] = torch.randn((3, H, W), generator=g) * alpha
    a[5] = depth + 0.1 * torch.rand((H, W), generator=g)
    a[6], a[7] = torch.rand((H, W), generator=g), (torch.rand((H, W), generator=g) > 0.5).float()
    a[1, 2, 3] = 1e-4                 # below the clamp
    a[1, 4, 4], a[0, 4, 4] = 0.0, 0.0
    a[2:5, 5, 5] = 0.0                # zero normal
    return a


@pytest.mark.parametrize("H,W,ratio", [(23, 31, 0.0), (16, 40, 1.0), (12, 12, 0.3)])
def test_surface_attributes_against_cpu_oracle(H, W, ratio):
    from b200gs.shade import _SurfaceAttributes, surface_camera_constants
    from oracle import shade_oracle as so
    wvt, fpt = _view(H, W, H)
    a = _allmap(H, W, H + W)
    ups = [torch.randn(s, generator=torch.Generator().manual_seed(k)) for k, s in enumerate(((H, W, 3), (1, H, W), (3, H, W)))]
    ao = a.double().requires_grad_(True)
    outs_o = so.surface_attributes(ao, wvt.double(), fpt.double(), ratio)
    sum((o * u.double()).sum() for o, u in zip(outs_o, ups)).backward()
    ad = a.cuda().requires_grad_(True)
    outs_n = _SurfaceAttributes.apply(ad, surface_camera_constants(wvt, fpt, W, H), ratio)
    sum((o * u.cuda()).sum() for o, u in zip(outs_n, ups)).backward()
    for name, n_, o_ in zip(("render_normal", "surf_depth", "surf_normal"), outs_n, outs_o):
        err = float((n_.detach().cpu().double() - o_.detach()).abs().max())
        assert err <= 2e-4 * max(1.0, float(o_.detach().abs().max())), (name, err)
    ref = ao.grad
    for ch in range(8):
        err = float((ad.grad[ch].cpu().double() - ref[ch]).abs().max())
        assert err <= 2e-3 * float(ref[ch].abs().max()) + 1e-6, (ch, err, float(ref[ch].abs().max()))


def test_surface_attributes_against_torch_lines_on_device():
    """The reference's lines (restated in oracle/shade_oracle.py) run as fp32 torch ops on the GPU at 800x800."""
    from b200gs.shade import _SurfaceAttributes, surface_camera_constants
    from oracle import shade_oracle as so
    H = W = 800
    wvt, fpt = _view(H, W, 1)
    a = _allmap(H, W, 9)
    ups = [torch.randn(s, generator=torch.Generator().manual_seed(k)).cuda() for k, s in enumerate(((H, W, 3), (1, H, W), (3, H, W)))]
    torch.set_default_device("cuda")
    try:
        ar = a.cuda().requires_grad_(True)
        outs_r = so.surface_attributes(ar, wvt.cuda(), fpt.cuda(), 0.5)
        sum((o * u).sum() for o, u in zip(outs_r, ups)).backward()
    finally:
        torch.set_default_device("cpu")
    ad = a.cuda().requires_grad_(True)
    outs_n = _SurfaceAttributes.apply(ad, surface_camera_constants(wvt, fpt, W, H), 0.5)
    sum((o * u).sum() for o, u in zip(outs_n, ups)).backward()
    for name, n_, r_ in zip(("render_normal", "surf_depth", "surf_normal"), outs_n, outs_r):
        assert float((n_ - r_).abs().max()) <= 1e-3 * max(1.0, float(r_.abs().max())), name    # fp32 cross products of nearby points
    for ch in range(8):
        d = (ad.grad[ch] - ar.grad[ch]).abs()
        tol = 1e-2 * float(ar.grad[ch].abs().max()) + 1e-6
        assert float((d > tol).float().mean()) <= 1e-3, (ch, float(d.max()), tol)
