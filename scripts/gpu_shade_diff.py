"""GPU diagnostic: fused reflection pass (b200gs.shade.reflect_shade) against the reference's torch lines on the
reference cubemap extension -- relative L2 and max differences of the image and of every gradient, 800x800, L = 128."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
import torch
import harness as hz  # noqa: E402
from test_shade_gpu import _Env, _camera, _inputs

refC = hz.ref_module("ref_cubemap_C")
from b200gs.shade import reflect_shade
H = W = 800
L = 128
HWK, R, T = _camera(H, W, 5)
n, base, r, tex, fail = _inputs(H, W, L, 6)
up = torch.randn((3, H, W), generator=torch.Generator().manual_seed(7)).cuda()


class RefLookup(torch.autograd.Function):
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


def ref_run(perturb=False):
    nn = n * (1.0 + 1.2e-7) if perturb else n       # every component moved by about one unit in the last place
    lr = [t.cuda().requires_grad_(True) for t in (nn, base, r, tex, fail)]
    Rt, Tt = torch.from_numpy(R).cuda().T, torch.from_numpy(T).cuda()
    i, j = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32), indexing="xy")
    pixel_camera = torch.tensor(np.dot(np.stack([i, j, np.ones_like(i)], axis=2), np.linalg.inv(HWK[2]).T)).cuda()
    rays_o = (-Rt.T @ Tt.unsqueeze(-1)).flatten()
    rays_d = (pixel_camera - Tt[None, None]).reshape(-1, 3) @ Rt - rays_o[None]
    rays_d = (rays_d / torch.norm(rays_d, dim=1, keepdim=True)).reshape(H, W, 3)
    refl = rays_d - 2 * lr[0] * torch.sum(rays_d * lr[0], dim=-1, keepdim=True)
    refl.retain_grad()
    lookup = RefLookup.apply(refl.reshape(-1, 3), lr[3], lr[4])
    refl_color = torch.sigmoid(lookup).reshape(H, W, 3).permute(2, 0, 1)
    final_ref = (1 - lr[2]) * lr[1] + lr[2] * refl_color
    (final_ref * up).sum().backward()
    return final_ref.detach(), refl_color.detach(), [t.grad for t in lr], refl.detach(), refl.grad


f1, e1, g1, d1, dg1 = ref_run()
f1b, e1b, g1b, _, _ = ref_run()
f1p, e1p, g1p, _, _ = ref_run(perturb=True)
print("REFERENCE lines against THEMSELVES with the normals moved by one ulp:")
for name, b, bp in zip(("normal", "base", "refl", "texture"), g1, g1p):
    print(f"  {name:8s} grad rel l2 {hz.rel_l2(bp, b):.3g}  fraction of entries > 1e-4 rel {1 - hz.grad_close_fraction(bp, b, rtol=1e-4):.3g}")
print("fused kernels against the reference lines:")
ln = [t.cuda().requires_grad_(True) for t in (n, base, r, tex, fail)]
fn, en = reflect_shade(_Env(ln[3], ln[4]), HWK, R, T, ln[0], ln[1], ln[2])
(fn * up).sum().backward()
print("final image   max abs", float((fn - f1).abs().max()), "rel l2", hz.rel_l2(fn, f1))
print("refl colour   max abs", float((en - e1).abs().max()), "rel l2", hz.rel_l2(en, e1))
for name, a, b, b2 in zip(("normal", "base", "refl", "texture"), ln, g1, g1b):
    print(f"{name:8s} grad rel l2 {hz.rel_l2(a.grad, b):.3g}  (reference run to run {hz.rel_l2(b2, b):.3g})  max abs diff / max entry "
          f"{float((a.grad - b).abs().max()) / float(b.abs().max()):.3g}  fraction of entries > 1e-4 rel "
          f"{1 - hz.grad_close_fraction(a.grad, b, rtol=1e-4):.3g}")
