"""GPU harness: the fused reflection pass (csrc/cubemap.cu reflect_shade_*) next to what the reference runs for it --
the torch lines of gaussian_renderer/__init__.py:22-35, :201 and utils/general_utils.py:177-197 (restated, with the
`pixel_camera` grid cached as the reference caches it) driving the UNMODIFIED reference cubemap extension
(oracle/_ref/ref_cubemap_C) when it is built, else this repo's drop-in cubemap backend.

  python tests/run_shade.py [H W ...] > gpurun_out/shade.jsonl

Per size: ms per forward+backward with calls queued back to back (CUDA events), the kernels' own times from the
library's stage events, algorithmic bytes (forward 12+12+4 in, 24 out; backward 12+12+4+12+12 in, 12+4+12 out per
pixel, texture traffic not counted) against the measured HBM copy peak, and the parity gate of the same run."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import harness as hz  # noqa: E402
from run_update import batch_timed, hbm_peak  # noqa: E402
from test_shade_gpu import _Env, _camera, _inputs  # noqa: E402


def run(H, W, L=128):
    import b200gs
    import _cubemapencoder as new_cm
    from b200gs.shade import camera_constants, reflect_shade
    backend = hz.ref_module("ref_cubemap_C")
    backend_name = "reference extension (oracle/_ref/ref_cubemap_C)"
    if backend is None:
        backend, backend_name = new_cm, "this repo's _cubemapencoder"
    HWK, R, T = _camera(H, W, 5)
    n, base, r, tex, fail = _inputs(H, W, L, 6)
    up = torch.randn((3, H, W), generator=torch.Generator().manual_seed(7)).cuda()
    out = {"H": H, "W": W, "L": L, "reference_lookup_backend": backend_name}

    class Lookup(torch.autograd.Function):          # cubemapencoder/cubemap_encoder.py:19-64
        @staticmethod
        def forward(ctx, inputs, emb, failv):
            inputs, emb = inputs.contiguous(), emb.contiguous()
            B = inputs.shape[0]
            o = torch.empty((3, B), device="cuda")
            backend.cubemap_encode_forward(inputs, emb, failv, o, 1, 1, B, 3, L)
            ctx.save_for_backward(inputs, emb)
            return o.permute(1, 0)

        @staticmethod
        def backward(ctx, g):
            inputs, emb = ctx.saved_tensors
            go = g.permute(1, 0).contiguous()
            ge, gi, gf = torch.zeros_like(emb), torch.zeros_like(inputs), torch.zeros(3, device="cuda")
            backend.cubemap_encode_backward(go, inputs, emb, ge, gi, gf, 1, 1, inputs.shape[0], 3, L)
            return gi, ge, gf

    lr = [t.cuda().requires_grad_(True) for t in (n, base, r, tex, fail)]
    Rs, Ts = torch.from_numpy(R).cuda(), torch.from_numpy(T).cuda()
    i, j = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32), indexing="xy")
    pixel_camera = torch.tensor(np.dot(np.stack([i, j, np.ones_like(i)], axis=2), np.linalg.inv(HWK[2]).T)).cuda()   # cached global

    def ref_path():
        for t in lr:
            t.grad = None
        Rt = Rs.T
        rays_o = (-Rt.T @ Ts.unsqueeze(-1)).flatten()
        rays_d = (pixel_camera - Ts[None, None]).reshape(-1, 3) @ Rt - rays_o[None]
        rays_d = (rays_d / torch.norm(rays_d, dim=1, keepdim=True)).reshape(H, W, 3)
        refl = rays_d - 2 * lr[0] * torch.sum(rays_d * lr[0], dim=-1, keepdim=True)
        refl_color = torch.sigmoid(Lookup.apply(refl.reshape(-1, 3), lr[3], lr[4])).reshape(H, W, 3).permute(2, 0, 1)
        final = (1 - lr[2]) * lr[1] + lr[2] * refl_color
        (final * up).sum().backward()
        return final

    ln = [t.cuda().requires_grad_(True) for t in (n, base, r, tex, fail)]
    env, cam = _Env(ln[3], ln[4]), camera_constants(HWK, R, T)

    def new_path():
        for t in ln:
            t.grad = None
        final, _ = reflect_shade(env, HWK, R, T, ln[0], ln[1], ln[2], cam=cam)
        (final * up).sum().backward()
        return final

    f_ref, f_new = ref_path(), new_path()
    out["final_max_abs"] = float((f_ref - f_new).abs().max())
    gr = {}
    for name, a, b in zip(("normal", "base", "refl", "texture", "failv"), ln, lr):
        d = (a.grad - b.grad).abs()
        tol = 1e-3 * float(b.grad.abs().max()) + 1e-6
        gr[name] = {"max_abs": float(d.max()), "ref_max": float(b.grad.abs().max()), "frac_over_tol": float((d > tol).float().mean())}
    out["grads"] = gr
    out["ref_ms"] = batch_timed(ref_path)
    out["new_ms"] = batch_timed(new_path)
    out["speedup"] = out["ref_ms"] / out["new_ms"]
    b200gs.timing_enable(1)
    b200gs.timing_read()
    for _ in range(10):
        new_path()
    torch.cuda.synchronize()
    st = b200gs.timing_read()
    b200gs.timing_enable(0)
    out["fwd_kernel_ms"], out["bwd_kernel_ms"] = st["shade_fwd"][0] / st["shade_fwd"][1], st["shade_bwd"][0] / st["shade_bwd"][1]
    N = H * W
    peak, _ = hbm_peak()
    out["fwd_gbps"], out["bwd_gbps"] = 52 * N / out["fwd_kernel_ms"] / 1e6, 80 * N / out["bwd_kernel_ms"] / 1e6
    out["fwd_hbm_frac"], out["bwd_hbm_frac"], out["hbm_peak_gbps"] = out["fwd_gbps"] / peak, out["bwd_gbps"] / peak, peak
    return out


def run_surface(H, W):
    """Surface attributes (gaussian_renderer/__init__.py:143-183): fused kernels vs the torch lines on the device."""
    import b200gs
    from b200gs.shade import _SurfaceAttributes, surface_camera_constants
    from oracle import shade_oracle as so
    from test_shade_gpu import _allmap, _view
    wvt, fpt = _view(H, W, 1)
    a = _allmap(H, W, 9)
    ups = [torch.randn(s, generator=torch.Generator().manual_seed(k)).cuda() for k, s in enumerate(((H, W, 3), (1, H, W), (3, H, W)))]
    ar, ad = a.cuda().requires_grad_(True), a.cuda().requires_grad_(True)
    wv, fp = wvt.cuda(), fpt.cuda()
    cam = surface_camera_constants(wvt, fpt, W, H)

    def ref_path():
        ar.grad = None
        torch.set_default_device("cuda")
        try:
            outs = so.surface_attributes(ar, wv, fp, 0.5)
        finally:
            torch.set_default_device("cpu")
        sum((o * u).sum() for o, u in zip(outs, ups)).backward()

    def new_path():
        ad.grad = None
        outs = _SurfaceAttributes.apply(ad, cam, 0.5)
        sum((o * u).sum() for o, u in zip(outs, ups)).backward()

    ref_path(); new_path()
    out = {"what": "surface_attributes", "H": H, "W": W}
    d = (ad.grad - ar.grad).abs()
    out["grad_frac_over_1e-2"] = float((d > 1e-2 * ar.grad.abs().max() + 1e-6).float().mean())
    out["ref_ms"], out["new_ms"] = batch_timed(ref_path), batch_timed(new_path)
    out["speedup"] = out["ref_ms"] / out["new_ms"]
    b200gs.timing_enable(1)
    b200gs.timing_read()
    for _ in range(10):
        new_path()
    torch.cuda.synchronize()
    st = b200gs.timing_read()
    b200gs.timing_enable(0)
    out["fwd_kernel_ms"], out["bwd_kernel_ms"] = st["surface_fwd"][0] / st["surface_fwd"][1], st["surface_bwd"][0] / st["surface_bwd"][1]
    N = H * W
    peak, _ = hbm_peak()
    # forward: 6 planes in, 3 + 1 + 3 floats out (+ surf_depth re-read); backward: 6 planes + depth + 7 upstream in, 8 planes out
    out["fwd_gbps"], out["bwd_gbps"] = 56 * N / out["fwd_kernel_ms"] / 1e6, 88 * N / out["bwd_kernel_ms"] / 1e6
    out["fwd_hbm_frac"], out["bwd_hbm_frac"] = out["fwd_gbps"] / peak, out["bwd_gbps"] / peak
    return out


if __name__ == "__main__":
    a = [int(v) for v in sys.argv[1:]]
    for H, W in (list(zip(a[::2], a[1::2])) or [(1080, 1920), (800, 800)]):
        print(json.dumps(run(H, W)), flush=True)
        torch.cuda.empty_cache()
        print(json.dumps(run_surface(H, W)), flush=True)
        torch.cuda.empty_cache()
