"""Reflection shading around the rasterizer as one kernel each way ("next" row 2,
first half; csrc/cubemap.cu reflect_shade_*): what gaussian_renderer/__init__.py does at :22-35 and :199-201

    refl_color  = get_refl_color(pc.get_envmap, camera.HWK, camera.R, camera.T, render_normal)
    final_image = (1 - refl_strength_map) * base_color + refl_strength_map * refl_color

with ~12 full-image torch kernels plus the cubemap extension.  Same argument meaning: `env_map` is the
CubemapEncoder module (drop-in `cubemapencoder` package or anything exposing `.params['Cubemap_texture']`,
`.params['Cubemap_failv']`, `.interp_id`, `.seamless`), `HWK = (H, W, K)` with K the 3x3 intrinsics,
`R`, `T` the camera's stored rotation / translation, `normal_map` [H, W, 3] (already normalised, as the
renderer passes it).  No CPU path.
"""
import ctypes as C

import numpy as np
import torch

import b200gs as _b


def _host9(a, n):
    a = np.ascontiguousarray(np.asarray(a.detach().cpu() if isinstance(a, torch.Tensor) else a, dtype=np.float32).reshape(-1))
    if a.size != n:
        raise ValueError(f"expected {n} floats, got {a.size}")
    return a, a.ctypes.data_as(C.POINTER(C.c_float))


class _ReflectShade(torch.autograd.Function):
    @staticmethod
    def forward(ctx, normal_map, base_color, refl_strength_map, texture, failv, cam, interp, seamless):
        for name, t in (("normal_map", normal_map), ("base_color", base_color), ("refl_strength_map", refl_strength_map),
                        ("Cubemap_texture", texture), ("Cubemap_failv", failv)):
            if not (t.is_cuda and t.dtype == torch.float32):
                raise RuntimeError(f"{name} must be an fp32 CUDA tensor (there is no CPU path)")
        H, W = normal_map.shape[0], normal_map.shape[1]
        if tuple(normal_map.shape) != (H, W, 3) or tuple(base_color.shape) != (3, H, W) or refl_strength_map.numel() != H * W:
            raise RuntimeError("expected normal_map [H,W,3], base_color [3,H,W], refl_strength_map [1,H,W]")
        if texture.dim() != 4 or texture.shape[0] != 6 or texture.shape[1] != 3 or texture.shape[2] != texture.shape[3]:
            raise RuntimeError("the fused reflection pass needs a [6, 3, L, L] cubemap")
        normal_map, base_color, refl = normal_map.contiguous(), base_color.contiguous(), refl_strength_map.contiguous()
        texture, failv = texture.contiguous(), failv.contiguous()
        final, env = torch.empty_like(base_color), torch.empty_like(base_color)
        (kin, kp), (rot, rp), (tr, tp) = cam
        _b.check(_b.reflect_shade_forward(H, W, kp, rp, tp, _b.ptr(normal_map), _b.ptr(base_color), _b.ptr(refl), _b.ptr(texture),
                                          _b.ptr(failv), texture.shape[2], interp, seamless, _b.ptr(final), _b.ptr(env),
                                          torch.cuda.current_stream().cuda_stream))
        ctx.save_for_backward(normal_map, base_color, refl, texture, env)
        ctx.cam, ctx.interp, ctx.seamless, ctx.refl_shape = cam, interp, seamless, refl_strength_map.shape
        return final, env

    @staticmethod
    def backward(ctx, g_final, g_env):
        normal_map, base_color, refl, texture, env = ctx.saved_tensors
        H, W = normal_map.shape[0], normal_map.shape[1]
        g_final = g_final.contiguous()
        g_env = g_env.contiguous() if g_env is not None else None
        g_base, g_refl = torch.empty_like(base_color), torch.empty_like(refl)
        g_normal = torch.empty_like(normal_map)
        g_tex, g_fail = torch.zeros_like(texture), torch.zeros((3,), dtype=torch.float32, device=texture.device)
        (kin, kp), (rot, rp), (tr, tp) = ctx.cam
        _b.check(_b.reflect_shade_backward(H, W, kp, rp, tp, _b.ptr(normal_map), _b.ptr(base_color), _b.ptr(refl), _b.ptr(texture),
                                           texture.shape[2], ctx.interp, ctx.seamless, _b.ptr(env), _b.ptr(g_final), _b.ptr(g_env),
                                           _b.ptr(g_base), _b.ptr(g_refl), _b.ptr(g_normal), _b.ptr(g_tex), _b.ptr(g_fail),
                                           torch.cuda.current_stream().cuda_stream))
        return g_normal, g_base, g_refl.reshape(ctx.refl_shape), g_tex, g_fail, None, None, None


def camera_constants(HWK, R, T):
    """Host-side constants of one camera: inv(K), R, T as float32 (keep the result; it owns the memory)."""
    K = np.asarray(HWK[2], dtype=np.float32)
    return _host9(np.linalg.inv(K), 9), _host9(R, 9), _host9(T, 3)


def reflect_shade(env_map, HWK, R, T, normal_map, base_color, refl_strength_map, cam=None):
    """-> (final_image [3,H,W], refl_color [3,H,W]) == ((1-r)*base + r*get_refl_color(...), get_refl_color(...))."""
    params = env_map.params
    cam = cam if cam is not None else camera_constants(HWK, R, T)
    return _ReflectShade.apply(normal_map, base_color, refl_strength_map, params["Cubemap_texture"], params["Cubemap_failv"], cam,
                               int(getattr(env_map, "interp_id", 1)), int(getattr(env_map, "seamless", 1)))
