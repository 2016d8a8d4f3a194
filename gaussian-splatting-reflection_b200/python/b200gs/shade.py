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


# ---------------------------------------------------------------------------------------------------------------
# Surface attributes (second half of row 2): gaussian_renderer/__init__.py:143-183
class _SurfaceAttributes(torch.autograd.Function):
    @staticmethod
    def forward(ctx, allmap, cam, depth_ratio):
        if not (allmap.is_cuda and allmap.dtype == torch.float32 and allmap.dim() == 3 and allmap.shape[0] == 8):
            raise RuntimeError("allmap must be an fp32 CUDA tensor [8, H, W] (there is no CPU path)")
        allmap = allmap.contiguous()
        H, W = allmap.shape[1], allmap.shape[2]
        normal = torch.empty((H, W, 3), dtype=torch.float32, device=allmap.device)
        surf_depth = torch.empty((1, H, W), dtype=torch.float32, device=allmap.device)
        surf_normal = torch.empty((3, H, W), dtype=torch.float32, device=allmap.device)
        (_, mp), (_, ap), (_, op) = cam
        _b.check(_b.surface_forward(H, W, _b.ptr(allmap), mp, ap, op, float(depth_ratio), _b.ptr(normal), _b.ptr(surf_depth),
                                    _b.ptr(surf_normal), torch.cuda.current_stream().cuda_stream))
        ctx.save_for_backward(allmap, surf_depth)
        ctx.cam, ctx.ratio = cam, float(depth_ratio)
        return normal, surf_depth, surf_normal

    @staticmethod
    def backward(ctx, g_normal, g_depth, g_sn):
        allmap, surf_depth = ctx.saved_tensors
        H, W = allmap.shape[1], allmap.shape[2]
        g_all = torch.empty_like(allmap)
        (_, mp), (_, ap), (_, op) = ctx.cam
        cont = lambda t: None if t is None else t.contiguous()   # noqa: E731
        g_normal, g_depth, g_sn = cont(g_normal), cont(g_depth), cont(g_sn)
        _b.check(_b.surface_backward(H, W, _b.ptr(allmap), mp, ap, op, ctx.ratio, _b.ptr(surf_depth), _b.ptr(g_normal),
                                     _b.ptr(g_depth), _b.ptr(g_sn), _b.ptr(g_all), torch.cuda.current_stream().cuda_stream))
        return g_all, None, None


def surface_camera_constants(world_view_transform, full_proj_transform, W, H):
    """Host constants of depths_to_points (utils/point_utils.py:9-24) for one camera: the stored view rotation,
    the ray matrix A (rays_d = (x, y, 1) @ A) and rays_o.  Keep the result; it owns the memory."""
    wvt = np.asarray(world_view_transform.detach().cpu() if isinstance(world_view_transform, torch.Tensor) else world_view_transform, np.float64)
    fpt = np.asarray(full_proj_transform.detach().cpu() if isinstance(full_proj_transform, torch.Tensor) else full_proj_transform, np.float64)
    c2w = np.linalg.inv(wvt.T)
    ndc2pix = np.array([[W / 2, 0, 0, W / 2], [0, H / 2, 0, H / 2], [0, 0, 0, 1]], np.float64).T
    intrins = ((c2w.T @ fpt) @ ndc2pix)[:3, :3].T
    A = np.linalg.inv(intrins).T @ c2w[:3, :3].T
    return _host9(wvt[:3, :3], 9), _host9(A, 9), _host9(c2w[:3, 3], 3)


def surface_attributes(allmap, viewpoint_camera, depth_ratio, cam=None):
    """-> dict with the renderer's names: rend_alpha [1,H,W], rend_normal [3,H,W] (normalised, world space; the [H,W,3]
    layout the reflection pass wants is `render_normal_hw3`), rend_dist, surf_depth, surf_normal, env_scope_mask."""
    if cam is None:
        cam = surface_camera_constants(viewpoint_camera.world_view_transform, viewpoint_camera.full_proj_transform,
                                       int(viewpoint_camera.image_width), int(viewpoint_camera.image_height))
    normal, surf_depth, surf_normal = _SurfaceAttributes.apply(allmap, cam, depth_ratio)
    return {"rend_alpha": allmap[1:2], "render_normal_hw3": normal, "rend_normal": normal.permute(2, 0, 1), "rend_dist": allmap[6:7],
            "surf_depth": surf_depth, "surf_normal": surf_normal, "env_scope_mask": allmap[7:8]}
