"""CPU restatement of the reflection pass around the rasterizer -- the ORACLE for the reflect_shade kernels of
csrc/cubemap.cu ("next" row 2, first half).

TEST INFRASTRUCTURE ONLY.  Nothing in the product package imports this file; only tests/ may.

What is restated, line by line in torch (CPU, float64 by default):
  sample_camera_rays()     utils/general_utils.py:177-197  (the R = R.T convention, rays_o, pixel_world, normalisation)
  reflection()             gaussian_renderer/__init__.py:22-24
  sample_cubemap_color()   gaussian_renderer/__init__.py:26-30  (sigmoid of the cubemap lookup)
  the final blend          gaussian_renderer/__init__.py:201
The cubemap lookup and its gradients are oracle/cubemap_oracle.py (pinned against the reference extension's
golden vectors, tests/golden/cubemap_ref.npz), wrapped in an autograd Function; everything else is
differentiated by autograd.

PARITY UNPINNED for the glue lines themselves: the reference functions above call .cuda() and import the CUDA
rasterizer, so they cannot run in the build container, and /root/reference does not travel to the GPU box;
the reference ships no test or golden vector for them.  The GPU tests therefore check the
kernel against this restatement and, on the device, against the same torch lines driving the UNMODIFIED
reference cubemap extension (oracle/_ref/ref_cubemap_C).
"""
import numpy as np
import torch

from oracle import cubemap_oracle as co


class _Cubemap(torch.autograd.Function):
    @staticmethod
    def forward(ctx, inputs, texture, failv, interp, seamless):
        out = co.forward(inputs.detach().numpy(), texture.detach().numpy(), failv.detach().numpy(), interp, seamless)   # [C,B]
        ctx.save_for_backward(inputs, texture)
        ctx.mode = (interp, seamless)
        return torch.from_numpy(np.ascontiguousarray(out.T))                                                              # [B,C] like env_map(...)

    @staticmethod
    def backward(ctx, g):
        inputs, texture = ctx.saved_tensors
        gin, gtex, gfail = co.backward(np.ascontiguousarray(g.numpy().T), inputs.detach().numpy(), texture.detach().numpy(), *ctx.mode)
        return (torch.from_numpy(np.asarray(gin, dtype=inputs.numpy().dtype)), torch.from_numpy(np.asarray(gtex, dtype=texture.numpy().dtype)),
                torch.from_numpy(np.asarray(gfail, dtype=texture.numpy().dtype)), None, None)


def sample_camera_rays(HWK, R, T, dtype=torch.float64):
    H, W, K = HWK
    R = torch.as_tensor(np.asarray(R), dtype=dtype).T                    # "the R rot matrix is transposed save in 3DGS"
    T = torch.as_tensor(np.asarray(T), dtype=dtype)
    i, j = np.meshgrid(np.arange(W, dtype=np.float32), np.arange(H, dtype=np.float32), indexing="xy")
    xy1 = np.stack([i, j, np.ones_like(i)], axis=2)
    pixel_camera = torch.as_tensor(np.dot(xy1, np.linalg.inv(np.asarray(K, np.float32)).T), dtype=dtype)
    rays_o = (-R.T @ T.unsqueeze(-1)).flatten()
    pixel_world = (pixel_camera - T[None, None]).reshape(-1, 3) @ R
    rays_d = pixel_world - rays_o[None]
    rays_d = rays_d / torch.norm(rays_d, dim=1, keepdim=True)
    return rays_d.reshape(H, W, 3)


def reflect_shade(texture, failv, HWK, R, T, normal_map, base_color, refl_strength_map, interp=1, seamless=1):
    """torch tensors (same dtype) -> (final_image [3,H,W], refl_color [3,H,W]); differentiable in normal_map, base_color,
    refl_strength_map, texture, failv."""
    H, W = normal_map.shape[:2]
    rays_d = sample_camera_rays(HWK, R, T, normal_map.dtype)
    refl = rays_d - 2 * normal_map * torch.sum(rays_d * normal_map, dim=-1, keepdim=True)
    outcolor = torch.sigmoid(_Cubemap.apply(refl.reshape(-1, 3), texture, failv, interp, seamless))
    refl_color = outcolor.reshape(H, W, 3).permute(2, 0, 1)
    final_image = (1 - refl_strength_map) * base_color + refl_strength_map * refl_color
    return final_image, refl_color


# ---------------------------------------------------------------------------------------------------------------
# Surface attributes: gaussian_renderer/__init__.py:143-183 with utils/point_utils.py:9-37, restated line by line.
# The reference module imports cv2 / matplotlib / the CUDA rasterizer and calls .cuda(), so it cannot be imported
# here either: PARITY UNPINNED for these lines too (see the header).
def depths_to_points(world_view_transform, full_proj_transform, W, H, depthmap):
    dt = depthmap.dtype
    wvt, fpt = world_view_transform.to(dt), full_proj_transform.to(dt)
    c2w = (wvt.T).inverse()
    ndc2pix = torch.tensor([[W / 2, 0, 0, (W) / 2], [0, H / 2, 0, (H) / 2], [0, 0, 0, 1]], dtype=dt).T
    projection_matrix = c2w.T @ fpt
    intrins = (projection_matrix @ ndc2pix)[:3, :3].T
    grid_x, grid_y = torch.meshgrid(torch.arange(W, dtype=dt), torch.arange(H, dtype=dt), indexing="xy")
    points = torch.stack([grid_x, grid_y, torch.ones_like(grid_x)], dim=-1).reshape(-1, 3)
    rays_d = points @ intrins.inverse().T @ c2w[:3, :3].T
    rays_o = c2w[:3, 3]
    return depthmap.reshape(-1, 1) * rays_d + rays_o


def depth_to_normal(world_view_transform, full_proj_transform, W, H, depth):
    points = depths_to_points(world_view_transform, full_proj_transform, W, H, depth).reshape(*depth.shape[1:], 3)
    output = torch.zeros_like(points)
    dx = torch.cat([points[2:, 1:-1] - points[:-2, 1:-1]], dim=0)
    dy = torch.cat([points[1:-1, 2:] - points[1:-1, :-2]], dim=1)
    normal_map = torch.nn.functional.normalize(torch.cross(dx, dy, dim=-1), dim=-1)
    output[1:-1, 1:-1, :] = normal_map
    return output


def surface_attributes(allmap, world_view_transform, full_proj_transform, depth_ratio):
    """allmap [8,H,W] (torch, CPU) -> (render_normal [H,W,3], surf_depth [1,H,W], surf_normal [3,H,W])."""
    H, W = allmap.shape[1:]
    render_alpha = allmap[1:2]
    render_normal = allmap[2:5]
    render_normal = (render_normal.permute(1, 2, 0) @ (world_view_transform.to(allmap.dtype)[:3, :3].T)).permute(2, 0, 1)
    render_depth_median = torch.nan_to_num(allmap[5:6], 0, 0)
    render_depth_expected = allmap[0:1] / torch.clamp(render_alpha, min=1e-3)
    render_depth_expected = torch.nan_to_num(render_depth_expected, 0, 0)
    surf_depth = render_depth_expected * (1 - depth_ratio) + depth_ratio * render_depth_median
    surf_normal = depth_to_normal(world_view_transform, full_proj_transform, W, H, surf_depth).permute(2, 0, 1)
    surf_normal = surf_normal * render_alpha.detach()
    render_normal = render_normal.permute(1, 2, 0)
    render_normal = render_normal / (torch.norm(render_normal, dim=-1, keepdim=True) + 1e-6)
    return render_normal, surf_depth, surf_normal
