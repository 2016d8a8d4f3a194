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
