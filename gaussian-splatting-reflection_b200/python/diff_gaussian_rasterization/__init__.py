"""Drop-in `diff_gaussian_rasterization` package over libb200gs.so (sm_100a).

Public surface follows the reference package (G/diff_gaussian_rasterization/__init__.py):
`GaussianRasterizationSettings` with the 13th field `antialiasing` (:157-170),
`GaussianRasterizer.forward(means3D, means2D, opacities, shs, colors_precomp, normals,
refl_strengths, scales, rotations, cov3D_precomp)` -> (color, radii, invdepths, normal_map,
refl_strength_map) (:188-232), `rasterize_gaussians` and the autograd function (:48-155)."""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    antialiasing: bool


def _call_with_dump(fn, args, debug, dump_path, what):
    """debug=True: keep a CPU copy of the arguments and dump it if the call raises (reference :84-93, :141-148)."""
    if not debug:
        return fn(*args)
    saved = tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_path)
        print(f"\nAn error occured in {what}. Please forward {dump_path} for debugging.")
        raise


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, normals, refl_strengths, opacities, scales, rotations,
                cov3Ds_precomp, raster_settings):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, normals, refl_strengths, opacities, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width,
                sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.antialiasing, rs.debug)
        (num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, invdepths, normal_map,
         refl_strength_map) = _call_with_dump(_C.rasterize_gaussians, args, rs.debug, "snapshot_fw.dump", "forward")
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, normals, refl_strengths, means3D, scales, rotations, cov3Ds_precomp, radii,
                              sh, opacities, geomBuffer, binningBuffer, imgBuffer)
        return color, radii, invdepths, normal_map, refl_strength_map

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii, grad_out_depth, grad_out_normal_map, grad_out_strength_map):
        rs = ctx.raster_settings
        (colors_precomp, normals, refl_strengths, means3D, scales, rotations, cov3Ds_precomp, radii, sh, opacities,
         geomBuffer, binningBuffer, imgBuffer) = ctx.saved_tensors
        args = (rs.bg, means3D, radii, colors_precomp, normals, refl_strengths, opacities, scales, rotations,
                rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color,
                grad_out_depth, grad_out_normal_map, grad_out_strength_map, sh, rs.sh_degree, rs.campos, geomBuffer,
                ctx.num_rendered, binningBuffer, imgBuffer, rs.antialiasing, rs.debug)
        (grad_means2D, grad_colors_precomp, grad_normals, grad_refl, grad_opacities, grad_means3D, grad_cov3Ds_precomp,
         grad_sh, grad_scales, grad_rotations) = _call_with_dump(_C.rasterize_gaussians_backward, args, rs.debug,
                                                                 "snapshot_bw.dump", "backward")
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_normals, grad_refl, grad_opacities,
                grad_scales, grad_rotations, grad_cov3Ds_precomp, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, normals, refl_strengths, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, normals, refl_strengths, opacities, scales,
                                     rotations, cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, normals=None, refl_strengths=None,
                scales=None, rotations=None, cov3D_precomp=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        have_sr = scales is not None or rotations is not None
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (have_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        def empty():
            return torch.empty((0,), dtype=torch.float32, device=means3D.device)

        shs = empty() if shs is None else shs
        colors_precomp = empty() if colors_precomp is None else colors_precomp
        scales = empty() if scales is None else scales
        rotations = empty() if rotations is None else rotations
        cov3D_precomp = empty() if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, normals, refl_strengths, opacities, scales,
                                   rotations, cov3D_precomp, self.raster_settings)
