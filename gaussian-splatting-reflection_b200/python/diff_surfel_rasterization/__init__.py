"""Drop-in `diff_surfel_rasterization` package over libb200gs.so (sm_100a).

Public surface and behaviour follow the reference package
(S/diff_surfel_rasterization/__init__.py): `GaussianRasterizationSettings`
(:170-182), `GaussianRasterizer` with `.markVisible` and `.forward` (:184-239),
`rasterize_gaussians` (:21-46) and the autograd function behind it (:48-168), so
the reference's gaussian_renderer/__init__.py, train.py and render.py import and
call it unchanged.

Two keyword-only extras exist for callers that know this library (the reference's callers never pass
them): `exact_lists` (emit the reference's instance lists bit for bit, include/b200gs.h) and `grad_arena`
(a b200gs.dp.GradArena the backward accumulates the parameter gradients of a batch of views into, for
the one all-reduce of view-parallel training).  Both are per-call state of the module instance;
the library itself keeps no mode.  `GaussianRasterizer.prepare(...)` (optional) runs stage 1 of the module's next
forward() ahead of time, e.g. on a second stream while the previous view's compositing kernels run.
"""
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


def _snapshot(args):
    """CPU copy of the call arguments, taken before a debug-mode call so a failing
    call can be dumped (reference: cpu_deep_copy_tuple, :17-19)."""
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


def _call_with_dump(fn, args, debug, dump_path, what, **kw):
    if not debug:
        return fn(*args, **kw)
    saved = _snapshot(args)
    try:
        return fn(*args, **kw)
    except Exception:
        torch.save(saved, dump_path)
        print(f"\nAn error occured in {what}. Please forward {dump_path} for debugging.")
        raise


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, refl_strengths, opacities, scales, rotations,
                cov3Ds_precomp, raster_settings, env_scope_mask, opts=None):
        rs = raster_settings
        opts = opts or {}
        args = (rs.bg, means3D, env_scope_mask, colors_precomp, refl_strengths, opacities, scales, rotations,
                rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        prepared = opts.get("prepared")
        if prepared is not None:      # stage 1 already ran (GaussianRasterizer.prepare), possibly on another stream
            (num_rendered, color, depth, radii, geomBuffer, binningBuffer, imgBuffer, refl_strength_map,
             gaussian_weights) = _call_with_dump(_C.rasterize_gaussians_finish, (prepared,), rs.debug, "snapshot_fw.dump",
                                                 "forward")
        else:
            (num_rendered, color, depth, radii, geomBuffer, binningBuffer, imgBuffer, refl_strength_map,
             gaussian_weights) = _call_with_dump(_C.rasterize_gaussians, args, rs.debug, "snapshot_fw.dump", "forward",
                                                 exact_lists=bool(opts.get("exact_lists", False)),
                                                 depth_slices=int(opts.get("depth_slices", 0)))
        ctx.raster_settings = rs
        ctx.grad_arena = opts.get("grad_arena")
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, refl_strengths, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geomBuffer, binningBuffer, imgBuffer)
        return color, radii, depth, refl_strength_map, gaussian_weights

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii, grad_depth, grad_out_strength_map, _grad_weights):
        rs = ctx.raster_settings
        (colors_precomp, refl_strengths, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
         binningBuffer, imgBuffer) = ctx.saved_tensors
        args = (rs.bg, means3D, radii, colors_precomp, refl_strengths, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, grad_depth,
                grad_out_strength_map, sh, rs.sh_degree, rs.campos, geomBuffer, ctx.num_rendered, binningBuffer,
                imgBuffer, rs.debug)
        (grad_means2D, grad_colors_precomp, grad_refl, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh,
         grad_scales, grad_rotations) = _call_with_dump(_C.rasterize_gaussians_backward, args, rs.debug,
                                                        "snapshot_bw.dump", "backward", grad_arena=ctx.grad_arena)
        if ctx.grad_arena is not None:
            # The parameter gradients of this view now sit (summed with the batch's other views) in the arena.
            # Handing arena views to autograd would alias them with .grad / be added a second time, and would
            # run the activation backward once per view; the arena is propagated ONCE per batch instead
            # (b200gs.dp.GradArena.backward_through / bind_grads).  means2D (the densification statistic) and the
            # precomputed-input gradients stay ordinary per-view tensors.
            return (None, grad_means2D, None, grad_colors_precomp, None, None, None, None, grad_cov3Ds_precomp,
                    None, None, None)
        # input order of forward(); None for raster_settings, env_scope_mask and opts
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_refl, grad_opacities, grad_scales,
                grad_rotations, grad_cov3Ds_precomp, None, None, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, refl_strengths, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings, env_scope_mask, opts=None):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, refl_strengths, opacities, scales,
                                     rotations, cov3Ds_precomp, raster_settings, env_scope_mask, opts)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings, *, exact_lists=False, grad_arena=None, depth_slices=0):
        super().__init__()
        self.raster_settings = raster_settings
        self.exact_lists = bool(exact_lists)
        self.grad_arena = grad_arena
        self.depth_slices = int(depth_slices)
        self._prepared = None

    def markVisible(self, positions):
        """Boolean mask of the points in front of the camera's near plane."""
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    @staticmethod
    def _normalise(means3D, shs, colors_precomp, scales, rotations, cov3D_precomp, env_scope_mask):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        have_sr = scales is not None or rotations is not None
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (have_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        def empty():
            return torch.empty((0,), dtype=torch.float32, device=means3D.device)

        return tuple(empty() if t is None else t for t in (shs, colors_precomp, scales, rotations, cov3D_precomp,
                                                           env_scope_mask))

    def prepare(self, means3D, means2D, opacities, shs=None, colors_precomp=None, refl_strengths=None, scales=None,
                rotations=None, cov3D_precomp=None, env_scope_mask=None):
        """Optional, no reference counterpart: run stage 1 of the NEXT forward() of this module (preprocess, depth
        order, prefix sum, slice plan and the forward's one host sync) now, on the CURRENT stream -- typically a
        second stream, while the compositing kernels of the previous view still occupy the SMs.  forward() must
        then be called with the same tensors; it only runs stage 2.  The parameters must not change in between."""
        shs, colors_precomp, scales, rotations, cov3D_precomp, env_scope_mask = self._normalise(
            means3D, shs, colors_precomp, scales, rotations, cov3D_precomp, env_scope_mask)
        rs = self.raster_settings
        with torch.no_grad():
            prep = _C.rasterize_gaussians_prepare(
                rs.bg, means3D, env_scope_mask, colors_precomp, refl_strengths, opacities, scales, rotations,
                rs.scale_modifier, cov3D_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                rs.image_width, shs, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug, exact_lists=self.exact_lists,
                depth_slices=self.depth_slices)
        self._prepared = (prep, tuple(t.data_ptr() for t in (means3D, opacities, shs, colors_precomp, scales, rotations,
                                                             cov3D_precomp)))
        return self

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, refl_strengths=None, scales=None,
                rotations=None, cov3D_precomp=None, env_scope_mask=None):
        shs, colors_precomp, scales, rotations, cov3D_precomp, env_scope_mask = self._normalise(
            means3D, shs, colors_precomp, scales, rotations, cov3D_precomp, env_scope_mask)
        opts = None
        prepared, self._prepared = getattr(self, "_prepared", None), None
        if prepared is not None:
            prep, ptrs = prepared
            if ptrs != tuple(t.data_ptr() for t in (means3D, opacities, shs, colors_precomp, scales, rotations, cov3D_precomp)):
                raise RuntimeError("forward() was given other tensors than the prepare() call before it")
        if self.exact_lists or self.grad_arena is not None or self.depth_slices or prepared is not None:
            opts = {"exact_lists": self.exact_lists, "grad_arena": self.grad_arena, "depth_slices": self.depth_slices,
                    "prepared": prepared[0] if prepared is not None else None}
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, refl_strengths, opacities, scales, rotations,
                                   cov3D_precomp, self.raster_settings, env_scope_mask, opts)
