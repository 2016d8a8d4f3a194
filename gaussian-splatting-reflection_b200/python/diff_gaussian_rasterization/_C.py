"""`diff_gaussian_rasterization._C` -- the reference's pybind entry points (G/ext.cpp:15-19,
G/rasterize_points.h:19-78) over the C ABI of libb200gs.so: same argument order, same result
tuples.  This shim does what G/rasterize_points.cu does around the core (checks, `.contiguous()`,
output / scratch allocation); all arithmetic is in the sm_100a kernels.  There is no `fusedssim`
symbol here either (the reference's utils/loss_utils.py:16-19 probes for one under a bare except)."""
import ctypes as C

import torch

import b200gs as _b

NUM_CHANNELS = 3


def _cuda_f32(name, t):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if t.numel() and t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")
    return t.contiguous()


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def rasterize_gaussians(background, means3D, colors, normals, refl_strengths, opacity, scales, rotations, scale_modifier,
                        cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh,
                        degree, campos, prefiltered, antialiasing, debug, *, exact_lists=False, depth_slices=0,
                        slice_fractions=None):
    """-> (rendered, out_color[3,H,W], radii[P], geomBuffer, binningBuffer, imgBuffer, out_invdepth[1,H,W],
    normal_map[3,H,W], refl_strength_map[1,H,W])   (G/rasterize_points.cu:39-141)

    Keyword-only, per call (no reference counterpart): exact_lists, depth_slices, slice_fractions -- the fields of the
    same names of b200gs_gauss_fwd (include/b200gs.h), as in diff_surfel_rasterization._C.rasterize_gaussians."""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    background = _cuda_f32("background", background)
    means3D = _cuda_f32("means3D", means3D)
    colors = _cuda_f32("colors", colors)
    normals = _cuda_f32("normals", normals)
    refl_strengths = _cuda_f32("refl_strengths", refl_strengths)
    opacity = _cuda_f32("opacity", opacity)
    scales = _cuda_f32("scales", scales)
    rotations = _cuda_f32("rotations", rotations)
    cov3D_precomp = _cuda_f32("cov3D_precomp", cov3D_precomp)
    viewmatrix = _cuda_f32("viewmatrix", viewmatrix)
    projmatrix = _cuda_f32("projmatrix", projmatrix)
    sh = _cuda_f32("sh", sh)
    campos = _cuda_f32("campos", campos)
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    dev = means3D.device
    f32 = dict(dtype=torch.float32, device=dev)
    u8 = dict(dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        # the reference zero-fills these (G/rasterize_points.cu:78-84); the kernels write every pixel and every radius
        # themselves, so only the P == 0 short-circuit needs zeros.  One allocation for the eight output planes.
        new = torch.zeros if P == 0 else torch.empty
        planes = new((NUM_CHANNELS + 3 + 1 + 1, H, W), **f32)
        out_color, out_normal = planes[:NUM_CHANNELS], planes[NUM_CHANNELS:NUM_CHANNELS + 3]
        out_invdepth, out_refl = planes[NUM_CHANNELS + 3:NUM_CHANNELS + 4], planes[NUM_CHANNELS + 4:]
        radii = new((P,), dtype=torch.int32, device=dev)
        geom, binning, img = torch.empty((0,), **u8), torch.empty((0,), **u8), torch.empty((0,), **u8)
        rendered = 0
        if P != 0:
            M = sh.size(1) if sh.numel() else 0
            geom = torch.empty((_b.gauss_geom_bytes(P),), **u8)
            img = torch.empty((_b.gauss_image_bytes(W, H),), **u8)
            a = _b.GaussFwd(
                P=P, sh_degree=int(degree), sh_coeffs=M, width=W, height=H, tan_fovx=float(tan_fovx),
                tan_fovy=float(tan_fovy), scale_modifier=float(scale_modifier), prefiltered=int(bool(prefiltered)),
                antialiasing=int(bool(antialiasing)), debug=int(bool(debug)),
                background=_b.ptr(background), means3D=_b.ptr(means3D), shs=_b.ptr(sh), colors_precomp=_b.ptr(colors),
                normals=_b.ptr(normals), refl_strengths=_b.ptr(refl_strengths), opacities=_b.ptr(opacity),
                scales=_b.ptr(scales), rotations=_b.ptr(rotations), cov3D_precomp=_b.ptr(cov3D_precomp),
                viewmatrix=_b.ptr(viewmatrix), projmatrix=_b.ptr(projmatrix), campos=_b.ptr(campos),
                geom=geom.data_ptr(), geom_bytes=geom.numel(), img=img.data_ptr(), img_bytes=img.numel(),
                binning=None, binning_bytes=0, out_color=out_color.data_ptr(), out_invdepth=out_invdepth.data_ptr(),
                out_normal=out_normal.data_ptr(), out_refl=out_refl.data_ptr(), radii=radii.data_ptr(),
                exact_lists=int(bool(exact_lists)), depth_slices=int(depth_slices))
            if slice_fractions:
                for i, fr in enumerate(list(slice_fractions)[:7]):
                    a.slice_frac[i] = float(fr)
            plan = (C.c_int * _b.PLAN_INTS)()
            st = _stream()
            _b.check(_b.gauss_forward_stage1(C.byref(a), plan, st))
            rendered = plan[0]
            binning = torch.empty((_b.binning_bytes(rendered),), **u8)
            a.binning = binning.data_ptr()
            a.binning_bytes = binning.numel()
            _b.check(_b.gauss_forward_stage2(C.byref(a), plan, st))
    return rendered, out_color, radii, geom, binning, img, out_invdepth, out_normal, out_refl


def rasterize_gaussians_backward(background, means3D, radii, colors, normals, refl_strengths, opacities, scales, rotations,
                                 scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color,
                                 dL_dout_invdepth, dL_dout_normal_map, dL_dout_refl_strength_map, sh, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, antialiasing, debug):
    """-> (dL_dmeans2D[P,3] ("pixels" variant), dL_dcolors[P,3], dL_dnormals[P,3], dL_drefl[P,1], dL_dopacity[P,1],
    dL_dmeans3D[P,3], dL_dcov3D[P,6], dL_dsh[P,M,3], dL_dscales[P,3], dL_drotations[P,4])  (G/rasterize_points.cu:144-264)"""
    background = _cuda_f32("background", background)
    means3D = _cuda_f32("means3D", means3D)
    colors = _cuda_f32("colors", colors)
    opacities = _cuda_f32("opacities", opacities)
    scales = _cuda_f32("scales", scales)
    rotations = _cuda_f32("rotations", rotations)
    cov3D_precomp = _cuda_f32("cov3D_precomp", cov3D_precomp)
    viewmatrix = _cuda_f32("viewmatrix", viewmatrix)
    projmatrix = _cuda_f32("projmatrix", projmatrix)
    sh = _cuda_f32("sh", sh)
    campos = _cuda_f32("campos", campos)
    dL_dout_color = _cuda_f32("dL_dout_color", dL_dout_color)
    dL_dout_normal_map = _cuda_f32("dL_dout_normal_map", dL_dout_normal_map)
    has_inv = dL_dout_invdepth is not None and dL_dout_invdepth.size(0) != 0
    has_refl = dL_dout_refl_strength_map is not None and dL_dout_refl_strength_map.size(0) != 0
    if has_inv:
        dL_dout_invdepth = _cuda_f32("dL_dout_invdepth", dL_dout_invdepth)
    if has_refl:
        dL_dout_refl_strength_map = _cuda_f32("dL_dout_refl_strength_map", dL_dout_refl_strength_map)
    P, H, W = means3D.size(0), dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.numel() else 0
    o = dict(dtype=torch.float32, device=means3D.device)
    with torch.cuda.device(means3D.device):
        # The kernels only write the rows of Gaussians a pixel blended; everything else must read 0
        # (G/rasterize_points.cu:186-199 zero-fills eleven tensors).  One zero-filled allocation, eleven views.
        widths = [3, 3, NUM_CHANNELS, 3, 1, 6, 3 * M, 3, 4, 1 if has_inv else 0, 1 if has_refl else 0]
        flat = torch.zeros((P * sum(widths),), **o)
        blocks, off = [], 0
        for wd in widths:
            blocks.append(flat[off:off + P * wd].view(P if wd else 0, max(wd, 1)))
            off += P * wd
        (dL_dmeans3D, dL_dmeans2D, dL_dcolors, dL_dnormals, dL_dopacity, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations,
         dL_dinvdepths, dL_drefl) = blocks
        dL_dsh = dL_dsh.view(P, M, 3) if M else torch.zeros((P, 0, 3), **o)
        if P != 0:
            accum = torch.empty((_b.gauss_accum_bytes(P),), dtype=torch.uint8, device=means3D.device)
            a = _b.GaussBwd(
                P=P, sh_degree=int(degree), sh_coeffs=M, num_rendered=int(R), width=W, height=H,
                tan_fovx=float(tan_fovx), tan_fovy=float(tan_fovy), scale_modifier=float(scale_modifier),
                antialiasing=int(bool(antialiasing)), debug=int(bool(debug)),
                background=_b.ptr(background), means3D=_b.ptr(means3D), shs=_b.ptr(sh), colors_precomp=_b.ptr(colors),
                opacities=_b.ptr(opacities), scales=_b.ptr(scales), rotations=_b.ptr(rotations),
                cov3D_precomp=_b.ptr(cov3D_precomp), viewmatrix=_b.ptr(viewmatrix), projmatrix=_b.ptr(projmatrix),
                campos=_b.ptr(campos), radii=_b.ptr(radii.contiguous()), geom=_b.ptr(geomBuffer),
                binning=_b.ptr(binningBuffer), img=_b.ptr(imageBuffer),
                dL_dout_color=_b.ptr(dL_dout_color), dL_dout_invdepth=_b.ptr(dL_dout_invdepth) if has_inv else None,
                dL_dout_normal=_b.ptr(dL_dout_normal_map),
                dL_dout_refl=_b.ptr(dL_dout_refl_strength_map) if has_refl else None,
                accum=accum.data_ptr(), accum_bytes=accum.numel(),
                dL_dmeans2D=_b.ptr(dL_dmeans2D), dL_dconic=None, dL_dopacity=_b.ptr(dL_dopacity),
                dL_dcolors=_b.ptr(dL_dcolors), dL_dnormals=_b.ptr(dL_dnormals), dL_drefl=_b.ptr(dL_drefl),
                dL_dinvdepths=_b.ptr(dL_dinvdepths), dL_dmeans3D=_b.ptr(dL_dmeans3D), dL_dcov3D=_b.ptr(dL_dcov3D),
                dL_dsh=_b.ptr(dL_dsh), dL_dscales=_b.ptr(dL_dscales), dL_drotations=_b.ptr(dL_drotations),
                geom_bytes=geomBuffer.numel(), binning_bytes=binningBuffer.numel(), img_bytes=imageBuffer.numel())
            _b.check(_b.gauss_backward(C.byref(a), _stream()))
    return (dL_dmeans2D, dL_dcolors, dL_dnormals, dL_drefl, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales,
            dL_drotations)


def mark_visible(means3D, viewmatrix, projmatrix):
    """-> bool[P]   (G/rasterize_points.cu:266-285)"""
    P = means3D.size(0)
    present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
    if P != 0:
        means3D = _cuda_f32("means3D", means3D)
        viewmatrix = _cuda_f32("viewmatrix", viewmatrix)
        projmatrix = _cuda_f32("projmatrix", projmatrix)
        with torch.cuda.device(means3D.device):
            _b.check(_b.mark_visible(P, _b.ptr(means3D), _b.ptr(viewmatrix), _b.ptr(projmatrix), present.data_ptr(),
                                     _stream()))
    return present
