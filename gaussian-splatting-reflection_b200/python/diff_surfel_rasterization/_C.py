"""`diff_surfel_rasterization._C` -- same three entry points, argument order and
result tuples as the reference's pybind module (S/ext.cpp:15-19,
S/rasterize_points.h:18-72), implemented over the C ABI of libb200gs.so.

This shim does what S/rasterize_points.cu does around the core: argument checks,
`.contiguous()`, output / scratch allocation as torch tensors, and mapping error
codes to RuntimeError.  All arithmetic happens in the sm_100a kernels.
"""
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


def rasterize_gaussians_prepare(background, means3D, env_scope_mask, colors, refl_strengths, opacity, scales, rotations,
                        scale_modifier, transMat_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                        image_height, image_width, sh, degree, campos, prefiltered, debug, *, exact_lists=False,
                        cull_audit=False, depth_slices=0, slice_fractions=None):
    """Stage 1 of `rasterize_gaussians` (same arguments): checks, output / scratch allocation, preprocess, depth order,
    prefix sum and slice plan on the CURRENT stream, returning after that stream has drained (the one host sync of
    the forward) with a PreparedForward for `rasterize_gaussians_finish`.  No reference counterpart: the split lets a
    caller that knows its next camera run this part of the NEXT view on a second (high-priority) stream while the
    compositing kernels of the current one occupy the SMs (b200gs.dp.ViewPipeline).  The handle keeps every input
    tensor alive; the parameters must not change between the two calls."""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    background = _cuda_f32("background", background)
    means3D = _cuda_f32("means3D", means3D)
    colors = _cuda_f32("colors", colors)
    opacity = _cuda_f32("opacity", opacity)
    scales = _cuda_f32("scales", scales)
    rotations = _cuda_f32("rotations", rotations)
    transMat_precomp = _cuda_f32("transMat_precomp", transMat_precomp)
    viewmatrix = _cuda_f32("viewmatrix", viewmatrix)
    projmatrix = _cuda_f32("projmatrix", projmatrix)
    sh = _cuda_f32("sh", sh)
    campos = _cuda_f32("campos", campos)
    if refl_strengths is not None:
        refl_strengths = _cuda_f32("refl_strengths", refl_strengths)
    if env_scope_mask is not None and env_scope_mask.numel():
        if env_scope_mask.dtype != torch.bool:
            raise RuntimeError(f"expected scalar type Bool but found {env_scope_mask.dtype}")
        env_scope_mask = env_scope_mask.contiguous()
        if env_scope_mask.numel() < means3D.size(0):
            raise RuntimeError("env_scope_mask has fewer entries than there are points")

    P, H, W = means3D.size(0), int(image_height), int(image_width)
    dev = means3D.device
    f32 = dict(dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        # the reference zero-fills these (S/rasterize_points.cu:88-95); the kernels write every pixel and
        # every radius themselves, so only the P == 0 short-circuit and the atomicMax target need zeros
        # One allocation for the 12 output planes and one for the per-Gaussian outputs (views are handed out): on
        # small scenes the call is launch-bound and every allocation / fill is a launch or a host round trip.
        new = torch.zeros if P == 0 else torch.empty
        planes = new((NUM_CHANNELS + 8 + 1, H, W), **f32)
        out_color, out_others, out_refl = planes[:NUM_CHANNELS], planes[NUM_CHANNELS:NUM_CHANNELS + 8], planes[NUM_CHANNELS + 8:]
        per_gaussian = new((2, P), **f32)
        radii = per_gaussian[0].view(torch.int32)
        gaussian_weights = per_gaussian[1]              # zeroed by stage 1 (the atomicMax target of the compositing)
        u8 = dict(dtype=torch.uint8, device=dev)
        geom = torch.empty((0,), **u8)
        binning = torch.empty((0,), **u8)
        img = torch.empty((0,), **u8)
        rendered = 0
        if P != 0:
            M = sh.size(1) if sh.numel() else 0
            geom = torch.empty((_b.surfel_geom_bytes(P),), **u8)
            img = torch.empty((_b.surfel_image_bytes(W, H),), **u8)
            a = _b.SurfelFwd(
                P=P, sh_degree=int(degree), sh_coeffs=M, width=W, height=H,
                tan_fovx=float(tan_fovx), tan_fovy=float(tan_fovy), scale_modifier=float(scale_modifier),
                prefiltered=int(bool(prefiltered)), debug=int(bool(debug)),
                background=_b.ptr(background), means3D=_b.ptr(means3D), env_scope_mask=_b.ptr(env_scope_mask),
                shs=_b.ptr(sh), colors_precomp=_b.ptr(colors), refl_strengths=_b.ptr(refl_strengths),
                opacities=_b.ptr(opacity), scales=_b.ptr(scales), rotations=_b.ptr(rotations),
                transMat_precomp=_b.ptr(transMat_precomp), viewmatrix=_b.ptr(viewmatrix),
                projmatrix=_b.ptr(projmatrix), campos=_b.ptr(campos),
                geom=geom.data_ptr(), geom_bytes=geom.numel(), img=img.data_ptr(), img_bytes=img.numel(),
                binning=None, binning_bytes=0,
                out_color=out_color.data_ptr(), out_others=out_others.data_ptr(), out_refl=out_refl.data_ptr(),
                radii=radii.data_ptr(), gaussian_weights=gaussian_weights.data_ptr(),
                exact_lists=int(bool(exact_lists)), cull_audit=int(bool(cull_audit)), depth_slices=int(depth_slices))
            if slice_fractions:
                for i, fr in enumerate(list(slice_fractions)[:7]):
                    a.slice_frac[i] = float(fr)
            plan = (C.c_int * _b.PLAN_INTS)()
            st = _stream()
            _b.check(_b.surfel_forward_stage1(C.byref(a), plan, st))
            rendered = plan[0]
    prep = PreparedForward()
    prep.a, prep.plan, prep.P, prep.device = (a if P != 0 else None), (plan if P != 0 else None), P, dev
    prep.outputs = (planes, per_gaussian, out_color, out_others, out_refl, radii, gaussian_weights)
    prep.scratch = (geom, img)
    prep.keep = (background, means3D, env_scope_mask, colors, refl_strengths, opacity, scales, rotations,
                 transMat_precomp, viewmatrix, projmatrix, sh, campos)
    prep.rendered = rendered
    return prep


class PreparedForward:
    """What stage 1 of a forward leaves for stage 2: the argument struct, the plan, outputs and scratch."""
    __slots__ = ("a", "plan", "P", "device", "outputs", "scratch", "keep", "rendered")


def rasterize_gaussians_finish(prep):
    """Stage 2 of a prepared forward on the CURRENT stream (asynchronous): per depth slice emission, tile sort, tile
    ranges, block masks, compositing.  -> the 9-tuple of `rasterize_gaussians`."""
    planes, per_gaussian, out_color, out_others, out_refl, radii, gaussian_weights = prep.outputs
    geom, img = prep.scratch
    u8 = dict(dtype=torch.uint8, device=prep.device)
    binning = torch.empty((0,), **u8)
    if prep.P != 0:
        with torch.cuda.device(prep.device):
            cur = torch.cuda.current_stream()
            for t in (planes, per_gaussian, geom, img):      # allocated on stage 1's stream, used (and later freed) here
                t.record_stream(cur)
            a = prep.a
            binning = torch.empty((_b.binning_bytes(prep.rendered),), **u8)
            a.binning = binning.data_ptr()
            a.binning_bytes = binning.numel()
            _b.check(_b.surfel_forward_stage2(C.byref(a), prep.plan, _stream()))
    return prep.rendered, out_color, out_others, radii, geom, binning, img, out_refl, gaussian_weights


def rasterize_gaussians(background, means3D, env_scope_mask, colors, refl_strengths, opacity, scales, rotations,
                        scale_modifier, transMat_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                        image_height, image_width, sh, degree, campos, prefiltered, debug, *, exact_lists=False,
                        cull_audit=False, depth_slices=0, slice_fractions=None):
    """-> (rendered, out_color[3,H,W], out_others[8,H,W], radii[P], geomBuffer, binningBuffer,
    imgBuffer, out_refl_strength_map[1,H,W], gaussian_weights[P])   (S/rasterize_points.cu:39-151)

    The 21 positional arguments are the reference's.  Keyword-only extras of this library, per call (the
    library keeps no mode of its own): exact_lists -- emit the reference's instance lists bit for bit
    instead of only the reachable instances (include/b200gs.h); cull_audit -- test infrastructure; depth_slices
    (0 automatic, 1 off, n at most n) / slice_fractions -- the occlusion-aware binning of b200gs_surfel_forward_stage1."""
    return rasterize_gaussians_finish(rasterize_gaussians_prepare(
        background, means3D, env_scope_mask, colors, refl_strengths, opacity, scales, rotations, scale_modifier,
        transMat_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
        prefiltered, debug, exact_lists=exact_lists, cull_audit=cull_audit, depth_slices=depth_slices,
        slice_fractions=slice_fractions))


def rasterize_gaussians_backward(background, means3D, radii, colors, refl_strengths, scales, rotations, scale_modifier,
                                 transMat_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color,
                                 dL_dout_others, dL_dout_refl_strength_map, sh, degree, campos, geomBuffer, R,
                                 binningBuffer, imageBuffer, debug, *, grad_arena=None):
    """-> (dL_dmeans2D[P,3], dL_dcolors[P,3], dL_drefl[P,1], dL_dopacity[P,1], dL_dmeans3D[P,3],
    dL_dtransMat[P,9], dL_dsh[P,M,3], dL_dscales[P,2], dL_drotations[P,4])   (S/rasterize_points.cu:153-267)

    The 24 positional arguments are the reference's.  grad_arena (keyword-only, per call): a
    b200gs.dp.GradArena -- the parameter gradients dL_dmeans3D / dL_dsh / dL_dopacity / dL_dscales /
    dL_drotations / dL_drefl are then written into (accumulate mode: ADDED to) the blocks of that flat
    buffer, which NCCL reduces as it lies, and the returned tensors are views of it."""
    background = _cuda_f32("background", background)
    means3D = _cuda_f32("means3D", means3D)
    colors = _cuda_f32("colors", colors)
    scales = _cuda_f32("scales", scales)
    rotations = _cuda_f32("rotations", rotations)
    transMat_precomp = _cuda_f32("transMat_precomp", transMat_precomp)
    viewmatrix = _cuda_f32("viewmatrix", viewmatrix)
    projmatrix = _cuda_f32("projmatrix", projmatrix)
    sh = _cuda_f32("sh", sh)
    campos = _cuda_f32("campos", campos)
    for name, t in (("radii", radii), ("geomBuffer", geomBuffer), ("binningBuffer", binningBuffer),
                    ("imageBuffer", imageBuffer)):
        if not t.is_cuda:
            raise RuntimeError(f"{name} must be a CUDA tensor")
    dL_dout_color = _cuda_f32("dL_dout_color", dL_dout_color)
    dL_dout_others = _cuda_f32("dL_dout_others", dL_dout_others)
    has_refl = dL_dout_refl_strength_map is not None and dL_dout_refl_strength_map.size(0) != 0
    if has_refl:
        dL_dout_refl_strength_map = _cuda_f32("dL_dout_refl_strength_map", dL_dout_refl_strength_map)

    P, H, W = means3D.size(0), dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.numel() else 0
    o = dict(dtype=torch.float32, device=means3D.device)
    with torch.cuda.device(means3D.device):
        arena = grad_arena
        if arena is not None and (arena.P != P or arena.M != M or not has_refl or arena.flat.device != means3D.device):
            raise RuntimeError("gradient arena does not match this call (P, SH coefficients, refl map or device)")
        if arena is not None:
            if not arena.accumulate:
                arena.zero_()
            av = arena.views()
        # The kernels only write the rows of Gaussians a pixel blended; everything else must read 0
        # (S/rasterize_points.cu:186-197 zero-fills ten tensors).  One zero-filled allocation, ten views.
        widths = [3, NUM_CHANNELS, 3, 9] + ([] if arena is not None else [3, 1, 3 * M, 2, 4, 1 if has_refl else 0])
        flat = torch.zeros((P * sum(widths),), **o)
        blocks, off = [], 0
        for wd in widths:
            blocks.append(flat[off:off + P * wd].view(P, wd))
            off += P * wd
        dL_dmeans2D, dL_dcolors, dL_dnormal, dL_dtransMat = blocks[:4]
        if arena is not None:
            dL_dmeans3D, dL_dopacity, dL_dsh, dL_dscales = av["means3D"], av["opacity"], av["sh"], av["scales"]
            dL_drotations, dL_drefl = av["rotations"], av["refl"]
        else:
            dL_dmeans3D, dL_dopacity, dL_dsh, dL_dscales, dL_drotations, dL_drefl = blocks[4:]
            dL_dsh = dL_dsh.view(P, M, 3)
            if not has_refl:
                dL_drefl = dL_drefl.view(0, 1)
        if P != 0:
            accum = torch.empty((_b.surfel_accum_bytes(P),), dtype=torch.uint8, device=means3D.device)
            a = _b.SurfelBwd(
                P=P, sh_degree=int(degree), sh_coeffs=M, num_rendered=int(R), width=W, height=H,
                tan_fovx=float(tan_fovx), tan_fovy=float(tan_fovy), scale_modifier=float(scale_modifier),
                debug=int(bool(debug)),
                background=_b.ptr(background), means3D=_b.ptr(means3D), shs=_b.ptr(sh), colors_precomp=_b.ptr(colors),
                scales=_b.ptr(scales), rotations=_b.ptr(rotations), transMat_precomp=_b.ptr(transMat_precomp),
                viewmatrix=_b.ptr(viewmatrix), projmatrix=_b.ptr(projmatrix), campos=_b.ptr(campos),
                radii=_b.ptr(radii.contiguous()), geom=_b.ptr(geomBuffer), binning=_b.ptr(binningBuffer),
                img=_b.ptr(imageBuffer),
                dL_dout_color=_b.ptr(dL_dout_color), dL_dout_others=_b.ptr(dL_dout_others),
                dL_dout_refl=_b.ptr(dL_dout_refl_strength_map) if has_refl else None,
                accum=accum.data_ptr(), accum_bytes=accum.numel(),
                dL_dmeans2D=_b.ptr(dL_dmeans2D), dL_dnormal=_b.ptr(dL_dnormal), dL_dopacity=_b.ptr(dL_dopacity),
                dL_dcolors=_b.ptr(dL_dcolors), dL_drefl=_b.ptr(dL_drefl), dL_dmeans3D=_b.ptr(dL_dmeans3D),
                dL_dtransMat=_b.ptr(dL_dtransMat), dL_dsh=_b.ptr(dL_dsh), dL_dscales=_b.ptr(dL_dscales),
                dL_drotations=_b.ptr(dL_drotations),
                accumulate=1 if (arena is not None and arena.accumulate) else 0,
                geom_bytes=geomBuffer.numel(), binning_bytes=binningBuffer.numel(), img_bytes=imageBuffer.numel())
            _b.check(_b.surfel_backward(C.byref(a), _stream()))
    return (dL_dmeans2D, dL_dcolors, dL_drefl, dL_dopacity, dL_dmeans3D, dL_dtransMat, dL_dsh, dL_dscales,
            dL_drotations)


def mark_visible(means3D, viewmatrix, projmatrix):
    """-> bool[P]   (S/rasterize_points.cu:269-288)"""
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
