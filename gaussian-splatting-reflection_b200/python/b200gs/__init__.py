"""ctypes binding of libb200gs.so (the C ABI declared in include/b200gs.h).

This is the only place the shared library is loaded.  There is no fallback of
any kind: if the library is missing or a call fails, an exception is raised.
PyTorch is used by the callers of this module for device memory and streams
only; nothing here imports it.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.environ.get(
    "B200GS_LIB", os.path.normpath(os.path.join(_HERE, "..", "..", "lib", "libb200gs.so")))


class B200gsError(RuntimeError):
    pass


if not os.path.exists(_LIB_PATH):
    raise ImportError(
        f"libb200gs.so not found at {_LIB_PATH}; build it with "
        "`python gaussian-splatting-reflection_b200/build.py` (there is no CPU fallback)")
lib = C.CDLL(_LIB_PATH)

_p = C.c_void_p
_f = C.c_float
_i = C.c_int
_sz = C.c_size_t


class SurfelFwd(C.Structure):
    """struct b200gs_surfel_fwd"""
    _fields_ = [
        ("P", _i), ("sh_degree", _i), ("sh_coeffs", _i), ("width", _i), ("height", _i),
        ("tan_fovx", _f), ("tan_fovy", _f), ("scale_modifier", _f),
        ("prefiltered", _i), ("debug", _i),
        ("background", _p), ("means3D", _p), ("env_scope_mask", _p), ("shs", _p),
        ("colors_precomp", _p), ("refl_strengths", _p), ("opacities", _p), ("scales", _p),
        ("rotations", _p), ("transMat_precomp", _p), ("viewmatrix", _p), ("projmatrix", _p),
        ("campos", _p),
        ("geom", _p), ("geom_bytes", _sz), ("img", _p), ("img_bytes", _sz),
        ("binning", _p), ("binning_bytes", _sz),
        ("out_color", _p), ("out_others", _p), ("out_refl", _p), ("radii", _p),
        ("gaussian_weights", _p), ("exact_lists", _i), ("cull_audit", _i), ("depth_slices", _i),
        ("slice_frac", _f * 7),
    ]


class SurfelBwd(C.Structure):
    """struct b200gs_surfel_bwd"""
    _fields_ = [
        ("P", _i), ("sh_degree", _i), ("sh_coeffs", _i), ("num_rendered", _i),
        ("width", _i), ("height", _i),
        ("tan_fovx", _f), ("tan_fovy", _f), ("scale_modifier", _f), ("debug", _i),
        ("background", _p), ("means3D", _p), ("shs", _p), ("colors_precomp", _p), ("scales", _p),
        ("rotations", _p), ("transMat_precomp", _p), ("viewmatrix", _p), ("projmatrix", _p),
        ("campos", _p), ("radii", _p), ("geom", _p), ("binning", _p), ("img", _p),
        ("dL_dout_color", _p), ("dL_dout_others", _p), ("dL_dout_refl", _p),
        ("accum", _p), ("accum_bytes", _sz),
        ("dL_dmeans2D", _p), ("dL_dnormal", _p), ("dL_dopacity", _p), ("dL_dcolors", _p),
        ("dL_drefl", _p), ("dL_dmeans3D", _p), ("dL_dtransMat", _p), ("dL_dsh", _p),
        ("dL_dscales", _p), ("dL_drotations", _p), ("accumulate", _i),
        ("geom_bytes", _sz), ("binning_bytes", _sz), ("img_bytes", _sz),
    ]


class GaussFwd(C.Structure):
    """struct b200gs_gauss_fwd"""
    _fields_ = [
        ("P", _i), ("sh_degree", _i), ("sh_coeffs", _i), ("width", _i), ("height", _i),
        ("tan_fovx", _f), ("tan_fovy", _f), ("scale_modifier", _f),
        ("prefiltered", _i), ("antialiasing", _i), ("debug", _i),
        ("background", _p), ("means3D", _p), ("shs", _p), ("colors_precomp", _p), ("normals", _p),
        ("refl_strengths", _p), ("opacities", _p), ("scales", _p), ("rotations", _p), ("cov3D_precomp", _p),
        ("viewmatrix", _p), ("projmatrix", _p), ("campos", _p),
        ("geom", _p), ("geom_bytes", _sz), ("img", _p), ("img_bytes", _sz), ("binning", _p), ("binning_bytes", _sz),
        ("out_color", _p), ("out_invdepth", _p), ("out_normal", _p), ("out_refl", _p), ("radii", _p),
        ("exact_lists", _i), ("depth_slices", _i), ("slice_frac", _f * 7),
    ]


class GaussBwd(C.Structure):
    """struct b200gs_gauss_bwd"""
    _fields_ = [
        ("P", _i), ("sh_degree", _i), ("sh_coeffs", _i), ("num_rendered", _i), ("width", _i), ("height", _i),
        ("tan_fovx", _f), ("tan_fovy", _f), ("scale_modifier", _f), ("antialiasing", _i), ("debug", _i),
        ("background", _p), ("means3D", _p), ("shs", _p), ("colors_precomp", _p), ("opacities", _p), ("scales", _p),
        ("rotations", _p), ("cov3D_precomp", _p), ("viewmatrix", _p), ("projmatrix", _p), ("campos", _p),
        ("radii", _p), ("geom", _p), ("binning", _p), ("img", _p),
        ("dL_dout_color", _p), ("dL_dout_invdepth", _p), ("dL_dout_normal", _p), ("dL_dout_refl", _p),
        ("accum", _p), ("accum_bytes", _sz),
        ("dL_dmeans2D", _p), ("dL_dconic", _p), ("dL_dopacity", _p), ("dL_dcolors", _p), ("dL_dnormals", _p),
        ("dL_drefl", _p), ("dL_dinvdepths", _p), ("dL_dmeans3D", _p), ("dL_dcov3D", _p), ("dL_dsh", _p),
        ("dL_dscales", _p), ("dL_drotations", _p),
        ("geom_bytes", _sz), ("binning_bytes", _sz), ("img_bytes", _sz),
    ]


def _sig(name, restype, *argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = list(argtypes)
    return fn


last_error = _sig("b200gs_last_error", C.c_char_p)
version = _sig("b200gs_version", _i)
sm_arch = _sig("b200gs_sm_arch", _i)
launch_count = _sig("b200gs_launch_count", C.c_ulonglong)
timing_enable = _sig("b200gs_timing_enable", None, _i)
timing_only = _sig("b200gs_timing_only", None, _i)
timing_stages = _sig("b200gs_timing_stages", _i)
timing_stage_name = _sig("b200gs_timing_stage_name", C.c_char_p, _i)
_timing_read = _sig("b200gs_timing_read", _i, C.POINTER(_f), C.POINTER(C.c_ulonglong), _i)


def timing_read():
    """-> {stage name: (total ms, launches)} accumulated since the last read."""
    n = timing_stages()
    ms = (_f * n)()
    cnt = (C.c_ulonglong * n)()
    _timing_read(ms, cnt, n)
    return {timing_stage_name(i).decode(): (float(ms[i]), int(cnt[i])) for i in range(n)}


_cull_violations = _sig("b200gs_debug_cull_violations", _i, _p, _i)


_cull_stats = _sig("b200gs_debug_cull_stats", _i, _p, _i)


def cull_stats(reset=True):
    out = (C.c_ulonglong * 8)()
    check(_cull_stats(C.cast(out, _p), 1 if reset else 0))
    return [int(x) for x in out]


def cull_violations(reset=True):
    """-> (forward-threshold, backward-threshold) violation counts of the cull audit."""
    out = (C.c_ulonglong * 2)()
    check(_cull_violations(C.cast(out, _p), 1 if reset else 0))
    return int(out[0]), int(out[1])


mark_visible = _sig("b200gs_mark_visible", _i, _i, _p, _p, _p, _p, _p)
surfel_geom_bytes = _sig("b200gs_surfel_geom_bytes", _sz, _i)
surfel_image_bytes = _sig("b200gs_surfel_image_bytes", _sz, _i, _i)
binning_bytes = _sig("b200gs_binning_bytes", _sz, _i)
surfel_accum_bytes = _sig("b200gs_surfel_accum_bytes", _sz, _i)
MAX_SLICES, PLAN_INTS = 8, 24
surfel_forward_stage1 = _sig("b200gs_surfel_forward_stage1", _i, C.POINTER(SurfelFwd), C.POINTER(_i), _p)
surfel_forward_stage2 = _sig("b200gs_surfel_forward_stage2", _i, C.POINTER(SurfelFwd), C.POINTER(_i), _p)
surfel_backward = _sig("b200gs_surfel_backward", _i, C.POINTER(SurfelBwd), _p)
# the same backward with its accumulator in a static device array inside the library (no allocation anywhere on the
# call; P <= STATIC_ACCUM_MAX_POINTS; calls on one device share the array and must not overlap on the GPU)
STATIC_ACCUM_MAX_POINTS = 131072
surfel_backward_static = _sig("b200gs_surfel_backward_static", _i, C.POINTER(SurfelBwd), _p)
surfel_inspect_geom = _sig("b200gs_surfel_inspect_geom", _i, _p, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p)
inspect_binning = _sig("b200gs_inspect_binning", _i, _p, _i, _p, _p, _p, _p)
surfel_inspect_image = _sig("b200gs_surfel_inspect_image", _i, _p, _i, _i, _p, _p, _p, _p)


surfel_inspect_tiles = _sig("b200gs_surfel_inspect_tiles", _i, _p, _i, _i, _p, _p, _p)


gauss_geom_bytes = _sig("b200gs_gauss_geom_bytes", _sz, _i)
gauss_image_bytes = _sig("b200gs_gauss_image_bytes", _sz, _i, _i)
gauss_accum_bytes = _sig("b200gs_gauss_accum_bytes", _sz, _i)
gauss_forward_stage1 = _sig("b200gs_gauss_forward_stage1", _i, C.POINTER(GaussFwd), C.POINTER(_i), _p)
gauss_forward_stage2 = _sig("b200gs_gauss_forward_stage2", _i, C.POINTER(GaussFwd), C.POINTER(_i), _p)
gauss_backward = _sig("b200gs_gauss_backward", _i, C.POINTER(GaussBwd), _p)
gauss_inspect_geom = _sig("b200gs_gauss_inspect_geom", _i, _p, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p)
gauss_inspect_image = _sig("b200gs_gauss_inspect_image", _i, _p, _i, _i, _p, _p, _p, _p)
_u = C.c_uint32
cubemap_forward = _sig("b200gs_cubemap_forward", _i, _p, _p, _p, _p, _u, _u, _u, _u, _u, _p)
cubemap_backward = _sig("b200gs_cubemap_backward", _i, _p, _p, _p, _p, _p, _p, _u, _u, _u, _u, _u, _p)


ADAM_MAX_GROUPS = 16


class AdamGroup(C.Structure):
    """struct b200gs_adam_group"""
    _fields_ = [("param", _p), ("grad", _p), ("exp_avg", _p), ("exp_avg_sq", _p), ("n", C.c_longlong),
                ("param_row", _i), ("grad_row", _i), ("grad_col", _i), ("lr", C.c_double), ("step", C.c_longlong)]


class Adam(C.Structure):
    """struct b200gs_adam"""
    _fields_ = [("groups", _i), ("group", AdamGroup * ADAM_MAX_GROUPS), ("beta1", C.c_double), ("beta2", C.c_double),
                ("eps", C.c_double), ("step", C.c_longlong)]


adam_step = _sig("b200gs_adam_step", _i, C.POINTER(Adam), _p)
densify_stats = _sig("b200gs_densify_stats", _i, _i, _p, _p, _p, _p, _p, _p, _p, _p, _p)


peer_allreduce_f32 = _sig("b200gs_peer_allreduce_f32", _i, _p, C.POINTER(_p), _i, _i, _sz, _i, _p)


ssim_forward = _sig("b200gs_ssim_forward", _i, _f, _f, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p)
ssim_backward = _sig("b200gs_ssim_backward", _i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p)


_fp = C.POINTER(_f)
reflect_shade_forward = _sig("b200gs_reflect_shade_forward", _i, _i, _i, _fp, _fp, _fp, _p, _p, _p, _p, _p, _u, _u, _u, _p, _p, _p)
reflect_shade_backward = _sig("b200gs_reflect_shade_backward", _i, _i, _i, _fp, _fp, _fp, _p, _p, _p, _p, _u, _u, _u, _p, _p, _p,
                              _p, _p, _p, _p, _p, _p)


surface_forward = _sig("b200gs_surface_forward", _i, _i, _i, _p, _fp, _fp, _fp, _f, _p, _p, _p, _p)
surface_backward = _sig("b200gs_surface_backward", _i, _i, _i, _p, _fp, _fp, _fp, _f, _p, _p, _p, _p, _p, _p)


PLY_MAX_FIELDS, PLY_MAX_COLUMNS = 8, 128


class PlyField(C.Structure):
    """struct b200gs_ply_field"""
    _fields_ = [("data", _p), ("cols", _i), ("row_stride", _i), ("first", _i)]


class PlyLayout(C.Structure):
    """struct b200gs_ply_layout"""
    _fields_ = [("stride", _i), ("n_fields", _i), ("field", PlyField * PLY_MAX_FIELDS), ("map", C.c_ubyte * PLY_MAX_COLUMNS)]


ply_unpack = _sig("b200gs_ply_unpack", _i, C.c_longlong, _p, C.POINTER(PlyLayout), _p)
ply_pack = _sig("b200gs_ply_pack", _i, C.c_longlong, C.POINTER(PlyLayout), _p, _p)


struct_bytes = _sig("b200gs_struct_bytes", _sz, _i)
for _k, _cls in enumerate((SurfelFwd, SurfelBwd, GaussFwd, GaussBwd, Adam, PlyLayout)):
    if struct_bytes(_k) != C.sizeof(_cls):
        raise ImportError(f"libb200gs.so and its Python binding disagree on the layout of {_cls.__name__} "
                          f"({struct_bytes(_k)} vs {C.sizeof(_cls)} bytes): rebuild the library")


def check(rc):
    """Map a C return code to an exception (the reference throws std::runtime_error
    -> Python RuntimeError; S/rasterize_points.cu:63-65, auxiliary.h:300-307)."""
    if rc != 0:
        raise B200gsError(last_error().decode("utf-8", "replace"))


def ptr(t):
    """Device pointer of a tensor, or None for a missing / empty one (the reference
    treats an empty tensor as 'not provided': S/diff_surfel_rasterization/__init__.py:210-223)."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()
