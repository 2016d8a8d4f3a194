"""Shared drivers for the GPU parity tests, smoke() and bench.py.

`run_surfel` drives ANY module that has the reference's `_C` signature
(rasterize_gaussians / rasterize_gaussians_backward), so exactly the same call
runs this repo's sm_100a path and the unmodified reference extension from
oracle/_ref on identical inputs.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_PY = os.path.join(ROOT, "gaussian-splatting-reflection_b200", "python")
for p in (ROOT, PKG_PY):
    if p not in sys.path:
        sys.path.insert(0, p)


import contextlib


_SLICES = [0]        # depth slices the drivers pass (0: the library's automatic choice)
_EXACT = [False]     # list mode the drivers below pass to THIS library's `_C` calls (a per-call keyword there)


@contextlib.contextmanager
def exact_lists(on=True):
    """Drive the new library in exact-lists mode (the reference's instance lists bit for bit) or in the
    default culled-emission mode (include/b200gs.h: `exact_lists` of the forward struct).  The library has
    no mode of its own: this only sets what surfel_forward / gauss_forward below pass per call."""
    old = _EXACT[0]
    _EXACT[0] = bool(on)
    try:
        yield
    finally:
        _EXACT[0] = old


def _is_new(Cmod):
    return getattr(Cmod, "__name__", "").startswith(("diff_surfel_rasterization", "diff_gaussian_rasterization"))


def new_surfel_C():
    import diff_surfel_rasterization
    return diff_surfel_rasterization._C


def ref_module(name):
    """Prebuilt reference extension from oracle/_ref, or None."""
    from oracle import build_ref
    try:
        return build_ref.load_ref(name)
    except Exception as ex:  # noqa: BLE001
        print(f"[harness] cannot load {name}: {ex}")
        return None


_BG = {}


def _bg(bg, dev):
    """Background colour on the device, uploaded once per value (a per-call torch.tensor(bg, device=...) is a
    pageable H2D copy + sync that would be timed with every small-scene call)."""
    key = (tuple(float(x) for x in bg), str(dev))
    if key not in _BG:
        _BG[key] = torch.tensor(key[0], dtype=torch.float32, device=dev)
    return _BG[key]


def to_dev(d, device="cuda"):
    return {k: (torch.from_numpy(np.ascontiguousarray(v)).to(device) if isinstance(v, np.ndarray) else v)
            for k, v in d.items()}


def surfel_forward(Cmod, scene, cam, bg=(0.0, 0.0, 0.0), sh_degree=3, mask=None, scale_modifier=1.0,
                   colors_precomp=None, transMat_precomp=None, debug=False, prefiltered=False, exact=None, cull_audit=False,
                   depth_slices=None, slice_fractions=None, prepare_only=False):
    """scene/cam: dicts of CUDA tensors (see scenes.py).  Returns the 9-tuple as a dict."""
    dev = scene["means3D"].device
    e = torch.empty((0,), dtype=torch.float32, device=dev)
    P = scene["means3D"].shape[0]
    if mask is None:
        mask = torch.ones((P, 3), dtype=torch.bool, device=dev)
    bgt = _bg(bg, dev)
    use_sh = colors_precomp is None
    kw = {}
    if _is_new(Cmod):
        kw = dict(exact_lists=_EXACT[0] if exact is None else bool(exact), cull_audit=cull_audit,
                  depth_slices=_SLICES[0] if depth_slices is None else depth_slices, slice_fractions=slice_fractions)
    fn = Cmod.rasterize_gaussians_prepare if prepare_only else Cmod.rasterize_gaussians
    out = fn(
        bgt, scene["means3D"], mask, e if use_sh else colors_precomp, scene["refl"], scene["opacities"],
        e if transMat_precomp is not None else scene["scales"], e if transMat_precomp is not None else scene["rotations"],
        scale_modifier, e if transMat_precomp is None else transMat_precomp, cam["viewmatrix"], cam["projmatrix"],
        cam["tanfovx"], cam["tanfovy"], cam["height"], cam["width"], scene["shs"] if use_sh else e, sh_degree,
        cam["campos"], prefiltered, debug, **kw)
    extra = dict(bg=bgt, sh_degree=sh_degree, scale_modifier=scale_modifier, colors_precomp=colors_precomp,
                 transMat_precomp=transMat_precomp)
    if prepare_only:
        return out, extra
    r = dict(zip(_SURFEL_KEYS, out))
    r.update(extra)
    return r


_SURFEL_KEYS = ("rendered", "color", "others", "radii", "geom", "binning", "img", "refl_map", "weights")


def surfel_prepare(Cmod, scene, cam, **kw):
    """Stage 1 of surfel_forward on the current stream (this library only): -> handle for surfel_finish."""
    return surfel_forward(Cmod, scene, cam, prepare_only=True, **kw)


def surfel_finish(Cmod, handle):
    """Stage 2 of a prepared forward on the current stream -> the dict surfel_forward returns."""
    prep, extra = handle
    r = dict(zip(_SURFEL_KEYS, Cmod.rasterize_gaussians_finish(prep)))
    r.update(extra)
    return r


def surfel_backward(Cmod, scene, cam, fwd, grads, grad_arena=None):
    dev = scene["means3D"].device
    kw = dict(grad_arena=grad_arena) if grad_arena is not None else {}
    e = torch.empty((0,), dtype=torch.float32, device=dev)
    use_sh = fwd["colors_precomp"] is None
    tp = fwd["transMat_precomp"]
    out = Cmod.rasterize_gaussians_backward(
        fwd["bg"], scene["means3D"], fwd["radii"], e if use_sh else fwd["colors_precomp"], scene["refl"],
        e if tp is not None else scene["scales"], e if tp is not None else scene["rotations"], fwd["scale_modifier"],
        e if tp is None else tp, cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"],
        grads["color"], grads["others"], grads["refl"], scene["shs"] if use_sh else e, fwd["sh_degree"],
        cam["campos"], fwd["geom"], fwd["rendered"], fwd["binning"], fwd["img"], False, **kw)
    keys = ("means2D", "colors", "refl", "opacity", "means3D", "transMat", "sh", "scales", "rotations")
    return dict(zip(keys, out))


# ---------------------------------------------------------------- state extraction
def _al(off, a=128):
    return (off + a - 1) // a * a


def ref_surfel_state(fwd, P, W, H):
    """Parse the reference's geomBuffer / binningBuffer / imgBuffer (SURVEY Appendix B;
    S/cuda_rasterizer/rasterizer_impl.cu:155-194).  point_offsets lies behind the
    CUB temp block and is recomputed from tiles_touched."""
    g = fwd["geom"]
    base = g.data_ptr() % 128
    assert base == 0
    off = 0
    def take(buf, off, nbytes, dtype, shape):
        off = _al(off)
        t = buf[off:off + nbytes].view(dtype).reshape(shape)
        return t, off + nbytes
    depths, off = take(g, off, 4 * P, torch.float32, (P,))
    clamped, off = take(g, off, 3 * P, torch.uint8, (P, 3))
    iradii, off = take(g, off, 4 * P, torch.int32, (P,))
    means2D, off = take(g, off, 8 * P, torch.float32, (P, 2))
    transMat, off = take(g, off, 36 * P, torch.float32, (P, 9))
    normal_opacity, off = take(g, off, 16 * P, torch.float32, (P, 4))
    rgb, off = take(g, off, 12 * P, torch.float32, (P, 3))
    tiles, off = take(g, off, 4 * P, torch.int32, (P,))
    R = fwd["rendered"]
    b = fwd["binning"]
    off = 0
    point_list, off = take(b, off, 4 * R, torch.int32, (R,))
    _, off = take(b, off, 4 * R, torch.int32, (R,))
    keys, off = take(b, off, 8 * R, torch.int64, (R,))
    im = fwd["img"]
    N = W * H
    T = ((W + 15) // 16) * ((H + 15) // 16)
    off = 0
    accum, off = take(im, off, 12 * N, torch.float32, (3, N))
    ncon, off = take(im, off, 8 * N, torch.int32, (2, N))
    ranges, off = take(im, off, 8 * N, torch.int32, (N, 2))
    return dict(depths=depths, clamped=clamped, means2D=means2D, transMat=transMat, normal_opacity=normal_opacity,
                rgb=rgb, tiles_touched=tiles, point_offsets=torch.cumsum(tiles, 0).to(torch.int32),
                point_list=point_list, keys=keys, accum=accum, n_contrib=ncon, ranges=ranges[:T])


def new_surfel_state(fwd, P, W, H):
    """Unpack this library's private scratch through the C ABI inspect helpers."""
    import ctypes as C
    import b200gs as _b
    dev = fwd["geom"].device
    f = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)  # noqa: E731
    i = lambda *s: torch.zeros(s, dtype=torch.int32, device=dev)  # noqa: E731
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    depths, means2D, transMat, no, rgb = f(P), f(P, 2), f(P, 9), f(P, 4), f(P, 3)
    tiles, offs = i(P), i(P)
    clamped = torch.zeros((P, 3), dtype=torch.uint8, device=dev)
    _b.check(_b.surfel_inspect_geom(fwd["geom"].data_ptr(), P, depths.data_ptr(), means2D.data_ptr(),
                                    transMat.data_ptr(), no.data_ptr(), rgb.data_ptr(), tiles.data_ptr(),
                                    offs.data_ptr(), clamped.data_ptr(), st))
    R = fwd["rendered"]
    keys = torch.zeros((R,), dtype=torch.int64, device=dev)
    pl = i(R)
    if R:
        _b.check(_b.inspect_binning(fwd["binning"].data_ptr(), R, depths.data_ptr(), keys.data_ptr(), pl.data_ptr(), st))
    N = W * H
    T = ((W + 15) // 16) * ((H + 15) // 16)
    accum, ncon, ranges = f(3, N), i(2, N), i(T, 2)
    _b.check(_b.surfel_inspect_image(fwd["img"].data_ptr(), W, H, accum.data_ptr(), ncon.data_ptr(),
                                     ranges.data_ptr(), st))
    return dict(depths=depths, clamped=clamped, means2D=means2D, transMat=transMat, normal_opacity=no, rgb=rgb,
                tiles_touched=tiles, point_offsets=offs, point_list=pl, keys=keys, accum=accum, n_contrib=ncon,
                ranges=ranges)


# ---------------------------------------------------------------- comparison metrics
def max_abs(a, b):
    return float((a.double() - b.double()).abs().max()) if a.numel() else 0.0


def rel_l2(a, b):
    d = (a.double() - b.double()).norm()
    n = b.double().norm()
    return float(d / n) if n > 0 else float(d)


def grad_close_fraction(a, b, rtol=1e-4, scale_floor=1e-4):
    """Fraction of elements with |a-b| <= rtol * max(|a|,|b|) + rtol*scale_floor*rms(b)."""
    a, b = a.double(), b.double()
    rms = b.pow(2).mean().sqrt() if b.numel() else torch.tensor(0.0)
    tol = rtol * torch.maximum(a.abs(), b.abs()) + rtol * max(float(rms), 1e-30)
    ok = (a - b).abs() <= tol
    return float(ok.double().mean()) if ok.numel() else 1.0


# ================================================================ 3DGS (EWA) path
def new_gauss_C():
    import diff_gaussian_rasterization
    return diff_gaussian_rasterization._C


def gauss_scene(scene):
    """3DGS inputs from a surfel test scene: a third scale, per-Gaussian unit normals."""
    g = dict(scene)
    s = scene["scales"]
    g["scales3"] = torch.cat([s, 0.3 * s[:, :1]], 1).contiguous()
    q = scene["rotations"]
    n = torch.stack([2 * (q[:, 1] * q[:, 3] + q[:, 0] * q[:, 2]), 2 * (q[:, 2] * q[:, 3] - q[:, 0] * q[:, 1]),
                     1 - 2 * (q[:, 1] ** 2 + q[:, 2] ** 2)], 1)
    g["normals"] = n.contiguous()
    return g


def gauss_forward(Cmod, scene, cam, bg=(0.0, 0.0, 0.0), sh_degree=3, scale_modifier=1.0, antialiasing=False,
                  colors_precomp=None, cov3D_precomp=None, debug=False, exact=None, depth_slices=None,
                  slice_fractions=None):
    dev = scene["means3D"].device
    e = torch.empty((0,), dtype=torch.float32, device=dev)
    bgt = _bg(bg, dev)
    use_sh = colors_precomp is None
    kw = {}
    if _is_new(Cmod):
        kw = dict(exact_lists=_EXACT[0] if exact is None else bool(exact),
                  depth_slices=_SLICES[0] if depth_slices is None else depth_slices, slice_fractions=slice_fractions)
    out = Cmod.rasterize_gaussians(
        bgt, scene["means3D"], e if use_sh else colors_precomp, scene["normals"], scene["refl"], scene["opacities"],
        e if cov3D_precomp is not None else scene["scales3"], e if cov3D_precomp is not None else scene["rotations"],
        scale_modifier, e if cov3D_precomp is None else cov3D_precomp, cam["viewmatrix"], cam["projmatrix"],
        cam["tanfovx"], cam["tanfovy"], cam["height"], cam["width"], scene["shs"] if use_sh else e, sh_degree,
        cam["campos"], False, antialiasing, debug, **kw)
    keys = ("rendered", "color", "radii", "geom", "binning", "img", "invdepth", "normal_map", "refl_map")
    r = dict(zip(keys, out))
    r.update(bg=bgt, sh_degree=sh_degree, scale_modifier=scale_modifier, antialiasing=antialiasing,
             colors_precomp=colors_precomp, cov3D_precomp=cov3D_precomp)
    return r


def gauss_backward(Cmod, scene, cam, fwd, grads):
    """grads: color[3,H,W], invdepth[1,H,W], normal[3,H,W], refl[1,H,W]."""
    dev = scene["means3D"].device
    e = torch.empty((0,), dtype=torch.float32, device=dev)
    use_sh = fwd["colors_precomp"] is None
    cp = fwd["cov3D_precomp"]
    out = Cmod.rasterize_gaussians_backward(
        fwd["bg"], scene["means3D"], fwd["radii"], e if use_sh else fwd["colors_precomp"], scene["normals"],
        scene["refl"], scene["opacities"], e if cp is not None else scene["scales3"],
        e if cp is not None else scene["rotations"], fwd["scale_modifier"], e if cp is None else cp, cam["viewmatrix"],
        cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], grads["color"], grads["invdepth"], grads["normal"],
        grads["refl"], scene["shs"] if use_sh else e, fwd["sh_degree"], cam["campos"], fwd["geom"], fwd["rendered"],
        fwd["binning"], fwd["img"], fwd["antialiasing"], False)
    keys = ("means2D", "colors", "normals", "refl", "opacity", "means3D", "cov3D", "sh", "scales", "rotations")
    return dict(zip(keys, out))


def gauss_grads(W, H, seed=0):
    rng = np.random.default_rng(991 + seed)
    return dict(color=rng.normal(0, 1, (3, H, W)).astype(np.float32), invdepth=rng.normal(0, 1, (1, H, W)).astype(np.float32),
                normal=rng.normal(0, 1, (3, H, W)).astype(np.float32), refl=rng.normal(0, 1, (1, H, W)).astype(np.float32))


def ref_gauss_state(fwd, P, W, H):
    """Parse the reference's buffers of the 3DGS path (G/cuda_rasterizer/rasterizer_impl.cu:155-194)."""
    def take(buf, off, nbytes, dtype, shape):
        off = _al(off)
        return buf[off:off + nbytes].view(dtype).reshape(shape), off + nbytes
    g, off = fwd["geom"], 0
    depths, off = take(g, off, 4 * P, torch.float32, (P,))
    clamped, off = take(g, off, 3 * P, torch.uint8, (P, 3))
    iradii, off = take(g, off, 4 * P, torch.int32, (P,))
    means2D, off = take(g, off, 8 * P, torch.float32, (P, 2))
    cov3D, off = take(g, off, 24 * P, torch.float32, (P, 6))
    conic_opacity, off = take(g, off, 16 * P, torch.float32, (P, 4))
    rgb, off = take(g, off, 12 * P, torch.float32, (P, 3))
    tiles, off = take(g, off, 4 * P, torch.int32, (P,))
    R, b, off = fwd["rendered"], fwd["binning"], 0
    point_list, off = take(b, off, 4 * R, torch.int32, (R,))
    _, off = take(b, off, 4 * R, torch.int32, (R,))
    keys, off = take(b, off, 8 * R, torch.int64, (R,))
    im, N, off = fwd["img"], W * H, 0
    T = ((W + 15) // 16) * ((H + 15) // 16)
    accum, off = take(im, off, 4 * N, torch.float32, (N,))
    ncon, off = take(im, off, 4 * N, torch.int32, (N,))
    ranges, off = take(im, off, 8 * N, torch.int32, (N, 2))
    return dict(depths=depths, clamped=clamped, means2D=means2D, cov3D=cov3D, conic_opacity=conic_opacity, rgb=rgb,
                tiles_touched=tiles, point_offsets=torch.cumsum(tiles, 0).to(torch.int32), point_list=point_list,
                keys=keys, accum=accum, n_contrib=ncon, ranges=ranges[:T])


def new_gauss_state(fwd, P, W, H):
    import ctypes as C
    import b200gs as _b
    dev = fwd["geom"].device
    f = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)  # noqa: E731
    i = lambda *s: torch.zeros(s, dtype=torch.int32, device=dev)  # noqa: E731
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    depths, means2D, cov3D, co, rgb = f(P), f(P, 2), f(P, 6), f(P, 4), f(P, 3)
    tiles, offs = i(P), i(P)
    clamped = torch.zeros((P, 3), dtype=torch.uint8, device=dev)
    _b.check(_b.gauss_inspect_geom(fwd["geom"].data_ptr(), P, depths.data_ptr(), means2D.data_ptr(), cov3D.data_ptr(),
                                   co.data_ptr(), rgb.data_ptr(), tiles.data_ptr(), offs.data_ptr(), clamped.data_ptr(), st))
    R = fwd["rendered"]
    keys, pl = torch.zeros((R,), dtype=torch.int64, device=dev), i(R)
    if R:
        _b.check(_b.inspect_binning(fwd["binning"].data_ptr(), R, depths.data_ptr(), keys.data_ptr(), pl.data_ptr(), st))
    N, T = W * H, ((W + 15) // 16) * ((H + 15) // 16)
    accum, ncon, ranges = f(N), i(N), i(T, 2)
    _b.check(_b.gauss_inspect_image(fwd["img"].data_ptr(), W, H, accum.data_ptr(), ncon.data_ptr(), ranges.data_ptr(), st))
    return dict(depths=depths, clamped=clamped, means2D=means2D, cov3D=cov3D, conic_opacity=co, rgb=rgb,
                tiles_touched=tiles, point_offsets=offs, point_list=pl, keys=keys, accum=accum, n_contrib=ncon,
                ranges=ranges)
