"""Test infrastructure: run the reference's OWN, UNMODIFIED Python (gaussian_renderer.render / render_fast,
scene.gaussian_model.GaussianModel, scene.cameras.Camera) on the GPU box, twice:

  arm "ref"  with the reference's compiled extensions behind its imports (oracle/_ref/ref_surfel_C.so as
             diff_surfel_rasterization._C, ref_cubemap_C.so as _cubemapencoder, ref_ssim_C.so as fused_ssim_cuda)
  arm "new"  with THIS repo's drop-in packages (gaussian-splatting-reflection_b200/python) behind the same imports

north_star: "so gaussian_renderer/__init__.py, train.py and render.py call it unchanged".  The Python files are
byte-for-byte copies made by oracle/build_ref.py:copy_python() into the git-ignored oracle/_ref/py*/ (they travel to
the GPU box like the .so files; /root/reference does not exist there).  Third-party modules the fork imports at
module scope but the rendering path never calls (plyfile, simple_knn, kornia, matplotlib -- SURVEY.md section 7 step 2)
are stubbed in sys.modules for the duration of the import.

Each arm gets its own module universe: the packages are imported under a clean sys.modules / sys.path and the
previous state is restored afterwards, so both arms coexist in one process (the functions keep their own globals).
"""
import importlib
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_OUT = os.path.join(ROOT, "oracle", "_ref")
PY_GLUE = os.path.join(REF_OUT, "py")
PY_EXT = os.path.join(REF_OUT, "py_ext")
PKG_PY = os.path.join(ROOT, "gaussian-splatting-reflection_b200", "python")

_OWNED = ("diff_surfel_rasterization", "diff_gaussian_rasterization", "cubemapencoder", "_cubemapencoder", "fused_ssim",
          "fused_ssim_cuda", "gaussian_renderer", "scene", "utils", "arguments")
_STUBS = ("plyfile", "simple_knn", "simple_knn._C", "kornia", "kornia.filters", "matplotlib", "matplotlib.pyplot")


def available():
    return os.path.exists(os.path.join(PY_GLUE, "gaussian_renderer", "__init__.py"))


def _stub(name):
    m = types.ModuleType(name)
    m.__dict__["__stub__"] = True

    def _missing(*a, **k):
        raise RuntimeError(f"{name} is stubbed in this test harness (not on the rendering path)")
    for attr in ("PlyData", "PlyElement", "distCUDA2", "UnsharpMask", "figure", "close"):
        setattr(m, attr, _missing)
    return m


def _owned(name):
    return name in _OWNED or name.split(".")[0] in _OWNED


_ARMS = {}


def load_arm(kind):
    """-> namespace with render, render_fast, GaussianModel, Camera, MiniCam, PipelineParams-like `pipe()`,
    OptimizationParams `opt()`, the arm's `diff_surfel_rasterization` and `cubemapencoder` modules."""
    if kind in _ARMS:
        return _ARMS[kind]
    assert kind in ("ref", "new")
    if not available():
        return None
    saved_modules = {k: v for k, v in sys.modules.items() if _owned(k) or k in _STUBS}
    saved_path = list(sys.path)
    for k in saved_modules:
        del sys.modules[k]
    try:
        for s in _STUBS:
            try:
                importlib.import_module(s)
            except Exception:  # noqa: BLE001
                sys.modules[s] = _stub(s)
        sys.modules["simple_knn"].__path__ = []          # packages, so `from simple_knn._C import ...` resolves
        sys.modules["kornia"].__path__ = []
        sys.modules["matplotlib"].__path__ = []
        if kind == "ref":
            sys.path.insert(0, ROOT)
            from oracle import build_ref
            ext = {n: build_ref.load_ref(n) for n in ("ref_surfel_C", "ref_gauss_C", "ref_cubemap_C", "ref_ssim_C")}
            if ext["ref_surfel_C"] is None or ext["ref_cubemap_C"] is None:
                return None
            sys.modules["diff_surfel_rasterization._C"] = ext["ref_surfel_C"]
            if ext["ref_gauss_C"] is not None:
                sys.modules["diff_gaussian_rasterization._C"] = ext["ref_gauss_C"]
            sys.modules["_cubemapencoder"] = ext["ref_cubemap_C"]
            if ext["ref_ssim_C"] is not None:
                sys.modules["fused_ssim_cuda"] = ext["ref_ssim_C"]
            sys.path[:0] = [PY_EXT, PY_GLUE]
        else:
            sys.path[:0] = [PKG_PY, PY_GLUE]
        ns = types.SimpleNamespace(kind=kind)
        ns.diff_surfel_rasterization = importlib.import_module("diff_surfel_rasterization")
        ns.cubemapencoder = importlib.import_module("cubemapencoder")
        want = PY_EXT if kind == "ref" else PKG_PY
        assert os.path.abspath(ns.diff_surfel_rasterization.__file__).startswith(want), ns.diff_surfel_rasterization.__file__
        assert os.path.abspath(ns.cubemapencoder.__file__).startswith(want), ns.cubemapencoder.__file__
        gr = importlib.import_module("gaussian_renderer")
        assert os.path.abspath(gr.__file__).startswith(PY_GLUE)
        assert gr.GaussianRasterizer is ns.diff_surfel_rasterization.GaussianRasterizer
        gm = importlib.import_module("scene.gaussian_model")
        assert gm.CubemapEncoder is ns.cubemapencoder.CubemapEncoder
        cams = importlib.import_module("scene.cameras")
        args = importlib.import_module("arguments")
        ns.gaussian_renderer, ns.gaussian_model, ns.cameras, ns.arguments = gr, gm, cams, args
        ns.general_utils = importlib.import_module("utils.general_utils")
        ns.loss_utils = importlib.import_module("utils.loss_utils")
        ns.render, ns.render_fast = gr.render, gr.render_fast
        ns.GaussianModel, ns.Camera, ns.MiniCam = gm.GaussianModel, cams.Camera, cams.MiniCam
        try:
            ns.fused_ssim = importlib.import_module("fused_ssim")
        except Exception:  # noqa: BLE001
            ns.fused_ssim = None

        def _params(cls):
            from argparse import ArgumentParser
            return cls(ArgumentParser())
        ns.pipe = lambda: _params(args.PipelineParams)
        ns.opt = lambda: _params(args.OptimizationParams)
        _ARMS[kind] = ns
        return ns
    finally:
        for k in [k for k in sys.modules if _owned(k) or k in _STUBS]:
            del sys.modules[k]
        sys.modules.update(saved_modules)
        sys.path[:] = saved_path


# ---------------------------------------------------------------- model / camera builders shared by the arms
def make_camera(arm, cam_np, image=None, uid=0):
    """scene.cameras.Camera from a tests/scenes.py camera: R is stored transposed (COLMAP convention of the fork:
    utils/graphics_utils.py:38-49), HWK carries the intrinsics the reflection pass reads."""
    import numpy as np
    W, H = cam_np["width"], cam_np["height"]
    w2c = np.asarray(cam_np["w2c"], np.float64)
    R, T = np.ascontiguousarray(w2c[:3, :3].T), np.ascontiguousarray(w2c[:3, 3])
    # float64 like a dataset reader's intrinsics: getProjectionMatrixCorrect (utils/graphics_utils.py:72-91) assigns
    # expressions of K's entries into a torch tensor, which this torch accepts for numpy.float64 (a Python float) only
    K = np.array([[W / (2 * cam_np["tanfovx"]), 0, W / 2], [0, H / (2 * cam_np["tanfovy"]), H / 2], [0, 0, 1]], np.float64)
    return arm.Camera(resolution=(W, H), colmap_id=uid, R=R, T=T, FoVx=cam_np["fovx"], FoVy=cam_np["fovy"], image=image,
                      gt_alpha_mask=None, image_name=f"view{uid}", uid=uid, HWK=(H, W, K))


def make_model(arm, sc_np, env_tex=None, sh_degree=3, cubemap_resol=16):
    """GaussianModel of the arm with the RAW parameters that activate to the scene's values (inverse sigmoid / log,
    scene/gaussian_model.py:34-52), set the way create_from_pcd / load_ply leave them (:168-187)."""
    import torch
    from torch import nn
    gm = arm.GaussianModel(sh_degree, init_cubemap_resol=cubemap_resol)
    dev = "cuda"
    t = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    inv_sig = lambda x: torch.log(x / (1 - x))  # noqa: E731
    shs = t(sc_np["shs"])
    gm._xyz = nn.Parameter(t(sc_np["means3D"]).clone().requires_grad_(True))
    gm._features_dc = nn.Parameter(shs[:, :1, :].contiguous().requires_grad_(True))
    gm._features_rest = nn.Parameter(shs[:, 1:, :].contiguous().requires_grad_(True))
    gm._scaling = nn.Parameter(torch.log(t(sc_np["scales"])).requires_grad_(True))
    gm._rotation = nn.Parameter(t(sc_np["rotations"]).clone().requires_grad_(True))
    gm._opacity = nn.Parameter(inv_sig(t(sc_np["opacities"])).requires_grad_(True))
    gm._refl_strength = nn.Parameter(inv_sig(t(sc_np["refl"]).clamp(1e-4, 1 - 1e-4)).requires_grad_(True))
    gm.max_radii2D = torch.zeros((gm._xyz.shape[0],), device=dev)
    gm.active_sh_degree = sh_degree
    env = arm.cubemapencoder.CubemapEncoder(output_dim=3, resolution=cubemap_resol).cuda()
    if env_tex is not None:
        with torch.no_grad():
            env.params["Cubemap_texture"].copy_(env_tex.to(dev))
    gm.env_map = env
    return gm
