#!/usr/bin/env python
"""Build recipe for oracle/_ref/: the UNMODIFIED reference CUDA extensions.

TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is imported by the product
package; only tests/, __graft_entry__.smoke() and bench.py's reference /
cpu_baseline legs may touch it.

The reference extensions are compiled straight from the sources where
they lie under /root/reference/submodules (no source is copied into this repo);
the only outputs are the shared objects and ninja scratch under
oracle/_ref/ (git-ignored, NOT gpurun-ignored, so the .so files travel to the
GPU box, where /root/reference does not exist).

  ref_surfel_C   <- diff-surfel-rasterization   (S/setup.py:21-30 source list)
  ref_gauss_C    <- diff-gaussian-rasterization (G/setup.py:21-29)
  ref_cubemap_C  <- cubemapencoder              (C/setup.py:19-27, -use_fast_math)
  ref_ssim_C     <- fused-ssim                  (submodules/fused-ssim/setup.py: ssim.cu, ext.cpp)

`-include cstdint` is required on gcc 13: S|G/cuda_rasterizer/rasterizer_impl.h
uses std::uintptr_t / uint32_t without including <cstdint>.
"""
import os
import sys

REF = os.environ.get("B200GS_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo"]

TARGETS = {
    "ref_surfel_C": dict(
        root="submodules/diff-surfel-rasterization",
        sources=["cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu",
                 "cuda_rasterizer/backward.cu", "rasterize_points.cu", "ext.cpp"],
        include=["third_party/glm"],
        nvcc=["-include", "cstdint"],
        cxx=[],
    ),
    "ref_gauss_C": dict(
        root="submodules/diff-gaussian-rasterization",
        sources=["cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu",
                 "cuda_rasterizer/backward.cu", "rasterize_points.cu", "ext.cpp"],
        include=["third_party/glm"],
        nvcc=["-include", "cstdint"],
        cxx=[],
    ),
    "ref_cubemap_C": dict(
        root="submodules/cubemapencoder",
        sources=["src/cubemapencoder.cu", "src/bindings.cpp"],
        include=[],
        nvcc=["-O3", "-std=c++17", "-DCUDA_HAS_FP16=1", "-D__CUDA_NO_HALF_OPERATORS__",
              "-D__CUDA_NO_HALF_CONVERSIONS__", "-D__CUDA_NO_HALF2_OPERATORS__", "-use_fast_math"],
        cxx=["-O3", "-std=c++17"],
    ),
    "ref_ssim_C": dict(
        root="submodules/fused-ssim",
        sources=["ssim.cu", "ext.cpp"],
        include=[],
        nvcc=[],
        cxx=[],
    ),
}


# The reference's Python around the extensions, copied VERBATIM (byte for byte, checked below) next to the .so
# files so that tests can run the reference's own `gaussian_renderer.render()` / `GaussianModel` on the GPU box,
# where /root/reference does not exist.  Same treatment as the compiled extensions: outputs only under
# oracle/_ref/ (git-ignored, not gpurun-ignored); nothing is copied into the tracked tree.
#   oracle/_ref/py/      the fork's glue: gaussian_renderer, scene.{cameras,gaussian_model}, utils.*, arguments
#   oracle/_ref/py_ext/  the Python halves of the reference EXTENSION packages (their `_C` / `_cubemapencoder`
#                        compiled halves are the ref_*_C.so above, put into sys.modules by tests/refpy.py)
PY_GLUE = ["gaussian_renderer/__init__.py", "scene/cameras.py", "scene/gaussian_model.py", "arguments/__init__.py",
           "utils/general_utils.py", "utils/point_utils.py", "utils/graphics_utils.py", "utils/sh_utils.py",
           "utils/loss_utils.py", "utils/system_utils.py", "utils/image_utils.py"]
PY_EXT = {"submodules/diff-surfel-rasterization/diff_surfel_rasterization/__init__.py": "diff_surfel_rasterization/__init__.py",
          "submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py": "diff_gaussian_rasterization/__init__.py",
          "submodules/cubemapencoder/cubemapencoder/__init__.py": "cubemapencoder/__init__.py",
          "submodules/cubemapencoder/cubemapencoder/cubemap_encoder.py": "cubemapencoder/cubemap_encoder.py",
          "submodules/fused-ssim/fused_ssim/__init__.py": "fused_ssim/__init__.py"}


def copy_python():
    """-> list of files now present under oracle/_ref/py*.  No-op (returns what is there) without /root/reference."""
    import filecmp
    import shutil
    pairs = [(os.path.join(REF, f), os.path.join(OUT, "py", f)) for f in PY_GLUE]
    pairs += [(os.path.join(REF, a), os.path.join(OUT, "py_ext", b)) for a, b in PY_EXT.items()]
    have = []
    for src, dst in pairs:
        if os.path.isdir(REF):
            if not os.path.exists(src):
                print(f"[build_ref] missing in the reference: {src}", file=sys.stderr)
                continue
            os.makedirs(os.path.dirname(dst), exist_ok=True)
            if not (os.path.exists(dst) and filecmp.cmp(src, dst, shallow=False)):
                shutil.copyfile(src, dst)
            assert filecmp.cmp(src, dst, shallow=False), dst
        if os.path.exists(dst):
            have.append(os.path.relpath(dst, OUT))
    return have


def built(name):
    return os.path.exists(os.path.join(OUT, name, name + ".so"))


def build(names=None, verbose=False):
    """Build the requested reference extensions (all by default).  Returns the
    list of names that are available afterwards.  A missing /root/reference is
    not an error: prebuilt .so files (if any) are used as they are."""
    names = list(names or TARGETS)
    copy_python()
    if not os.path.isdir(REF):
        return [n for n in names if built(n)]
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0a")
    from torch.utils.cpp_extension import load
    ok = []
    for name in names:
        t = TARGETS[name]
        root = os.path.join(REF, t["root"])
        bdir = os.path.join(OUT, name)
        os.makedirs(bdir, exist_ok=True)
        try:
            load(name=name,
                 sources=[os.path.join(root, s) for s in t["sources"]],
                 extra_include_paths=[os.path.join(root, i) for i in t["include"]],
                 extra_cuda_cflags=ARCH + t["nvcc"],
                 extra_cflags=t["cxx"],
                 build_directory=bdir,
                 is_python_module=False,   # build only; importing needs no GPU but keep it lazy
                 verbose=verbose)
            ok.append(name)
        except Exception as ex:  # noqa: BLE001 - report and carry on with the others
            print(f"[build_ref] {name}: FAILED: {ex}", file=sys.stderr)
    return ok


def load_ref(name):
    """Import a prebuilt reference extension from oracle/_ref (GPU box or here)."""
    import importlib.util
    import torch  # noqa: F401  (libtorch symbols must be resident first)
    path = os.path.join(OUT, name, name + ".so")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    got = build(sys.argv[1:] or None, verbose=True)
    print("built:", got)
