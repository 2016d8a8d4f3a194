#!/usr/bin/env python
"""Generate tests/golden/gauss_tiny_ref.npz: outputs of the UNMODIFIED reference 3DGS extension
(oracle/_ref/ref_gauss_C, built by oracle/build_ref.py from
/root/reference/submodules/diff-gaussian-rasterization) on the deterministic `tiny` scene of
tests/scenes.py (third scale and per-Gaussian normals as in tests/harness.py:gauss_scene).
CUDA-only, so this runs on the GPU box:

    gpurun -- 'python tests/golden/make_golden_gauss.py'     # writes gpurun_out/gauss_tiny_ref.npz
    cp gpurun_out/gauss_tiny_ref.npz tests/golden/

The fixture pins the CPU oracle (tests/test_oracle_cpu.py::gauss, no GPU needed).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import harness as hz  # noqa: E402
import scenes  # noqa: E402


def main(name="tiny", bg=(0.1, 0.2, 0.3)):
    refC = hz.ref_module("ref_gauss_C")
    assert refC is not None, "oracle/_ref/ref_gauss_C.so missing"
    sc_np, cam_np = scenes.build(name)
    sc, cam = hz.gauss_scene(hz.to_dev(sc_np)), hz.to_dev(cam_np)
    W, H, P = cam_np["width"], cam_np["height"], sc_np["means3D"].shape[0]
    grads = hz.to_dev(hz.gauss_grads(W, H))
    out = dict(bg=np.float32(bg), scales3=sc["scales3"].cpu().numpy(), normals=sc["normals"].cpu().numpy())
    for aa in (False, True):
        f = hz.gauss_forward(refC, sc, cam, bg=bg, antialiasing=aa)
        b = hz.gauss_backward(refC, sc, cam, f, grads)
        torch.cuda.synchronize()
        st = hz.ref_gauss_state(f, P, W, H)
        tag = "aa_" if aa else ""
        out[tag + "rendered"] = np.int64(f["rendered"])
        for k in ("color", "invdepth", "normal_map", "refl_map", "radii"):
            out[tag + k] = f[k].cpu().numpy()
        for k in ("tiles_touched", "keys", "point_list", "ranges", "n_contrib"):
            out[tag + "state_" + k] = st[k].cpu().numpy()
        for k, v in b.items():
            out[tag + "grad_" + k] = v.cpu().numpy()
    dst = os.path.join(hz.ROOT, "gpurun_out", f"gauss_{name}_ref.npz")
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    np.savez_compressed(dst, **out)
    print("wrote", dst, os.path.getsize(dst), "bytes")


if __name__ == "__main__":
    main()
