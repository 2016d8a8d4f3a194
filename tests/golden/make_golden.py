#!/usr/bin/env python
"""Generate tests/golden/surfel_tiny_ref.npz: outputs of the UNMODIFIED reference CUDA
extension (oracle/_ref/ref_surfel_C, built by oracle/build_ref.py from
/root/reference/submodules/diff-surfel-rasterization) on the deterministic `tiny` scene of
tests/scenes.py.  The reference is CUDA-only, so this runs on the GPU box:

    gpurun -- 'python tests/golden/make_golden.py'     # writes gpurun_out/surfel_tiny_ref.npz
    cp gpurun_out/surfel_tiny_ref.npz tests/golden/

The fixture pins the CPU oracle (tests/test_oracle_cpu.py, no GPU needed) and the sm_100a
kernels (tests/test_surfel_gpu.py) to the reference's own results.
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
    refC = hz.ref_module("ref_surfel_C")
    assert refC is not None, "oracle/_ref/ref_surfel_C.so missing"
    sc_np, cam_np = scenes.build(name)
    sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
    W, H, P = cam_np["width"], cam_np["height"], sc_np["means3D"].shape[0]
    grads = hz.to_dev(scenes.upstream_grads(W, H))
    f = hz.surfel_forward(refC, sc, cam, bg=bg)
    b = hz.surfel_backward(refC, sc, cam, f, grads)
    torch.cuda.synchronize()
    st = hz.ref_surfel_state(f, P, W, H)
    out = dict(bg=np.float32(bg), rendered=np.int64(f["rendered"]))
    for k in ("color", "others", "refl_map", "radii", "weights"):
        out[k] = f[k].cpu().numpy()
    for k in ("tiles_touched", "keys", "point_list", "ranges", "n_contrib", "accum", "depths", "means2D", "transMat",
              "normal_opacity", "rgb", "clamped"):
        out["state_" + k] = st[k].cpu().numpy()
    for k, v in b.items():
        out["grad_" + k] = v.cpu().numpy()
    dst = os.path.join(hz.ROOT, "gpurun_out", f"surfel_{name}_ref.npz")
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    np.savez_compressed(dst, **out)
    print("wrote", dst, os.path.getsize(dst), "bytes")


if __name__ == "__main__":
    main()
