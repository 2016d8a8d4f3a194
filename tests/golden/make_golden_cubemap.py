#!/usr/bin/env python
"""Generate tests/golden/cubemap_ref.npz: outputs of the UNMODIFIED reference cubemap extension
(oracle/_ref/ref_cubemap_C, built by oracle/build_ref.py with the reference's own flags incl.
-use_fast_math) for the three lookup modes on seeded inputs.  CUDA-only -> run on the GPU box:
    gpurun -- 'python tests/golden/make_golden_cubemap.py' ; cp gpurun_out/cubemap_ref.npz tests/golden/
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import harness as hz  # noqa: E402


def case_inputs(B=4096, C=3, L=16, seed=5):
    rng = np.random.default_rng(seed)
    d = rng.normal(size=(B, 3)).astype(np.float32)
    d[:64] = np.sign(d[:64]) * (1.0 + 0.01 * rng.uniform(-1, 1, (64, 3)).astype(np.float32))   # near cube vertices
    d[64:256, 0] = np.sign(d[64:256, 0]) * np.abs(d[64:256, 1])                                  # near cube edges
    d[300] = 0.0                                                                                  # fail value path
    tex = rng.uniform(-0.5, 0.5, (6, C, L, L)).astype(np.float32)
    fail = rng.uniform(-1, 1, C).astype(np.float32)
    g = rng.normal(size=(C, B)).astype(np.float32)
    return d, tex, fail, g


def run(mod, d, tex, fail, g, interp, seamless):
    dev = "cuda"
    d_, tex_, fail_, g_ = (torch.from_numpy(a).to(dev) for a in (d, tex, fail, g))
    B, C, L = d.shape[0], tex.shape[1], tex.shape[2]
    out = torch.empty((C, B), device=dev)
    mod.cubemap_encode_forward(d_, tex_, fail_, out, interp, seamless, B, C, L)
    gt, gi, gf = torch.zeros_like(tex_), torch.empty_like(d_), torch.zeros(C, device=dev)
    mod.cubemap_encode_backward(g_, d_, tex_, gt, gi, gf, interp, seamless, B, C, L)
    torch.cuda.synchronize()
    return out, gi, gt, gf


def main():
    ref = hz.ref_module("ref_cubemap_C")
    assert ref is not None
    d, tex, fail, g = case_inputs()
    out = dict(inputs=d, cubemap=tex, fail=fail, grad_outputs=g)
    for interp, seamless in ((0, 0), (1, 0), (1, 1)):
        o, gi, gt, gf = run(ref, d, tex, fail, g, interp, seamless)
        tag = f"m{interp}{seamless}_"
        out.update({tag + "out": o.cpu().numpy(), tag + "gin": gi.cpu().numpy(), tag + "gtex": gt.cpu().numpy(),
                    tag + "gfail": gf.cpu().numpy()})
    dst = os.path.join(hz.ROOT, "gpurun_out", "cubemap_ref.npz")
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    np.savez_compressed(dst, **out)
    print("wrote", dst, os.path.getsize(dst))


if __name__ == "__main__":
    main()
