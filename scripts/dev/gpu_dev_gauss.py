#!/usr/bin/env python
"""Developer tool (GPU box): 3DGS path, this repo vs the reference extension (oracle/_ref/ref_gauss_C).
  python tests/gpu_dev_gauss.py tiny small c1 [--aa]"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import sys

import torch

import harness as hz
import scenes
from gpu_dev_surfel import timed


def compare(name, aa):
    sc_np, cam_np = scenes.build(name)
    P, W, H = sc_np["means3D"].shape[0], cam_np["width"], cam_np["height"]
    sc, cam = hz.gauss_scene(hz.to_dev(sc_np)), hz.to_dev(cam_np)
    grads = hz.to_dev(hz.gauss_grads(W, H))
    newC, refC = hz.new_gauss_C(), hz.ref_module("ref_gauss_C")
    fn = hz.gauss_forward(newC, sc, cam, antialiasing=aa)
    bn = hz.gauss_backward(newC, sc, cam, fn, grads)
    torch.cuda.synchronize()
    print(f"[{name}] aa={aa} P={P} {W}x{H} R={fn['rendered']} V={int((fn['radii'] > 0).sum())}", flush=True)
    if refC is not None:
        fr = hz.gauss_forward(refC, sc, cam, antialiasing=aa)
        br = hz.gauss_backward(refC, sc, cam, fr, grads)
        br2 = hz.gauss_backward(refC, sc, cam, fr, grads)
        torch.cuda.synchronize()
        sn, sr = hz.new_gauss_state(fn, P, W, H), hz.ref_gauss_state(fr, P, W, H)
        vis = fr["radii"] > 0
        bits = lambda a, b: int((a.view(torch.int32) != b.view(torch.int32)).sum())  # noqa: E731
        print("   rendered", fn["rendered"], fr["rendered"], "radii_mismatch", int((fr["radii"] != fn["radii"]).sum()),
              "tiles_mismatch", int((sr["tiles_touched"] != sn["tiles_touched"]).sum()))
        for k in ("depths", "means2D", "cov3D", "conic_opacity", "rgb"):
            print(f"   {k}_bits_mismatch", bits(sr[k][vis], sn[k][vis]), "maxabs", hz.max_abs(sr[k][vis], sn[k][vis]))
        if fr["rendered"] == fn["rendered"]:
            print("   keys_mismatch", int((sr["keys"] != sn["keys"]).sum()), "point_list_mismatch",
                  int((sr["point_list"] != sn["point_list"]).sum()))
        print("   ranges_mismatch", int((sr["ranges"] != sn["ranges"]).sum()), "n_contrib_mismatch",
              int((sr["n_contrib"] != sn["n_contrib"]).sum()), "accum_bits", bits(sr["accum"], sn["accum"]))
        for k in ("color", "invdepth", "normal_map", "refl_map"):
            print(f"   {k}_bits_mismatch", bits(fr[k], fn[k]), "maxabs", hz.max_abs(fr[k], fn[k]))
        for k in bn:
            print(f"   grad_{k} rel_l2 {hz.rel_l2(bn[k], br[k]):.3e}  refnoise {hz.rel_l2(br2[k], br[k]):.3e}  "
                  f"close {hz.grad_close_fraction(bn[k], br[k]):.5f}")

    def new_fb():
        f = hz.gauss_forward(newC, sc, cam, antialiasing=aa)
        hz.gauss_backward(newC, sc, cam, f, grads)
    res = {"new_fwd_ms": timed(lambda: hz.gauss_forward(newC, sc, cam, antialiasing=aa), 5), "new_fwdbwd_ms": timed(new_fb, 5)}
    if refC is not None:
        def ref_fb():
            f = hz.gauss_forward(refC, sc, cam, antialiasing=aa)
            hz.gauss_backward(refC, sc, cam, f, grads)
        res.update(ref_fwd_ms=timed(lambda: hz.gauss_forward(refC, sc, cam, antialiasing=aa), 5), ref_fwdbwd_ms=timed(ref_fb, 5))
    print("  ", res, flush=True)


if __name__ == "__main__":
    names = [a for a in sys.argv[1:] if not a.startswith("--")] or ["tiny"]
    for n in names:
        compare(n, "--aa" in sys.argv)
