#!/usr/bin/env python
"""Developer tool (GPU box): run this repo's surfel path and the reference extension
(oracle/_ref) on identical synthetic scenes, print parity statistics and timings.

  python tests/gpu_dev_surfel.py small c1 c3 [--no-ref] [--iters 5]
"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import json
import sys
import time

import torch

import harness as hz
import scenes


def timed(fn, iters, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def compare(name, iters, use_ref=True):
    cfg = scenes.CONFIGS[name]
    sc_np, cam_np = scenes.build(name)
    P, W, H = sc_np["means3D"].shape[0], cam_np["width"], cam_np["height"]
    sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
    sc["refl"] = sc["refl"].contiguous()
    grads = hz.to_dev(scenes.upstream_grads(W, H))
    newC = hz.new_surfel_C()
    refC = hz.ref_module("ref_surfel_C") if use_ref else None
    res = dict(config=name, P=P, W=W, H=H)

    fn = hz.surfel_forward(newC, sc, cam)
    torch.cuda.synchronize()
    bn = hz.surfel_backward(newC, sc, cam, fn, grads)
    torch.cuda.synchronize()
    R = fn["rendered"]
    res.update(R=R, V=int((fn["radii"] > 0).sum()))
    print(f"[{name}] P={P} {W}x{H} R={R} V={res['V']}", flush=True)

    if refC is not None:
        fr = hz.surfel_forward(refC, sc, cam)
        br = hz.surfel_backward(refC, sc, cam, fr, grads)
        torch.cuda.synchronize()
        sn, sr = hz.new_surfel_state(fn, P, W, H), hz.ref_surfel_state(fr, P, W, H)
        vis = fr["radii"] > 0
        par = {}
        par["rendered_equal"] = fr["rendered"] == R
        par["radii_mismatch"] = int((fr["radii"] != fn["radii"]).sum())
        par["tiles_mismatch"] = int((sr["tiles_touched"] != sn["tiles_touched"]).sum())
        par["depth_bits_mismatch"] = int((sr["depths"].view(torch.int32)[vis] != sn["depths"].view(torch.int32)[vis]).sum())
        par["transMat_bits_mismatch"] = int((sr["transMat"].view(torch.int32)[vis] != sn["transMat"].view(torch.int32)[vis]).sum())
        par["means2D_bits_mismatch"] = int((sr["means2D"].view(torch.int32)[vis] != sn["means2D"].view(torch.int32)[vis]).sum())
        par["normal_opacity_maxabs"] = hz.max_abs(sr["normal_opacity"][vis], sn["normal_opacity"][vis])
        par["rgb_maxabs"] = hz.max_abs(sr["rgb"][vis], sn["rgb"][vis])
        par["clamped_mismatch"] = int((sr["clamped"][vis] != sn["clamped"][vis]).sum())
        if fr["rendered"] == R:
            par["keys_mismatch"] = int((sr["keys"] != sn["keys"]).sum())
            par["point_list_mismatch"] = int((sr["point_list"] != sn["point_list"]).sum())
        par["ranges_mismatch"] = int((sr["ranges"] != sn["ranges"]).sum())
        par["n_contrib_mismatch"] = int((sr["n_contrib"] != sn["n_contrib"]).sum())
        par["accum_maxabs"] = hz.max_abs(sr["accum"], sn["accum"])
        for k in ("color", "others", "refl_map"):
            par[f"{k}_maxabs"] = hz.max_abs(fr[k], fn[k])
            par[f"{k}_bits_mismatch"] = int((fr[k].view(torch.int32) != fn[k].view(torch.int32)).sum())
        par["weights_ref_gt_new"] = int((fr["weights"] > fn["weights"]).sum())
        par["weights_neq"] = int((fr["weights"] != fn["weights"]).sum())
        # run-to-run noise of the reference backward (atomics) as the yardstick
        br2 = hz.surfel_backward(refC, sc, cam, fr, grads)
        for k in bn:
            par[f"grad_{k}_rel_l2"] = hz.rel_l2(bn[k], br[k])
            par[f"grad_{k}_close_1e-4"] = hz.grad_close_fraction(bn[k], br[k])
            par[f"grad_{k}_refnoise_rel_l2"] = hz.rel_l2(br2[k], br[k])
            par[f"grad_{k}_refnoise_close_1e-4"] = hz.grad_close_fraction(br2[k], br[k])
        res["parity"] = par
        for k, v in par.items():
            print(f"   {k}: {v}")

    def new_fwd():
        return hz.surfel_forward(newC, sc, cam)

    def new_fb():
        f = hz.surfel_forward(newC, sc, cam)
        hz.surfel_backward(newC, sc, cam, f, grads)

    res["new_fwd_ms"] = timed(new_fwd, iters)
    res["new_fwdbwd_ms"] = timed(new_fb, iters)
    if refC is not None:
        def ref_fwd():
            return hz.surfel_forward(refC, sc, cam)

        def ref_fb():
            f = hz.surfel_forward(refC, sc, cam)
            hz.surfel_backward(refC, sc, cam, f, grads)
        res["ref_fwd_ms"] = timed(ref_fwd, iters)
        res["ref_fwdbwd_ms"] = timed(ref_fb, iters)
    print({k: v for k, v in res.items() if k.endswith("_ms")}, flush=True)
    return res


if __name__ == "__main__":
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    iters = 5
    if "--iters" in sys.argv:
        iters = int(sys.argv[sys.argv.index("--iters") + 1])
        args = [a for a in args if a != str(iters)]
    out = []
    for name in args or ["small"]:
        out.append(compare(name, iters, use_ref="--no-ref" not in sys.argv))
    import os
    os.makedirs(os.path.join(hz.ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(hz.ROOT, "gpurun_out", "dev_surfel.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("done", time.strftime("%H:%M:%S"))
