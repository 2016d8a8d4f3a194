"""Dev harness (GPU): gradient of the DISTORTION plane alone (others[6]) -- this repo vs the reference extension vs the
fp64 CPU oracle, on small scenes.  Which side is closer to the true derivative?"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import numpy as np
import torch
import harness as hz
import scenes
from oracle import surfel_oracle as so

for name in sys.argv[1:] or ["tiny", "small"]:
    sc_np, cam_np = scenes.build(name)
    sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
    W, H = cam_np["width"], cam_np["height"]
    for plane in (6, 0, 5):
        g = {k: np.zeros_like(v) for k, v in scenes.upstream_grads(W, H).items()}
        g["others"][plane] = 1.0 / (W * H)          # d mean(plane)
        gd = hz.to_dev(g)
        newC, refC = hz.new_surfel_C(), hz.ref_module("ref_surfel_C")
        fn = hz.surfel_forward(newC, sc, cam); bn = hz.surfel_backward(newC, sc, cam, fn, gd)
        fr = hz.surfel_forward(refC, sc, cam); br = hz.surfel_backward(refC, sc, cam, fr, gd)
        _, go = so.forward_backward(sc_np, cam_np, g, dtype=torch.float64)
        row = {}
        for k in ("means3D", "opacity", "scales", "rotations"):
            o = torch.from_numpy(np.asarray(go[k], np.float64)).reshape(bn[k].shape)
            row[k] = (round(hz.rel_l2(bn[k].cpu().double(), o), 6), round(hz.rel_l2(br[k].cpu().double(), o), 6),
                      round(hz.rel_l2(bn[k], br[k]), 6))
        print(name, "plane", plane, "(new vs fp64 oracle, ref vs fp64 oracle, new vs ref):", row, flush=True)
