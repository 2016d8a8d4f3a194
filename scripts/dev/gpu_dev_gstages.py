"""Dev harness (GPU): per-stage device time of one 3DGS fwd+bwd on a config."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import sys
import torch
import harness as hz
import scenes
import b200gs

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
newC = hz.new_gauss_C()
sc_np, cam_np = scenes.build(name)
sc, cam = hz.gauss_scene(hz.to_dev(sc_np)), hz.to_dev(cam_np)
grads = hz.to_dev(hz.gauss_grads(cam_np["width"], cam_np["height"]))


def step():
    f = hz.gauss_forward(newC, sc, cam)
    return hz.gauss_backward(newC, sc, cam, f, grads)


for mode in (0, 1):
    b200gs.set_exact_lists(mode)
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    b200gs.timing_enable(1)
    for _ in range(10):
        step()
    torch.cuda.synchronize()
    st = b200gs.timing_read()
    b200gs.timing_enable(0)
    print("exact" if mode else "culled", {k: round(v[0] / max(v[1], 1), 4) for k, v in st.items() if v[1]},
          "sum", round(sum(v[0] / max(v[1], 1) for v in st.values() if v[1]), 3))
