"""Dev harness (GPU): host enqueue time vs device time of one fwd+bwd on a small config."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import sys, time
import torch
import harness as hz
import scenes

name = sys.argv[1] if len(sys.argv) > 1 else "c1"
newC = hz.new_surfel_C()
sc_np, cam_np = scenes.build(name)
sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
grads = hz.to_dev(scenes.upstream_grads(cam_np["width"], cam_np["height"]))
import b200gs


def step():
    f = hz.surfel_forward(newC, sc, cam)
    return hz.surfel_backward(newC, sc, cam, f, grads)


for _ in range(5):
    step()
torch.cuda.synchronize()
n0 = b200gs.launch_count()
t0 = time.perf_counter()
N = 50
for _ in range(N):
    step()
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"{name}: host enqueue {1e3*(t1-t0)/N:.3f} ms/step, wall {1e3*(t2-t0)/N:.3f} ms/step, own launches/step {(b200gs.launch_count()-n0)/N:.0f}")
b200gs.timing_enable(1)
for _ in range(10):
    step()
torch.cuda.synchronize()
st = b200gs.timing_read()
b200gs.timing_enable(0)
print({k: round(v[0] / max(v[1], 1), 4) for k, v in st.items() if v[1]}, "sum", round(sum(v[0] / max(v[1], 1) for v in st.values() if v[1]), 3))
# forward only host time split
t0 = time.perf_counter()
for _ in range(N):
    f = hz.surfel_forward(newC, sc, cam)
t1 = time.perf_counter()
torch.cuda.synchronize()
print(f"forward only: host {1e3*(t1-t0)/N:.3f} ms")
