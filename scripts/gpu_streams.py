"""GPU experiment: a batch of independent views issued round-robin over K CUDA streams (one host thread) against the
same batch on one stream.  While the host waits for one stream's stage-1 plan, the other streams have a whole
stage 2 + backward queued, and the small latency-bound kernels of one view (scan, sort passes, ranges) run in the
shadow of another view's compositing kernels.
  python scripts/gpu_streams.py [config] [views]"""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import torch
import harness as hz  # noqa: E402 (sets sys.path)
import scenes

name = sys.argv[1] if len(sys.argv) > 1 else "bench"
V = int(sys.argv[2]) if len(sys.argv) > 2 else 16
sc_np, cam_np = scenes.build(name)
sc = hz.to_dev(sc_np)
W, H = cam_np["width"], cam_np["height"]
cams = [hz.to_dev(scenes.camera(name, i % 32, 32)) for i in range(V)]
grads = hz.to_dev(scenes.upstream_grads(W, H))
C = hz.new_surfel_C()


def batch(K):
    streams = [torch.cuda.Stream() for _ in range(K)] if K > 1 else [torch.cuda.current_stream()]
    main = torch.cuda.current_stream()
    if K > 1:
        for s in streams:
            s.wait_stream(main)
    for i, cam in enumerate(cams):
        with torch.cuda.stream(streams[i % K]):
            f = hz.surfel_forward(C, sc, cam)
            hz.surfel_backward(C, sc, cam, f, grads)
    if K > 1:
        for s in streams:
            main.wait_stream(s)


def timed(K, n=3, w=2):
    for _ in range(w):
        batch(K)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        batch(K)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n / V


out = {"config": name, "views": V}
for K in (1, 2, 3, 4):
    out[f"ms_per_view_{K}_streams"] = round(timed(K), 4)
print(json.dumps(out), flush=True)
