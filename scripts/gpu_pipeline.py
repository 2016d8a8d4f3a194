"""GPU experiment: stage 1 of the NEXT view (preprocess, count, depth order, scan, plan + the host sync) on a
high-priority stream while stage 2 + backward of the current view run on the main stream.
  python scripts/gpu_pipeline.py [config] [views]"""
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
dev = sc["means3D"].device
e = torch.empty((0,), dtype=torch.float32, device=dev)
mask = torch.ones((sc["means3D"].shape[0], 3), dtype=torch.bool, device=dev)
bg = torch.zeros(3, device=dev)
KEYS = ("rendered", "color", "others", "radii", "geom", "binning", "img", "refl_map", "weights")


def prepare(cam):
    return C.rasterize_gaussians_prepare(bg, sc["means3D"], mask, e, sc["refl"], sc["opacities"], sc["scales"], sc["rotations"],
                                         1.0, e, cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"],
                                         cam["height"], cam["width"], sc["shs"], 3, cam["campos"], False, False)


def finish(prep):
    r = dict(zip(KEYS, C.rasterize_gaussians_finish(prep)))
    r.update(bg=bg, sh_degree=3, scale_modifier=1.0, colors_precomp=None, transMat_precomp=None)
    return r


def plain():
    for cam in cams:
        f = hz.surfel_forward(C, sc, cam)
        hz.surfel_backward(C, sc, cam, f, grads)


def make_pipelined(priority, before_bwd=False):
    side = torch.cuda.Stream(priority=priority)

    def run():
        main = torch.cuda.current_stream()
        side.wait_stream(main)
        with torch.cuda.stream(side):
            prep = prepare(cams[0])
        for i, cam in enumerate(cams):
            f = finish(prep)                           # stage 2 of view i on the main stream (asynchronous)
            if before_bwd and i + 1 < len(cams):
                with torch.cuda.stream(side):          # stage 1 of view i+1 next to stage 2 of view i
                    prep = prepare(cams[i + 1])
            hz.surfel_backward(C, sc, cam, f, grads)   # backward of view i, queued behind it
            if not before_bwd and i + 1 < len(cams):
                with torch.cuda.stream(side):          # stage 1 of view i+1 next to them; the host waits for ITS end only
                    prep = prepare(cams[i + 1])
        main.wait_stream(side)
    return run


def timed(fn, n=3, w=2):
    for _ in range(w):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n / V


out = {"config": name, "views": V, "plain_ms_per_view": round(timed(plain), 4)}
lo, hi = torch.cuda.Stream.priority_range() if hasattr(torch.cuda.Stream, "priority_range") else (0, -1)
out["pipelined_high_priority"] = round(timed(make_pipelined(-1)), 4)
out["pipelined_same_priority"] = round(timed(make_pipelined(0)), 4)
out["pipelined_before_bwd"] = round(timed(make_pipelined(0, True)), 4)
out["pipelined_before_bwd_high_priority"] = round(timed(make_pipelined(-1, True)), 4)
out["plain_again"] = round(timed(plain), 4)
print(json.dumps(out), flush=True)

# ---- timeline of the pipelined batch: event timestamps on both streams, ms from the start of the batch
def timeline():
    side = torch.cuda.Stream()
    main = torch.cuda.current_stream()
    ev = lambda: torch.cuda.Event(enable_timing=True)   # noqa: E731
    marks = []
    t0 = ev(); t0.record(main)
    side.wait_stream(main)
    with torch.cuda.stream(side):
        a = ev(); a.record(side); prep = prepare(cams[0]); b = ev(); b.record(side)
    marks.append(("prep", 0, a, b))
    for i, cam in enumerate(cams):
        a = ev(); a.record(main)
        f = finish(prep)
        b = ev(); b.record(main)
        hz.surfel_backward(C, sc, cam, f, grads)
        c = ev(); c.record(main)
        marks.append(("stage2", i, a, b)); marks.append(("bwd", i, b, c))
        if i + 1 < len(cams):
            with torch.cuda.stream(side):
                a = ev(); a.record(side); prep = prepare(cams[i + 1]); b = ev(); b.record(side)
            marks.append(("prep", i + 1, a, b))
    main.wait_stream(side)
    torch.cuda.synchronize()
    return [(n, i, round(t0.elapsed_time(a), 3), round(t0.elapsed_time(b), 3)) for n, i, a, b in marks]


for _ in range(2):
    timeline()
tl = timeline()
print(json.dumps({"timeline_first_6_views": [m for m in tl if m[1] < 6]}), flush=True)
