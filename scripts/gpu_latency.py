"""GPU diagnostic: where a small view's time goes -- stage sum on the device, event-timed step, host time to issue
the step (wall clock up to the last call's return, GPU still busy), launches per view.
  python scripts/gpu_latency.py c1 c2"""
import json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import torch
import harness as hz  # noqa: E402 (sets sys.path)
import b200gs
import scenes

N = 50
for name in sys.argv[1:] or ["c1", "c2"]:
    gauss = name.endswith("g")             # c3g: c3 through the 3DGS (EWA) rasterizer
    sc_np, cam_np = scenes.build(name[:-1] if gauss else name)
    sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
    W, H = cam_np["width"], cam_np["height"]
    if gauss:
        sc, grads, C = hz.gauss_scene(sc), hz.to_dev(hz.gauss_grads(W, H)), hz.new_gauss_C()
        fwd = lambda: hz.gauss_forward(C, sc, cam)                          # noqa: E731
        fb = lambda: hz.gauss_backward(C, sc, cam, fwd(), grads)            # noqa: E731
    else:
        grads, C = hz.to_dev(scenes.upstream_grads(W, H)), hz.new_surfel_C()
        fwd = lambda: hz.surfel_forward(C, sc, cam)                         # noqa: E731
        fb = lambda: hz.surfel_backward(C, sc, cam, fwd(), grads)           # noqa: E731

    out = {"config": name}
    for label, fn in (("fwd", fwd), ("fwd_bwd", fb)):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        l0 = b200gs.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(N):
            fn()
        e1.record()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        out[label] = {"event_ms": round(e0.elapsed_time(e1) / N, 4), "host_issue_ms": round((t1 - t0) * 1e3 / N, 4),
                      "wall_ms": round((t2 - t0) * 1e3 / N, 4), "launches": (b200gs.launch_count() - l0) / N}
    b200gs.timing_enable(1); b200gs.timing_read(); fb(); torch.cuda.synchronize(); st = b200gs.timing_read(); b200gs.timing_enable(0)
    out["stages"] = {k: (round(v[0], 4), v[1]) for k, v in st.items() if v[1]}
    out["stage_sum_ms"] = round(sum(v[0] for v in st.values()), 4)
    print(json.dumps(out), flush=True)
