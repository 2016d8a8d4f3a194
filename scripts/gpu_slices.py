"""GPU diagnostic: depth-slice schedules on a config -- tiles left unfinished, stage times, step time."""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import torch
import harness as hz  # noqa: E402 (sets sys.path)
import b200gs
import scenes

def timed(fn, n=10, w=3):
    for _ in range(w): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

for name in sys.argv[1:] or ["bench"]:
    sc_np, cam_np = scenes.build(name)
    sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
    W, H = cam_np["width"], cam_np["height"]
    grads = hz.to_dev(scenes.upstream_grads(W, H))
    C = hz.new_surfel_C()
    T = ((W + 15) // 16) * ((H + 15) // 16)
    for sched in ([1], None, [0.03], [0.03, 0.12], [0.015, 0.03, 0.06, 0.12, 0.25, 0.5], [0.01, 0.03, 0.1, 0.3], [0.02, 0.06], [0.05, 0.25]):
        kw = dict(depth_slices=1) if sched == [1] else (dict(depth_slices=0) if sched is None else dict(depth_slices=len(sched) + 1, slice_fractions=sched))
        f = hz.surfel_forward(C, sc, cam, **kw)
        done = torch.zeros(T, dtype=torch.uint8, device="cuda"); live = torch.zeros(1, dtype=torch.int32, device="cuda")
        b200gs.check(b200gs.surfel_inspect_tiles(f["img"].data_ptr(), W, H, done.data_ptr(), live.data_ptr(), None))
        torch.cuda.synchronize()
        def fb():
            ff = hz.surfel_forward(C, sc, cam, **kw)
            hz.surfel_backward(C, sc, cam, ff, grads)
        def fo():
            hz.surfel_forward(C, sc, cam, **kw)
        t_fb, t_f = timed(fb), timed(fo)
        b200gs.timing_enable(1); b200gs.timing_read(); fb(); torch.cuda.synchronize(); st = b200gs.timing_read(); b200gs.timing_enable(0)
        print(json.dumps({"config": name, "schedule": sched, "live_after": int(live[0]) if sched != [1] else None, "tiles": T,
                          "fwd_bwd_ms": round(t_fb, 3), "fwd_ms": round(t_f, 3),
                          "stages": {k: (round(v[0], 3), v[1]) for k, v in st.items() if v[1]}}), flush=True)
