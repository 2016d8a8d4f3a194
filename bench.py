#!/usr/bin/env python
"""bench.py -- fwd+bwd throughput of the surfel (2DGS, reflection channels) differentiable rasterizer on the
workload BASELINE.json's metric is quoted on: 3M Gaussians, SH degree 3, 1920x1080.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl new|reference] [--workload bench|c5|...]

One process per GPU (torchrun for N > 1).  A "step" is one BATCH of --views-per-step (default 32) ring-camera views
of the replicated scene, dealt to the ranks (rank r renders views r, r+N, ...): per view forward (preprocess, depth
order, scan, emission, tile sort, tile ranges, block masks, compositing) + backward (compositing backward, preprocess
backward) with non-zero upstream gradients on every differentiable output plane; the parameter gradients of the batch
are summed in place in one flat [59 P] arena and, for N > 1, all-reduced once per batch.  Total
work is fixed as N grows -> "scaling": "strong".

  value        Mpix/s over all ranks (views x W x H / step time), inputs resident in HBM, CUDA events, max over ranks.
  e2e          the same batch through the drop-in Python API (GaussianRasterizer autograd module + an L1 loss) with each
               view's host inputs (camera matrices + target image, pinned) copied H2D and the batch loss read back D2H
               inside the timed region; gradients land in the arena too (GaussianRasterizer(grad_arena=...)).
  roofline     the dominant kernel, timed with CUDA event pairs recorded by the library on the launching stream inside
               the timed region (b200gs_timing_only).  The compositing kernels are instruction-issue bound (DRAM at
               1-2 % of peak, the staged records are re-served from L2), so for them "bound" is "issue": achieved = warp
               instructions per second (instruction count per launch from the committed ncu capture,
               profiles/traffic.json) against 148 SMs x 4 schedulers x the SM clock sampled during the run.  The
               contract's HBM figure of the same kernel (algorithmic bytes / duration / measured copy peak) is kept in
               `hbm_view`, and `roofline_hbm` is the slowest HBM-bound stage with its own fraction.
  cpu_baseline the CPU oracle (oracle/surfel_oracle.py, kind "port": the reference has no CPU path) on the WHOLE c1 view
               (BASELINE.json configs[0]: 100k Gaussians, 256x256, fwd + autograd bwd) on all host cores, with this
               library's time for the same view next to it; rank 0 at N=1 only.
  scale        N > 1: per-rank compute ms (min / max over ranks), all-reduce ms, bus GB/s against NVLink 5, and the
               fraction of the all-reduce hidden behind compute.
  forward_only the render_fast()-shaped inference call (torch.no_grad, eval_fps.py:25-59): ms per view and fps.
  --impl reference  runs the UNMODIFIED reference CUDA extension (oracle/_ref/ref_surfel_C.so) on the same inputs
               through the same driver; this arm imports nothing of this repo's library.  The reference's
               implementation of this path exists only as CUDA, so "the reference on this box" is that, on the GPU.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "gaussian-splatting-reflection_b200", "python")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "surfel rasterizer fwd+bwd throughput, 3M Gaussians SH3 @1920x1080 (32-view batch sharded over the GPUs)"
UNIT = "Mpix/s"
WORKLOAD = "bench"   # tests/scenes.py CONFIGS["bench"]: room scene, P=3M, 1920x1080
NVLINK5_GBS = 900.0  # per direction per GPU


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="new", choices=["new", "reference"])
    ap.add_argument("--workload", default=WORKLOAD)
    ap.add_argument("--points", type=int, default=0, help="override P (debug only; result is then not the headline)")
    ap.add_argument("--views-per-step", type=int, default=32,
                    help="views of one step (a batch), sharded over the ranks: rank r renders views r, r+n, ... and the "
                         "per-Gaussian gradients of the batch are summed (BASELINE.json configs[4], SURVEY.md 8e)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-stage-timing", action="store_true",
                    help="diagnosis: no per-stage CUDA events inside the timed region (then no roofline / stage table)")
    ap.add_argument("--exact-lists", action="store_true",
                    help="time the exact-lists mode (the reference's instance lists bit for bit) instead of the "
                         "default culled emission")
    ap.add_argument("--no-pipeline", action="store_true",
                    help="diagnosis: every view start to end on one stream (no stage-1 prefetch of the next view)")
    ap.add_argument("--nccl-allreduce", action="store_true",
                    help="N > 1: reduce the arena with ncclAllReduce instead of the library's NVLS multimem kernel")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def __enter__(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:  # noqa: BLE001
            self.proc = None
        return self

    def __exit__(self, *a):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:  # noqa: BLE001
                self.proc.kill()

    def summary(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.path or not os.path.exists(self.path):
            return out
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.path)
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), samples=len(sm))
        out["reasons"] = sorted(reasons)
        return out


def algorithmic_bytes(P, V, R, N):
    """compulsory HBM bytes per launch of each stage (surfel path, SH degree 3, colours from
    SH), each datum moved once.  P Gaussians, V visible, R (Gaussian, tile) instances the stage works on (the emitted
    ones: all of the reference's in exact-lists mode, the reachable ones in the default culled-emission mode), N pixels."""
    return {
        "count_kept": 4 * P + 72 * V + 4 * P,           # tiles_touched; rect + cull record of visible ones; count out
        "preprocess": 237 * P + 87 * V + 8 * P,        # inputs; state of visible ones; radii + tiles_touched
        "depth_sort": 16 * P,                           # (depth bits, id) pairs read once, written once
        "scan": 8 * P,
        "duplicate": 16 * V + 8 * R,                    # id + offset + rect in, (tile, id) out
        "sort": 16 * R,                                 # read the 8-B pairs once, write them once
        "tile_ranges": 4 * R,
        "tile_hits": 8 * R + 64 * V + 9 * R // 16,      # (tile, id) in, cull record per Gaussian, survivors out (~1/16)
        "composite_fwd": 64 * R + 68 * N,               # staged record per tile entry; 12 planes + 5 state words out
        "composite_bwd": 80 * R + 68 * N + 60 * V,      # record + colour/refl; pixel state + 12 grad planes; partial sums
        "preprocess_bwd": 237 * P + 60 * V + 248 * V,   # inputs again, partial sums in, final gradients out
    }


ISSUE_BOUND = ("composite_fwd", "composite_bwd", "tile_hits", "count_kept")   # what bounds each stage


def cpu_baseline():
    """BASELINE.json configs[0]: the WHOLE c1 view (100k Gaussians, SH degree 3, 256x256), forward + autograd
    backward, through the CPU oracle on all host cores.  Tiles are processed in chunks so the autograd tape of a chunk
    fits in memory; every tile, every list entry and every Gaussian is processed -- nothing is extrapolated."""
    import numpy as np
    import torch
    from oracle import surfel_oracle as so
    import scenes
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sc_np, cam_np = scenes.build("c1")
    W, H = cam_np["width"], cam_np["height"]
    gx, gy = (W + 15) // 16, (H + 15) // 16
    grads = scenes.upstream_grads(W, H)
    gc, go, gr = (torch.from_numpy(grads[k]) for k in ("color", "others", "refl"))
    go = go.clone()
    go[7] = 0
    t0 = time.time()
    leaves = so._leaves(sc_np, torch.float32, True)
    pre = so.preprocess(leaves, cam_np)
    keys, plist, ranges = so.binning(pre, W, H)
    tiles = np.arange(gx * gy)
    chunk = 32
    for c0 in range(0, gx * gy, chunk):
        comp = so.composite(pre, plist, ranges, cam_np, (0.0, 0.0, 0.0), tiles=tiles[c0:c0 + chunk], collect_weights=False)
        color, others, refl, _, _ = so.assemble(comp, W, H, torch.float32)
        ((color * gc).sum() + (others * go).sum() + (refl * gr).sum()).backward(retain_graph=True)
    t_view = time.time() - t0
    return {"value": W * H / t_view / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"oracle/surfel_oracle.py (torch fp32 on the host, {cores} threads): the whole c1 view -- P={sc_np['means3D'].shape[0]}, "
                      f"{W}x{H}, R={int(keys.size)} instances, forward + autograd backward of all {gx * gy} tiles in chunks of {chunk} -- "
                      f"{t_view:.1f} s; the reference has no CPU implementation of this path",
            "workload": "c1", "seconds": round(t_view, 2)}


def gpu_c1_ms(Cmod):
    """This library on the same c1 view the CPU baseline renders (median of 20 fwd+bwd after 5 warm-ups)."""
    import torch
    import harness as hz
    import scenes
    sc_np, cam_np = scenes.build("c1")
    sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
    grads = hz.to_dev(scenes.upstream_grads(cam_np["width"], cam_np["height"]))
    ts = []
    for i in range(25):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        f = hz.surfel_forward(Cmod, sc, cam)
        hz.surfel_backward(Cmod, sc, cam, f, grads)
        e1.record()
        torch.cuda.synchronize()
        if i >= 5:
            ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    a = parse()
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    import harness as hz
    import scenes
    new = a.impl == "new"
    b200gs = None
    if new:
        import b200gs                                   # the reference arm never loads libb200gs.so
        Cmod = hz.new_surfel_C()
        impl_desc = "libb200gs.so (sm_100a) via diff_surfel_rasterization._C"
    else:
        Cmod = hz.ref_module("ref_surfel_C")
        impl_desc = "unmodified reference CUDA extension oracle/_ref/ref_surfel_C.so (built for sm_100a)"
        if Cmod is None:
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable":
                                  "oracle/_ref/ref_surfel_C.so not built (python oracle/build_ref.py needs /root/reference)"}))
            return

    # ---- synthetic workload: replicated Gaussians, a batch of ring-camera views sharded over the ranks ----
    VB = a.views_per_step
    if VB < world or VB % world:
        raise SystemExit(f"--views-per-step {VB} must be a positive multiple of the number of GPUs ({world})")
    my_views = list(range(rank, VB, world))               # r, r+n, ...  
    sc_np, cam_np = scenes.build(a.workload, P=a.points or None, view=my_views[0] % 32, n_views=32)
    P, W, H = sc_np["means3D"].shape[0], cam_np["width"], cam_np["height"]
    M = sc_np["shs"].shape[1]
    N = W * H
    sc = hz.to_dev(sc_np)
    cams_np = [cam_np] + [scenes.camera(a.workload, v % 32, 32) for v in my_views[1:]]
    cams = [hz.to_dev(c) for c in cams_np]
    cam = cams[0]
    grads = hz.to_dev(scenes.upstream_grads(W, H))
    GRAD_KEYS = ("means3D", "sh", "opacity", "scales", "rotations", "refl")   # the 59 floats / Gaussian of the model
    exact = bool(a.exact_lists)

    # The parameter gradients of the views of a batch are summed.  This library: the backward ADDS them into one
    # flat [59 P] arena in place (b200gs.dp.GradArena, accumulate mode, a per-call argument of the backward), which
    # is then reduced over the ranks as it lies.  Reference arm: what autograd does for it -- one add per gradient
    # tensor per view -- and one concatenation before the all-reduce.
    arena, ref_acc = None, None
    if new:
        from b200gs.dp import GradArena
        arena = GradArena(P, M, "cuda", accumulate=True, symmetric=(world > 1 and not a.nccl_allreduce))

    pipelined = new and not a.no_pipeline and len(cams) > 1
    side = torch.cuda.Stream() if pipelined else None
    ev = {}                                                # optional event marks of the scale pass

    def mark(name):
        if ev.get("on"):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            ev.setdefault("marks", []).append((name, e))

    def step():
        nonlocal ref_acc
        mark("begin")
        if new and pipelined:
            main = torch.cuda.current_stream()
            side.wait_stream(main)        # everything before this step; NOT the arena memset below, which stage 1 never reads
        if arena is not None:
            arena.zero_()
        f = b = None
        if new and pipelined:
            # software pipeline over the batch's views: stage 1 of view i+1 (preprocess, depth order, scan, plan and
            # the forward's host sync) runs on a side stream while stage 2 + backward of view i occupy the SMs
            with torch.cuda.stream(side):
                prep = hz.surfel_prepare(Cmod, sc, cams[0], exact=exact)
            for vi, c in enumerate(cams):
                f = hz.surfel_finish(Cmod, prep)
                b = hz.surfel_backward(Cmod, sc, c, f, grads, grad_arena=arena)
                if vi + 1 < len(cams):
                    with torch.cuda.stream(side):
                        prep = hz.surfel_prepare(Cmod, sc, cams[vi + 1], exact=exact)
            main.wait_stream(side)
        for vi, c in enumerate(cams if not (new and pipelined) else []):
            if new:
                f = hz.surfel_forward(Cmod, sc, c, exact=exact)
                b = hz.surfel_backward(Cmod, sc, c, f, grads, grad_arena=arena)
            else:
                f = hz.surfel_forward(Cmod, sc, c)
                b = hz.surfel_backward(Cmod, sc, c, f, grads)
                if vi == 0:
                    ref_acc = {k: b[k] for k in GRAD_KEYS}
                else:
                    for k in GRAD_KEYS:
                        ref_acc[k] += b[k]
        mark("compute_end")
        if world > 1:
            if arena is not None:
                arena.all_reduce()
            else:
                flat = torch.cat([ref_acc[k].reshape(P, -1) for k in GRAD_KEYS], 1)
                dist.all_reduce(flat)
        mark("end")
        return f, b

    def timed(fn, steps, warmup, after_warmup=None):
        for _ in range(warmup):
            fn()
        if after_warmup is not None:
            torch.cuda.synchronize()
            after_warmup()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # workload counts of this rank's first view (for the algorithmic-bytes figures of the stage table)
    R_ref = None
    if new:
        f = hz.surfel_forward(Cmod, sc, cam, exact=True)
        R_ref = int(f["rendered"])           # the reference's instance count
        del f
    f = hz.surfel_forward(Cmod, sc, cam, **({"exact": exact} if new else {}))
    R, V = int(f["rendered"]), int((f["radii"] > 0).sum())
    if R_ref is None:
        R_ref = R
    del f

    # ---- per-stage table: a separate instrumented pass ------------------------------------------
    # An event pair between every two kernels keeps a launch from overlapping the previous kernel's tail
    # (measured: +0.4 ms on a 6.1 ms view), so the all-stage table comes from its own short pass and the
    # timed region below carries events around the dominant kernel only.
    stages_all, top_stage = {}, -1
    if new and not a.no_stage_timing:
        step()
        torch.cuda.synchronize()
        b200gs.timing_enable(1)
        b200gs.timing_read()
        was_pipelined, pipelined = pipelined, False      # one stream for this pass: an event pair around a stage must not
        step()                                           # also cover another view's stage 1 running next to it
        pipelined = was_pipelined
        torch.cuda.synchronize()
        stages_all = b200gs.timing_read()
        b200gs.timing_enable(0)
        names = [b200gs.timing_stage_name(i).decode() for i in range(b200gs.timing_stages())]
        top_name = max((k for k, v in stages_all.items() if v[1]), key=lambda k: stages_all[k][0], default=None)
        top_stage = names.index(top_name) if top_name else -1

    # ---- device-resident timing (value), events around the dominant kernel, clock sampling -------
    if top_stage >= 0:
        b200gs.timing_only(top_stage)
    n0 = [0]

    def reset_counters():                 # warm-up launches are not part of the timed region
        if new:
            b200gs.timing_read()
            n0[0] = b200gs.launch_count()
    with ClockSampler(local) as cs:
        total_ms = timed(step, a.steps, a.warmup, reset_counters)
    launches = (b200gs.launch_count() - n0[0]) if new else None
    stages = b200gs.timing_read() if top_stage >= 0 else {}      # the dominant kernel, live in the timed region
    if new:
        b200gs.timing_enable(0)
    clocks = cs.summary()
    ms_per_step = total_ms / a.steps
    value = VB * N / (ms_per_step * 1e-3) / 1e6            # whole job: every rank's views

    # ---- N > 1: where the step's time goes on every rank -----------------------------------------
    scale = None
    if world > 1:
        ev["on"] = True
        per = []
        for _ in range(3):
            ev["marks"] = []
            dist.barrier()
            step()
            torch.cuda.synchronize()
            m = dict(ev["marks"])
            per.append((m["begin"].elapsed_time(m["compute_end"]), m["compute_end"].elapsed_time(m["end"])))
        ev["on"] = False
        comp_ms = sorted(p[0] for p in per)[1]
        tail_ms = sorted(p[1] for p in per)[1]             # all-reduce as seen by this rank: includes waiting for the slowest rank
        # the collective alone: every rank enters it at the same time
        ar = []
        probe = None if arena is not None else torch.zeros(P * 59, device="cuda")
        for _ in range(5):
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if arena is not None:
                arena.all_reduce()
            else:
                dist.all_reduce(probe)
            e1.record()
            torch.cuda.synchronize()
            ar.append(e0.elapsed_time(e1))
        ar_ms = sorted(ar)[len(ar) // 2]
        t = torch.tensor([comp_ms, -comp_ms, tail_ms, ar_ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        comp_max, comp_min, tail_max, ar_max = float(t[0]), -float(t[1]), float(t[2]), float(t[3])
        nbytes = P * 59 * 4
        scale = {"compute_ms_per_rank_min": round(comp_min, 3), "compute_ms_per_rank_max": round(comp_max, 3),
                 "allreduce_ms": round(ar_max, 3), "allreduce_bytes": nbytes,
                 "allreduce_algo": ("ncclAllReduce" if (arena is None or not arena.symmetric) else arena.algo),
                 "bus_gbs": round(2 * (world - 1) / world * nbytes / (ar_max * 1e-3) / 1e9, 1),
                 "alg_gbs": round(nbytes / (ar_max * 1e-3) / 1e9, 1), "nvlink5_gbs_per_direction": NVLINK5_GBS,
                 # time of the step not covered by the slowest rank's compute, against the collective's own time
                 "exposed_comm_ms": round(ms_per_step - comp_max, 3),
                 "overlap_fraction": round(max(0.0, 1.0 - max(ms_per_step - comp_max, 0.0) / ar_max), 3),
                 "imbalance_ms": round(comp_max - comp_min, 3),
                 "note": "compute = zero_() + this rank's views; allreduce timed alone after a barrier (median of 5, max over ranks); "
                         "bus_gbs = 2(n-1)/n x bytes / time"}

    # ---- forward only (render_fast-shaped inference, eval_fps.py:25-59) ---------------------------
    fwd_only = None
    if rank == 0 or world > 1:
        with torch.no_grad():
            def fstep():
                if new and pipelined:      # stage 1 of the next frame on the side stream, as in the training step
                    main = torch.cuda.current_stream()
                    side.wait_stream(main)
                    with torch.cuda.stream(side):
                        prep = hz.surfel_prepare(Cmod, sc, cams[0], exact=exact)
                    for vi in range(len(cams)):
                        hz.surfel_finish(Cmod, prep)
                        if vi + 1 < len(cams):
                            with torch.cuda.stream(side):
                                prep = hz.surfel_prepare(Cmod, sc, cams[vi + 1], exact=exact)
                    main.wait_stream(side)
                    return
                for c in cams:
                    hz.surfel_forward(Cmod, sc, c, **({"exact": exact} if new else {}))
            f_ms = timed(fstep, max(3, a.steps // 4), 2) / max(3, a.steps // 4) / len(cams)
        fwd_only = {"ms_per_view": round(f_ms, 4), "fps_per_gpu": round(1e3 / f_ms, 1),
                    "api": "_C.rasterize_gaussians under torch.no_grad(), device-resident inputs"
                           + (" (stage 1 of the next frame prefetched on a side stream: _prepare / _finish)" if new and pipelined else "")}

    # ---- end to end through the drop-in Python API, host inputs ------------------------------
    e2e = None
    if not a.no_e2e:
        if new:
            import diff_surfel_rasterization as dsr
        else:
            import refpy                                   # the reference's own python package on its own extension
            arm = refpy.load_arm("ref")
            dsr = arm.diff_surfel_rasterization if arm is not None else None
        if dsr is not None:
            leaves = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations", "refl")}
            m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
            mask = torch.ones((P, 3), dtype=torch.bool, device="cuda")
            hosts = [{k: torch.from_numpy(c[k]).pin_memory() for k in ("viewmatrix", "projmatrix", "campos")} for c in cams_np]
            target = torch.rand((3, H, W)).pin_memory()   # one pinned target image, copied H2D again for every view
            for hv in hosts:
                hv["target"] = target
                hv["bg"] = torch.zeros(3).pin_memory()
            h2d = sum(t.numel() * t.element_size() for t in hosts[0].values()) * VB     # whole job
            loss_host = torch.zeros(1).pin_memory()
            copy_stream = torch.cuda.Stream()
            kw = {}
            if new:
                # the leaves' .grad ARE the arena blocks: the backward of every view adds into them in place, the
                # all-reduce runs on that memory, the optimizer would read .grad -- no per-view adds, no concatenation
                arena.bind_grads({"means3D": leaves["means3D"], "sh": leaves["shs"], "opacity": leaves["opacities"],
                                  "scales": leaves["scales"], "rotations": leaves["rotations"], "refl": leaves["refl"]})
                kw = dict(grad_arena=arena, exact_lists=exact)

            def upload(host):
                """H2D of one view's inputs on the copy stream (a data loader's prefetch); still inside the timed region."""
                with torch.cuda.stream(copy_stream):
                    dev = {k: v.to("cuda", non_blocking=True) for k, v in host.items()}
                    e = torch.cuda.Event()
                    e.record(copy_stream)
                return dev, e

            def e2e_step():
                if new:
                    arena.zero_()
                    m2d.grad = None
                else:
                    for t in list(leaves.values()) + [m2d]:
                        t.grad = None
                total = None
                inputs = dict(means3D=leaves["means3D"], means2D=m2d, opacities=leaves["opacities"], shs=leaves["shs"],
                              refl_strengths=leaves["refl"], scales=leaves["scales"], rotations=leaves["rotations"],
                              env_scope_mask=mask)

                def module(dev):
                    rs = dsr.GaussianRasterizationSettings(
                        image_height=H, image_width=W, tanfovx=cam_np["tanfovx"], tanfovy=cam_np["tanfovy"], bg=dev["bg"],
                        scale_modifier=1.0, viewmatrix=dev["viewmatrix"], projmatrix=dev["projmatrix"], sh_degree=3,
                        campos=dev["campos"], prefiltered=False, debug=False)
                    return dsr.GaussianRasterizer(rs, **kw)

                def prepared(view):
                    """this library only: stage 1 of that view's forward on the side stream, now"""
                    dev, e = view
                    side.wait_event(e)
                    for t in dev.values():
                        t.record_stream(side)
                    with torch.cuda.stream(side):
                        return module(dev).prepare(**inputs)

                cur = torch.cuda.current_stream()
                nxt = upload(hosts[0])
                mod_next = None
                if pipelined:
                    side.wait_stream(cur)
                    mod_next = prepared(nxt)
                for vi in range(len(hosts)):             # this rank's views of the batch; the gradients are summed
                    dev, e = nxt
                    if vi + 1 < len(hosts):
                        nxt = upload(hosts[vi + 1])      # the next view's inputs travel while this one renders
                    cur.wait_event(e)
                    for t in dev.values():
                        t.record_stream(cur)
                    mod = mod_next if pipelined else module(dev)
                    color, radii, allmap, refl_map, _ = mod(**inputs)
                    loss = (color - dev["target"]).abs().mean() + 1e-3 * allmap[:7].mean() + 1e-3 * refl_map.mean()
                    loss.backward()
                    if pipelined and vi + 1 < len(hosts):
                        mod_next = prepared(nxt)         # stage 1 of the next view next to this view's kernels
                    total = loss.detach() if total is None else total + loss.detach()
                if pipelined:
                    cur.wait_stream(side)
                if world > 1:
                    if new:
                        arena.all_reduce()
                    else:
                        flat = torch.cat([leaves[k].grad.reshape(P, -1) for k in leaves], 1)
                        dist.all_reduce(flat)
                loss_host.copy_(total.reshape(1), non_blocking=True)
                torch.cuda.current_stream().synchronize()     # the caller reads the batch loss every step
                return float(loss_host[0])

            e2e_ms = timed(e2e_step, a.steps, a.warmup) / a.steps
            e2e = {"value": VB * N / (e2e_ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": e2e_ms, "ms_per_view": e2e_ms * world / VB,
                   "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4,
                   "api": "diff_surfel_rasterization.GaussianRasterizer (autograd) + L1 loss"
                          + (", gradients accumulated in the GradArena the leaves' .grad are bound to" if new else "")
                          + (", stage 1 of the next view prefetched on a side stream (GaussianRasterizer.prepare)" if pipelined else "")}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- rank 0: rooflines, CPU baseline, the JSON line ----------------
    peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
    except Exception:  # noqa: BLE001
        pass
    alg = algorithmic_bytes(P, V, R, N)
    roofline, roofline_hbm, stage_table = None, None, {}
    prof = {}
    try:   # per-launch dram bytes / warp instructions from the committed ncu captures of this workload
        prof = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:  # noqa: BLE001
        pass

    views_per_rank = VB // world

    def row(name, ms, cnt, steps):
        # a stage may launch several times per view (one launch per depth slice): `ms` is the stage's time PER VIEW
        per = ms / (steps * views_per_rank)
        return {"ms": round(per, 4), "launches_per_step": cnt / (steps + 0.0), "launches_per_view": round(cnt / (steps * views_per_rank + 0.0), 2),
                "alg_bytes": alg[name], "gbs": round(alg[name] / (per * 1e-3) / 1e9, 1),
                "frac_of_hbm_peak": round(alg[name] / (per * 1e-3) / 1e9 / peak, 4),
                "bound": "issue" if name in ISSUE_BOUND else "hbm"}
    for name, (ms, cnt) in stages_all.items():           # the separate all-stage pass (1 step)
        if cnt and name in alg:
            stage_table[name] = row(name, ms, cnt, 1)
    live = {name: row(name, ms, cnt, a.steps) for name, (ms, cnt) in stages.items() if cnt and name in alg}
    if live:
        top = max(live, key=lambda k: live[k]["ms"])   # one entry: the dominant kernel, timed inside the timed region
        t_s = live[top]["ms"] * 1e-3
        common = {"kernel": top, "ms_per_launch": live[top]["ms"], "launches_per_step": live[top]["launches_per_step"],
                  "share_of_step": round(live[top]["ms"] * views_per_rank / ms_per_step, 3),
                  "traffic": (prof.get(top) if not isinstance(prof.get(top), dict) else prof[top].get("dram_bytes")),
                  "hbm_view": {"alg_bytes_per_launch": live[top]["alg_bytes"], "achieved_gbs": live[top]["gbs"], "peak_gbs": peak,
                               "frac": live[top]["frac_of_hbm_peak"], "peak_source": peak_src}}
        inst = prof.get(top + "_warp_inst") if top in ISSUE_BOUND else None
        if inst:
            sm_mhz = clocks.get("sm_mhz") or 1965.0
            ipeak = 148 * 4 * sm_mhz * 1e6 / 1e9          # G warp-instructions / s at one issue per scheduler per cycle
            ach = inst / t_s / 1e9
            roofline = dict(common, bound="issue", achieved=round(ach, 1), peak=round(ipeak, 1), unit="Gwarp-inst/s",
                            frac=round(ach / ipeak, 4), warp_inst_per_launch=inst,
                            note="the kernel's DRAM traffic is 1-2 % of peak (staged records are re-served from L2): it is bound by "
                                 "instruction issue; achieved = warp instructions per launch (ncu smsp__inst_executed.sum, "
                                 "profiles/traffic.json, mean over the batch's views) / measured duration; peak = 148 SMs x 4 "
                                 "schedulers x the SM clock sampled during the timed region")
        else:
            roofline = dict(common, bound="hbm", achieved=live[top]["gbs"], peak=peak, unit="GB/s", frac=live[top]["frac_of_hbm_peak"],
                            peak_source=peak_src)
    hbm_rows = {k: v for k, v in stage_table.items() if v["bound"] == "hbm"}
    if hbm_rows:
        k = max(hbm_rows, key=lambda k: hbm_rows[k]["ms"])
        roofline_hbm = {"kernel": k, "bound": "hbm", "achieved": hbm_rows[k]["gbs"], "peak": peak, "unit": "GB/s",
                        "frac": hbm_rows[k]["frac_of_hbm_peak"], "ms_per_launch": hbm_rows[k]["ms"],
                        "traffic": (prof.get(k) if not isinstance(prof.get(k), dict) else prof[k].get("dram_bytes")),
                        "note": "the slowest HBM-bound stage (separate all-stage pass)"}

    cpu = None
    if new and world == 1 and not a.no_cpu_baseline:
        cpu = cpu_baseline()
        cpu["gpu_same_view_ms"] = round(gpu_c1_ms(Cmod), 4)
        cpu["gpu_same_view_mpix_s"] = round(256 * 256 / (cpu["gpu_same_view_ms"] * 1e-3) / 1e6, 2)
    elif not new:
        cpu = {"value": value, "unit": UNIT, "cores": 0, "kind": "reference",
               "sample": "the reference has no CPU implementation of this path (CUDA only): this arm is its "
                         "unmodified CUDA extension on the same B200, same inputs, whole workload"}

    line = {
        "metric": (METRIC if a.workload == WORKLOAD and not a.points and VB == 32 else
                   f"surfel rasterizer fwd+bwd throughput, {P / 1e6:g}M Gaussians SH3 @{W}x{H} ({VB}-view batch sharded over the GPUs)"),
        "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": round(ms_per_step, 4), "ms_per_view": round(ms_per_step * world / VB, 4),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        # workload only: identical for both arms
        "config": {"workload": f"{a.workload}: room scene, replicated Gaussians, batch of {VB} ring-camera views per step, "
                               f"{VB // world} per GPU (rank r renders views r, r+n, ...), gradients of the batch summed",
                   "gaussians": P, "width": W, "height": H, "sh_degree": 3, "visible": V, "instances": R_ref,
                   "views_per_step": VB, "views_per_gpu": VB // world,
                   "counts": "visible / instances are those of rank 0's first view",
                   "l2": "inputs (>= 0.7 GB of Gaussian data + 0.5-1.3 GB of sort keys per view) exceed the 126 MB L2; no flush",
                   "parallelism": f"view-parallel x{world}, replicated Gaussians" + (", one all-reduce of the [P,59] fp32 gradients per batch" if world > 1 else "")},
        "impl_detail": {"implementation": impl_desc, "instances_emitted": R,
                        "lists": ("exact (reference's lists bit for bit)" if (exact or not new)
                                  else "culled emission (only instances that can reach alpha >= 1/255)"),
                        "view_pipeline": ("stage 1 of view i+1 on a side stream next to stage 2 + backward of view i"
                                          if pipelined else "none: every view start to end on one stream")},
        "clocks": clocks, "e2e": e2e, "gpu_launches": (int(launches) if launches is not None else None),
        "roofline": roofline, "roofline_hbm": roofline_hbm, "cpu_baseline": cpu, "stages": stage_table,
        "scale": scale, "forward_only": fwd_only,
    }
    if not new:
        line["impl"] = "reference"
        line["gpu_launches"] = None
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
