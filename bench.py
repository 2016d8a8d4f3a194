#!/usr/bin/env python
"""bench.py -- fwd+bwd throughput of the surfel (2DGS, reflection channels) differentiable
rasterizer on the workload BASELINE.json's metric is quoted on: 3M Gaussians, SH degree 3,
1920x1080, one camera view per GPU per step.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl new|reference]

One process per GPU (torchrun for N>1).  A "step" = one pass of the hot path over one view per
rank: forward (preprocess, scan, duplicate, 64-bit sort, tile ranges, compositing) + backward
(compositing backward, preprocess backward) with non-zero upstream gradients on every
differentiable output plane, then -- for N>1 -- one NCCL all-reduce of the per-Gaussian gradients
(Gaussians replicated, views sharded: SURVEY.md section 8e).  Per-GPU work is fixed -> weak scaling.

  value        Mpix/s over all ranks, inputs resident in HBM, CUDA-event timed, max over ranks.
  e2e          same metric through the drop-in Python API (GaussianRasterizer autograd module)
               with the step's host inputs (camera matrices + target image, pinned) copied
               H2D and the loss read back D2H inside the timed region.
  roofline     dominant kernel: algorithmic bytes per launch / its mean duration, measured with
               CUDA event pairs recorded by the library on the launching stream during the
               timed region (b200gs_timing_*), against MEASURED_PEAKS.json's HBM copy figure.
  cpu_baseline the CPU oracle (oracle/surfel_oracle.py, "port": the reference has no CPU path)
               on a bounded sample of the same view, rank 0 at N=1 only.
  --impl reference  runs the UNMODIFIED reference CUDA extension (oracle/_ref/ref_surfel_C.so)
               on the same inputs through the same driver: the reference's implementation of
               this path exists only as CUDA, so "the reference on this box" is that, on the GPU.
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
    ap.add_argument("--cpu-sample-tiles", type=int, default=64)
    ap.add_argument("--cpu-sample-len", type=int, default=600, help="list positions walked per sampled tile")
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
    """compulsory HBM bytes per launch of each stage (surfel path, SH
    degree 3, colours from SH), each datum moved once.  P Gaussians, V visible, R (Gaussian,tile)
    instances the stage works on (the emitted ones: all of the reference's in exact-lists mode,
    the reachable ones in the default culled-emission mode), N pixels."""
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


def cpu_baseline(scene_np, cam_np, n_tiles, max_len, R_total):
    """Oracle (CPU, torch fp32) on a bounded sample of the same view: the full per-Gaussian
    preprocess, then binning + compositing forward + autograd backward over the first `max_len`
    list positions of `n_tiles` 16x16 tiles spread over the image.  Like a CTA of the CUDA kernels
    the vectorised oracle walks a tile's list for all 256 pixels together, so the cost of a whole
    view is extrapolated per (tile, list entry): R_total entries in the view."""
    import numpy as np
    import torch
    from oracle import surfel_oracle as so
    import scenes
    cores = 1      # [S,256]-sized tensors: intra-op threading only adds overhead (measured 3.5x slower with 8)
    torch.set_num_threads(cores)
    W, H = cam_np["width"], cam_np["height"]
    gx, gy = (W + 15) // 16, (H + 15) // 16
    tiles = sorted({int((gy * (2 * k + 1)) // (2 * n_tiles)) * gx + int((gx * (2 * k + 1)) // (2 * n_tiles))
                    for k in range(n_tiles)})
    grads = scenes.upstream_grads(W, H)
    t0 = time.time()
    with torch.no_grad():
        pre = so.preprocess(so._leaves(scene_np, torch.float32, False), cam_np)      # every Gaussian, forward
    t_pre = time.time() - t0
    keys, plist, ranges = so.binning(pre, W, H, tiles=tiles)
    ranges = ranges.copy()
    ranges[:, 1] = ranges[:, 0] + (ranges[:, 1] - ranges[:, 0]).clip(max=max_len)
    entries = int(sum(ranges[t, 1] - ranges[t, 0] for t in tiles))
    # The per-pair cost is timed on the Gaussians of the sampled lists only: autograd through the preprocess of
    # all P Gaussians makes every backward op a dense P-row scatter (60 s for 1200 entries at P = 3M), which
    # says nothing about the per-pair work being extrapolated.
    Pn = scene_np["means3D"].shape[0]
    ids = np.unique(np.concatenate([plist[ranges[t, 0]:ranges[t, 1]] for t in tiles])) if entries else np.zeros(0, np.int64)
    sub_np = {k: (v[ids] if isinstance(v, np.ndarray) and v.ndim >= 1 and v.shape[0] == Pn else v) for k, v in scene_np.items()}
    plist_sub = np.searchsorted(ids, np.clip(plist, ids.min() if len(ids) else 0, ids.max() if len(ids) else 0)).astype(plist.dtype)
    t0 = time.time()
    pre_sub = so.preprocess(so._leaves(sub_np, torch.float32, True), cam_np)
    comp = so.composite(pre_sub, plist_sub, ranges, cam_np, (0.0, 0.0, 0.0), tiles=tiles, collect_weights=False)
    color, others, refl, _, _ = so.assemble(comp, W, H, torch.float32)
    loss = ((color * torch.from_numpy(grads["color"])).sum() + (others[:7] * torch.from_numpy(grads["others"][:7])).sum()
            + (refl * torch.from_numpy(grads["refl"])).sum())
    loss.backward()
    t_tiles = time.time() - t0
    t_view = 2.0 * t_pre + t_tiles * R_total / max(entries, 1)     # preprocess forward + (about as much) backward
    return {"value": W * H / t_view / 1e6, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"oracle/surfel_oracle.py: full preprocess of P={scene_np['means3D'].shape[0]} ({t_pre:.1f} s) + "
                      f"compositing fwd + autograd bwd (through the preprocess of the {len(ids)} Gaussians involved) over the first <={max_len} list positions of {len(tiles)} "
                      f"of {gx * gy} tiles = {entries} (tile, entry) pairs ({t_tiles:.1f} s); whole view (R={R_total} "
                      f"pairs) extrapolated per pair to {t_view:.0f} s",
            "seconds_measured": round(t_pre + t_tiles, 1)}


def load_impl(impl):
    """-> (module with the reference's `_C` signature, description) or (None, reason)."""
    import harness as hz
    if impl == "new":
        return hz.new_surfel_C(), "libb200gs.so (sm_100a) via diff_surfel_rasterization._C"
    m = hz.ref_module("ref_surfel_C")
    if m is None:
        return None, "oracle/_ref/ref_surfel_C.so not built (python oracle/build_ref.py needs /root/reference)"
    return m, "unmodified reference CUDA extension oracle/_ref/ref_surfel_C.so (built for sm_100a)"


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

    import b200gs
    import harness as hz
    import scenes
    Cmod, impl_desc = load_impl(a.impl)
    if Cmod is None:
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": impl_desc}))
        return

    # ---- synthetic workload: replicated Gaussians, a batch of ring-camera views sharded over the ranks ----
    VB = a.views_per_step
    if VB < world or VB % world:
        raise SystemExit(f"--views-per-step {VB} must be a positive multiple of the number of GPUs ({world})")
    from b200gs.dp import GradArena, shard_views
    my_views = shard_views(VB, rank, world)               # r, r+n, ...  
    sc_np, cam_np = scenes.build(a.workload, P=a.points or None, view=my_views[0] % 32, n_views=32)
    P, W, H = sc_np["means3D"].shape[0], cam_np["width"], cam_np["height"]
    N = W * H
    sc = hz.to_dev(sc_np)
    cams_np = [cam_np] + [scenes.camera(a.workload, v % 32, 32) for v in my_views[1:]]
    cams = [hz.to_dev(c) for c in cams_np]
    cam = cams[0]
    grads = hz.to_dev(scenes.upstream_grads(W, H))
    GRAD_KEYS = ("means3D", "sh", "opacity", "scales", "rotations", "refl")   # the 59 floats / Gaussian of the model

    # The parameter gradients of the views of a batch are summed.  This library: the backward ADDS them into one
    # flat [59 P] arena in place (b200gs.dp.GradArena, accumulate mode), which NCCL then reduces as it lies.
    # Reference arm: what autograd does for it -- one add per gradient tensor per view -- and one concatenation
    # before the all-reduce.
    arena, ref_acc = None, None
    if a.impl == "new":
        arena = GradArena(P, sc["shs"].shape[1], "cuda", accumulate=True)
        Cmod.use_gradient_arena(arena)

    def step():
        nonlocal ref_acc
        if arena is not None:
            arena.zero_()
        f = b = None
        for vi, c in enumerate(cams):
            f = hz.surfel_forward(Cmod, sc, c)
            b = hz.surfel_backward(Cmod, sc, c, f, grads)
            if arena is None:
                if vi == 0:
                    ref_acc = {k: b[k] for k in GRAD_KEYS}
                else:
                    for k in GRAD_KEYS:
                        ref_acc[k] += b[k]
        if world > 1:
            if arena is not None:
                arena.all_reduce()
            else:
                flat = torch.cat([ref_acc[k].reshape(P, -1) for k in GRAD_KEYS], 1)
                dist.all_reduce(flat)
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
    if a.impl == "new":
        b200gs.set_exact_lists(1)
        f = hz.surfel_forward(Cmod, sc, cam)
        R_ref = int(f["rendered"])           # the reference's instance count
        del f
        b200gs.set_exact_lists(1 if a.exact_lists else 0)
    f = hz.surfel_forward(Cmod, sc, cam)
    R, V = int(f["rendered"]), int((f["radii"] > 0).sum())
    if R_ref is None:
        R_ref = R
    del f

    # ---- per-stage table: a separate instrumented pass ------------------------------------------
    # An event pair between every two kernels keeps a launch from overlapping the previous kernel's tail
    # (measured: +0.4 ms on a 6.1 ms view), so the all-stage table comes from its own short pass and the
    # timed region below carries events around the dominant kernel only.
    stages_all, top_stage = {}, -1
    if a.impl == "new" and not a.no_stage_timing:
        step()
        torch.cuda.synchronize()
        b200gs.timing_enable(1)
        b200gs.timing_read()
        step()
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
        b200gs.timing_read()
        n0[0] = b200gs.launch_count()
    with ClockSampler(local) as cs:
        total_ms = timed(step, a.steps, a.warmup, reset_counters)
    launches = b200gs.launch_count() - n0[0]
    stages = b200gs.timing_read() if top_stage >= 0 else {}      # the dominant kernel, live in the timed region
    b200gs.timing_enable(0)
    clocks = cs.summary()
    ms_per_step = total_ms / a.steps
    value = VB * N / (ms_per_step * 1e-3) / 1e6            # whole job: every rank's views

    if arena is not None:
        Cmod.use_gradient_arena(None)
    # ---- end to end through the drop-in Python API, host inputs ------------------------------
    e2e = None
    if not a.no_e2e:
        import diff_surfel_rasterization as dsr
        saved_C = dsr._C
        dsr._C = Cmod                                   # the reference arm drives the reference's _C
        try:
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

            def upload(host):
                """H2D of one view's inputs on the copy stream (a data loader's prefetch); still inside the timed region."""
                with torch.cuda.stream(copy_stream):
                    dev = {k: v.to("cuda", non_blocking=True) for k, v in host.items()}
                    ev = torch.cuda.Event()
                    ev.record(copy_stream)
                return dev, ev

            def e2e_step():
                for t in list(leaves.values()) + [m2d]:
                    t.grad = None
                total = None
                nxt = upload(hosts[0])
                for vi in range(len(hosts)):             # this rank's views of the batch; autograd sums the gradients
                    dev, ev = nxt
                    if vi + 1 < len(hosts):
                        nxt = upload(hosts[vi + 1])      # the next view's inputs travel while this one renders
                    cur = torch.cuda.current_stream()
                    cur.wait_event(ev)
                    for t in dev.values():
                        t.record_stream(cur)
                    rs = dsr.GaussianRasterizationSettings(
                        image_height=H, image_width=W, tanfovx=cam_np["tanfovx"], tanfovy=cam_np["tanfovy"], bg=dev["bg"],
                        scale_modifier=1.0, viewmatrix=dev["viewmatrix"], projmatrix=dev["projmatrix"], sh_degree=3,
                        campos=dev["campos"], prefiltered=False, debug=False)
                    color, radii, allmap, refl_map, _ = dsr.GaussianRasterizer(rs)(
                        means3D=leaves["means3D"], means2D=m2d, opacities=leaves["opacities"], shs=leaves["shs"],
                        refl_strengths=leaves["refl"], scales=leaves["scales"], rotations=leaves["rotations"],
                        env_scope_mask=mask)
                    loss = (color - dev["target"]).abs().mean() + 1e-3 * allmap[:7].mean() + 1e-3 * refl_map.mean()
                    loss.backward()
                    total = loss.detach() if total is None else total + loss.detach()
                if world > 1:
                    flat = torch.cat([leaves[k].grad.reshape(P, -1) for k in leaves], 1)
                    dist.all_reduce(flat)
                loss_host.copy_(total.reshape(1), non_blocking=True)
                torch.cuda.current_stream().synchronize()     # the caller reads the batch loss every step
                return float(loss_host[0])

            e2e_ms = timed(e2e_step, a.steps, a.warmup) / a.steps
            e2e = {"value": VB * N / (e2e_ms * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": e2e_ms, "ms_per_view": e2e_ms * world / VB,
                   "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4,
                   "api": "diff_surfel_rasterization.GaussianRasterizer (autograd) + L1 loss"}
        finally:
            dsr._C = saved_C

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- rank 0: roofline of the dominant kernel, CPU baseline, the JSON line ----------------
    peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)"
    except Exception:  # noqa: BLE001
        pass
    alg = algorithmic_bytes(P, V, R, N)
    roofline, stage_table = None, {}

    def row(name, ms, cnt, steps):
        per = ms / cnt
        return {"ms": round(per, 4), "launches_per_step": cnt / (steps + 0.0),
                "alg_bytes": alg[name], "gbs": round(alg[name] / (per * 1e-3) / 1e9, 1),
                "frac_of_hbm_peak": round(alg[name] / (per * 1e-3) / 1e9 / peak, 4)}
    for name, (ms, cnt) in stages_all.items():           # the separate all-stage pass (1 step)
        if cnt and name in alg:
            stage_table[name] = row(name, ms, cnt, 1)
    live = {name: row(name, ms, cnt, a.steps) for name, (ms, cnt) in stages.items() if cnt and name in alg}
    if live:
        top = max(live, key=lambda k: live[k]["ms"])   # one entry: the dominant kernel, timed inside the timed region
        traffic = None
        try:   # per-launch dram bytes of that kernel from the committed ncu --set full capture, if any
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(top)
        except Exception:  # noqa: BLE001
            pass
        ach = live[top]["gbs"]
        roofline = {"kernel": top, "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s",
                    "frac": round(ach / peak, 4), "traffic": traffic, "peak_source": peak_src,
                    "ms_per_launch": live[top]["ms"], "launches_per_step": live[top]["launches_per_step"],
                    "alg_bytes_per_launch": live[top]["alg_bytes"],
                    "share_of_step": round(live[top]["ms"] * live[top]["launches_per_step"] / ms_per_step, 3),
                    "note": "CUDA events around this kernel inside the timed region; compositing kernels are FP32/MUFU-issue "
                            "bound, not HBM bound (DESIGN.md), the HBM fraction is reported as the contract asks; `stages` "
                            "holds every stage from a separate instrumented pass (all-stage events perturb the step by ~6 %)"}

    cpu = None
    if a.impl == "new" and world == 1 and not a.no_cpu_baseline:
        cpu = cpu_baseline(sc_np, cam_np, a.cpu_sample_tiles, a.cpu_sample_len, R_ref)   # the oracle walks the reference's full lists
    elif a.impl == "reference":
        cpu = {"value": value, "unit": UNIT, "cores": 0, "kind": "reference",
               "sample": "the reference has no CPU implementation of this path (CUDA only): this arm is its "
                         "unmodified CUDA extension on the same B200, same inputs, whole workload"}

    line = {
        "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": round(ms_per_step, 4), "ms_per_view": round(ms_per_step * world / VB, 4),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{a.workload}: room scene, replicated Gaussians, batch of {VB} ring-camera views per step, "
                               f"{VB // world} per GPU (rank r renders views r, r+n, ...), gradients of the batch summed in place",
                   "gaussians": P, "width": W, "height": H, "sh_degree": 3, "visible": V, "instances": R_ref, "instances_emitted": R,
                   "lists": ("exact (reference's lists bit for bit)" if (a.exact_lists or a.impl == "reference")
                             else "culled emission (only instances that can reach alpha >= 1/255)"),
                   "views_per_step": VB, "views_per_gpu": VB // world,
                   "counts": "visible / instances are those of rank 0's first view", "l2": "inputs (>= 0.7 GB of Gaussian data + 0.5-1.3 GB of sort keys per view) exceed the 126 MB L2; no flush",
                   "parallelism": f"view-parallel x{world}, replicated Gaussians" + (", one NCCL all-reduce of the [P,59] fp32 gradient arena per batch" if world > 1 else "")},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
        "roofline": roofline, "cpu_baseline": cpu, "stages": stage_table, "implementation": impl_desc,
    }
    if a.impl == "reference":
        line["impl"] = "reference"
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
