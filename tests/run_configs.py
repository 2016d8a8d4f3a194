"""GPU harness: BASELINE.json's configs c1..c5 on one B200 -- this library (culled emission and
exact-lists mode) next to the unmodified reference extension from oracle/_ref, same inputs.

  python tests/run_configs.py [c1 c2 c3 c4 c5 c3g] > gpurun_out/configs.json

Per config: median fwd+bwd ms over 20 timed iterations after 5 warm-ups (CUDA events), the parity
gates on the same run (radii, images, gradients), and the workload counts P, V, R that make the
algorithmic-bytes formula of SURVEY.md section 8(d) reproducible.  c2 adds the cubemap environment
lookup (640 k rays, L=128, C=3) + its backward to the step; c5 is 32 views of the 6M scene;
c3g is c3 through the 3DGS (EWA) rasterizer.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import harness as hz  # noqa: E402
import scenes  # noqa: E402


def timed(fn, warmup=5, iters=20):
    for _ in range(warmup):
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


def cubemap_step(mod, B=640_000, C=3, L=128, seed=0):
    """-> callable doing one seamless-bilinear lookup forward + backward through `mod`
    (an object with the reference's _cubemapencoder signature)."""
    g = torch.Generator().manual_seed(20261104 + seed)
    d = torch.randn((B, 3), generator=g)
    d = (d / d.norm(dim=1, keepdim=True)).cuda().contiguous()
    tex = (torch.rand((6, C, L, L), generator=g) - 0.5).cuda().contiguous()
    fail = torch.zeros(C, device="cuda")
    out = torch.empty((C, B), device="cuda")
    gout = torch.randn((C, B), generator=g).cuda().contiguous()
    gtex, gin, gfail = torch.zeros_like(tex), torch.zeros_like(d), torch.zeros_like(fail)

    def step():
        mod.cubemap_encode_forward(d, tex, fail, out, 1, 1, B, C, L)
        gtex.zero_()
        mod.cubemap_encode_backward(gout, d, tex, gtex, gin, gfail, 1, 1, B, C, L)
    return step, out, gtex


def surfel_config(name, n_views=1, with_cubemap=False):
    import b200gs
    newC, refC = hz.new_surfel_C(), hz.ref_module("ref_surfel_C")
    out = {"config": name, "path": "surfel", "views": n_views}
    sc_np, cam0 = scenes.build(name, view=0, n_views=32)
    sc = hz.to_dev(sc_np)
    P, W, H = sc_np["means3D"].shape[0], cam0["width"], cam0["height"]
    cams = [hz.to_dev(scenes.camera(name, v, 32)) if v else hz.to_dev(cam0) for v in range(n_views)]
    grads = hz.to_dev(scenes.upstream_grads(W, H))
    out.update(P=P, W=W, H=H)

    cm_new = cm_ref = None
    if with_cubemap:
        import _cubemapencoder as new_cm
        cm_new = cubemap_step(new_cm)
        ref_cm = hz.ref_module("ref_cubemap_C")
        cm_ref = cubemap_step(ref_cm) if ref_cm is not None else None

    def make(Cmod, cm):
        def step():
            last = None
            for cam in cams:
                f = hz.surfel_forward(Cmod, sc, cam)
                b = hz.surfel_backward(Cmod, sc, cam, f, grads)
                last = (f, b)
            if cm is not None:
                cm[0]()
            return last
        return step

    with hz.exact_lists(True):
        f_ex, b_ex = make(newC, None)()
        out["R"], out["V"] = int(f_ex["rendered"]), int((f_ex["radii"] > 0).sum())
        out["new_exact_ms"] = timed(make(newC, cm_new))
    f_cu, b_cu = make(newC, None)()
    out["R_emitted"] = int(f_cu["rendered"])
    out["new_ms"] = timed(make(newC, cm_new))
    if refC is not None:
        f_r, b_r = make(refC, None)()
        out["ref_ms"] = timed(make(refC, cm_ref), warmup=3, iters=10)
        out["speedup"] = out["ref_ms"] / out["new_ms"]
        out["parity"] = {
            "rendered_equal_exact_mode": out["R"] == int(f_r["rendered"]),
            "radii_equal": bool(torch.equal(f_cu["radii"], f_r["radii"])),
            "image_max_abs": max(hz.max_abs(f_cu[k], f_r[k]) for k in ("color", "others", "refl_map")),
            "image_bit_identical": all(bool(torch.equal(f_cu[k], f_r[k])) for k in ("color", "others", "refl_map")),
            "grad_max_rel_l2": max(hz.rel_l2(b_cu[k], b_r[k]) for k in b_cu),
        }
    N = W * H
    out["mpix_per_s"] = n_views * N / (out["new_ms"] * 1e-3) / 1e6
    out["views_per_s"] = n_views / (out["new_ms"] * 1e-3)
    out["alg_bytes_per_view"] = 722 * P + 294 * out["V"] + 168 * out["R"] + 136 * N
    out["hbm_frac_of_alg_bytes"] = out["alg_bytes_per_view"] * n_views / (out["new_ms"] * 1e-3) / 1e9 / 6560.3
    return out


def gauss_config(name):
    newC, refC = hz.new_gauss_C(), hz.ref_module("ref_gauss_C")
    sc_np, cam_np = scenes.build(name)
    sc, cam = hz.gauss_scene(hz.to_dev(sc_np)), hz.to_dev(cam_np)
    P, W, H = sc_np["means3D"].shape[0], cam_np["width"], cam_np["height"]
    grads = hz.to_dev(hz.gauss_grads(W, H))
    out = {"config": name + "g", "path": "3dgs", "views": 1, "P": P, "W": W, "H": H}

    def make(Cmod):
        def step():
            f = hz.gauss_forward(Cmod, sc, cam)
            return f, hz.gauss_backward(Cmod, sc, cam, f, grads)
        return step
    with hz.exact_lists(True):
        f_ex, _ = make(newC)()
        out["R"], out["V"] = int(f_ex["rendered"]), int((f_ex["radii"] > 0).sum())
        out["new_exact_ms"] = timed(make(newC))
    f_n, b_n = make(newC)()
    out["R_emitted"] = int(f_n["rendered"])
    out["new_ms"] = timed(make(newC))
    if refC is not None:
        f_r, b_r = make(refC)()
        out["ref_ms"] = timed(make(refC), warmup=3, iters=10)
        out["speedup"] = out["ref_ms"] / out["new_ms"]
        out["parity"] = {
            "rendered_equal_exact_mode": out["R"] == int(f_r["rendered"]),
            "radii_equal": bool(torch.equal(f_n["radii"], f_r["radii"])),
            "image_max_abs": max(hz.max_abs(f_n[k], f_r[k]) for k in ("color", "invdepth", "normal_map", "refl_map")),
            "image_bit_identical": all(bool(torch.equal(f_n[k], f_r[k])) for k in ("color", "invdepth", "normal_map", "refl_map")),
            "grad_max_rel_l2": max(hz.rel_l2(b_n[k], b_r[k]) for k in b_n),
        }
    out["mpix_per_s"] = W * H / (out["new_ms"] * 1e-3) / 1e6
    return out


def main(argv):
    names = argv or ["c1", "c2", "c3", "c4", "c3g", "c5"]
    res = []
    for n in names:
        if n.endswith("g"):
            r = gauss_config(n[:-1])
        else:
            r = surfel_config(n, n_views=32 if n == "c5" else 1, with_cubemap=(n == "c2"))
        print(json.dumps(r), flush=True)
        res.append(r)
        torch.cuda.empty_cache()
    return res


if __name__ == "__main__":
    main(sys.argv[1:])
