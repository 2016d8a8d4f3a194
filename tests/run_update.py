"""GPU harness: the per-Gaussian update step (csrc/update.cu) next to what the reference runs for it --
torch.optim.Adam as built in scene/gaussian_model.py:196-207 (default implementation, and fused=True
for information) and the torch lines of train.py:242-243 + add_densification_stats.

  python tests/run_update.py [P ...] > gpurun_out/update.jsonl

Per size: median ms over 20 single timed steps after 5 warm-ups (CUDA events; includes the host's enqueue
work of that step), ms per step with 20 steps queued back to back, the kernel's own time from the library's
stage events, algorithmic bytes
(28 B per parameter float: read p, g, m, v; write p, m, v -- 59 floats per Gaussian) over time against the
measured HBM copy peak, and the parity gate of the same run (parameters after the timed steps)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from run_configs import timed  # noqa: E402
from test_update_oracle_cpu import LRS, SHAPES  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6553.6, "fallback"


def batch_timed(fn, n=20):
    """ms per call with n calls queued back to back (host enqueue hidden behind the device queue)."""
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def run(P):
    import b200gs
    from b200gs.update import DensifyStats, GaussianAdam, surfel_param_groups
    g = torch.Generator(device="cuda").manual_seed(1)
    params = {k: torch.randn((P,) + s, device="cuda", generator=g) for k, s in SHAPES.items()}
    vis = (torch.rand(P, device="cuda", generator=g) > 0.4).float()
    grads = {k: torch.randn((P,) + s, device="cuda", generator=g) * 1e-3 * vis.view((P,) + (1,) * len(s))
             for k, s in dict(xyz=(3,), sh=(16, 3), opacity=(1,), scaling=(2,), rotation=(4,), refl=(1,)).items()}
    floats = sum(v[0].numel() for v in params.values())
    out = {"P": P, "floats_per_gaussian": floats, "alg_bytes": 28 * floats * P}

    mine = {k: v.clone() for k, v in params.items()}
    opt = GaussianAdam(surfel_param_groups(**mine, lrs=LRS), lr=0.0, eps=1e-15)
    out["new_ms"] = timed(lambda: opt.step(grads))
    out["new_ms_back_to_back"] = batch_timed(lambda: opt.step(grads))
    b200gs.timing_enable(1)
    for _ in range(10):
        opt.step(grads)
    torch.cuda.synchronize()
    st = b200gs.timing_read()["adam"]
    b200gs.timing_enable(0)
    out["kernel_ms"] = st[0] / st[1]
    steps = int(opt.state[mine["xyz"]]["step"])

    for name, kw in (("torch_default", {}), ("torch_fused", {"fused": True})):
        tp = {k: torch.nn.Parameter(v.clone()) for k, v in params.items()}
        tg = dict(grads, f_dc=grads["sh"][:, :1, :].contiguous(), f_rest=grads["sh"][:, 1:, :].contiguous())
        for k in SHAPES:
            tp[k].grad = tg[k]
        topt = torch.optim.Adam([{"params": [tp[k]], "lr": LRS[k], "name": k} for k in SHAPES], lr=0.0, eps=1e-15, **kw)
        out[name + "_ms"] = timed(topt.step)
        out[name + "_ms_back_to_back"] = batch_timed(topt.step)
        for _ in range(10):
            topt.step()
        if name == "torch_default":     # same number of steps on both sides -> compare the parameters
            assert topt.state[tp["xyz"]]["step"] == steps
            out["parity_max_abs_over_lr"] = max(float((mine[k] - tp[k].detach()).abs().max()) / LRS[k] for k in SHAPES)
        del tp, topt
        torch.cuda.empty_cache()
    peak, src = hbm_peak()
    out["achieved_gbps"] = out["alg_bytes"] / (out["kernel_ms"] * 1e-3) / 1e9
    out["hbm_peak_gbps"], out["hbm_peak_source"] = peak, src
    out["hbm_frac"] = out["achieved_gbps"] / peak
    out["speedup_vs_torch_default"] = out["torch_default_ms_back_to_back"] / out["new_ms_back_to_back"]

    # densification statistics
    radii = (torch.randint(0, 60, (P,), device="cuda", generator=g) * (vis > 0)).int()
    vgrad = torch.randn((P, 3), device="cuda", generator=g) * 1e-3
    rw = torch.rand(P, device="cuda", generator=g) * vis
    ds = DensifyStats(P)
    out["stats_new_ms"] = timed(lambda: ds.add(vgrad, radii, rw))
    mr = torch.zeros(P, device="cuda")
    acc, den, aw, dw = (torch.zeros((P, 1), device="cuda") for _ in range(4))

    def ref_stats():
        v = radii > 0
        mr[v] = torch.max(mr[v], radii[v])
        acc[v] += torch.norm(vgrad[v], dim=-1, keepdim=True)
        den[v] += 1
        m = rw > 0.0
        aw[m] += rw[m].unsqueeze(-1)
        dw[m] += 1
    out["stats_torch_ms"] = timed(ref_stats)
    out["stats_alg_bytes"] = (4 + 12 + 4 + 5 * 8) * P
    out["stats_gbps"] = out["stats_alg_bytes"] / (out["stats_new_ms"] * 1e-3) / 1e9
    out["stats_equal"] = bool(torch.equal(ds.denom, den.reshape(-1)) and torch.equal(ds.max_radii2D, mr))
    return out


if __name__ == "__main__":
    for P in [int(x) for x in sys.argv[1:]] or [3_000_000, 6_000_000]:
        print(json.dumps(run(P)), flush=True)
        torch.cuda.empty_cache()
