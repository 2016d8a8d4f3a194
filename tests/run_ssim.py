"""GPU harness: the SSIM loss kernels (csrc/ssim.cu) next to what the reference runs for them -- the fused-ssim
CUDA extension (oracle/_ref/ref_ssim_C, unmodified) and the torch conv2d ssim() of utils/loss_utils.py:63-101
restated with torch ops on the device.

  python tests/run_ssim.py [H W ...] > gpurun_out/ssim.jsonl

Per size: kernel time from the library's stage events and ms per fwd+bwd call queued back to back (CUDA events),
algorithmic bytes (forward 24 B, backward 28 B per pixel and channel) against the measured HBM copy peak, and
the parity gate of the same run."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import harness as hz  # noqa: E402
from run_update import batch_timed, hbm_peak  # noqa: E402

C1, C2 = 0.01 ** 2, 0.03 ** 2


def torch_ssim(img1, img2, window):
    F = torch.nn.functional
    ch = img1.shape[1]
    mu1, mu2 = F.conv2d(img1, window, padding=5, groups=ch), F.conv2d(img2, window, padding=5, groups=ch)
    s1 = F.conv2d(img1 * img1, window, padding=5, groups=ch) - mu1.pow(2)
    s2 = F.conv2d(img2 * img2, window, padding=5, groups=ch) - mu2.pow(2)
    s12 = F.conv2d(img1 * img2, window, padding=5, groups=ch) - mu1 * mu2
    return (((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1.pow(2) + mu2.pow(2) + C1) * (s1 + s2 + C2))).mean()


def run(H, W, CH=3):
    import b200gs
    import fused_ssim as fs
    from oracle import ssim_oracle as so
    g = torch.Generator(device="cuda").manual_seed(1)
    img2 = torch.rand((1, CH, H, W), device="cuda", generator=g)
    img1 = (img2 + 0.2 * torch.randn((1, CH, H, W), device="cuda", generator=g)).clamp(0, 1)
    up = torch.full_like(img1, 1.0 / img1.numel())
    out = {"H": H, "W": W, "CH": CH}

    def mine():
        m, d1, d2, d3 = fs.fusedssim(C1, C2, img1, img2, True)
        return m, fs.fusedssim_backward(C1, C2, img1, img2, up, d1, d2, d3)
    m_new, g_new = mine()
    out["new_ms"] = batch_timed(mine)
    b200gs.timing_enable(1)
    b200gs.timing_read()
    for _ in range(10):
        mine()
    torch.cuda.synchronize()
    st = b200gs.timing_read()
    b200gs.timing_enable(0)
    out["fwd_kernel_ms"], out["bwd_kernel_ms"] = st["ssim_fwd"][0] / st["ssim_fwd"][1], st["ssim_bwd"][0] / st["ssim_bwd"][1]
    n = CH * H * W
    peak, src = hbm_peak()
    out["fwd_gbps"], out["bwd_gbps"] = 24 * n / out["fwd_kernel_ms"] / 1e6, 28 * n / out["bwd_kernel_ms"] / 1e6
    out["fwd_hbm_frac"], out["bwd_hbm_frac"], out["hbm_peak_gbps"] = out["fwd_gbps"] / peak, out["bwd_gbps"] / peak, peak

    ref = hz.ref_module("ref_ssim_C")
    if ref is not None:
        def theirs():
            m, d1, d2, d3 = ref.fusedssim(C1, C2, img1, img2, True)
            return m, ref.fusedssim_backward(C1, C2, img1, img2, up, d1, d2, d3)
        m_ref, g_ref = theirs()
        out["ref_ext_ms"] = batch_timed(theirs)
        out["speedup_vs_ref_ext"] = out["ref_ext_ms"] / out["new_ms"]
        out["map_max_abs_vs_ref_ext"] = float((m_ref - m_new).abs().max())
        out["grad_max_rel_vs_ref_ext"] = float((g_ref - g_new).abs().max() / g_ref.abs().max())

    window = torch.from_numpy(so.window()).float().cuda()
    window = (window[:, None] @ window[None, :])[None, None].expand(CH, 1, 11, 11).contiguous()
    x = img1.clone().requires_grad_(True)

    def torch_path():
        x.grad = None
        torch_ssim(x, img2, window).backward()
    torch_path()
    out["torch_ms"] = batch_timed(torch_path)
    out["grad_max_rel_vs_torch"] = float((x.grad - g_new).abs().max() / x.grad.abs().max())
    return out


if __name__ == "__main__":
    a = [int(v) for v in sys.argv[1:]]
    sizes = list(zip(a[::2], a[1::2])) or [(1080, 1920), (822, 1237), (800, 800)]
    for H, W in sizes:
        print(json.dumps(run(H, W)), flush=True)
