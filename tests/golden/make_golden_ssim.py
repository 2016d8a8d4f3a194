#!/usr/bin/env python
"""Generate tests/golden/ssim_ref.npz from the reference's OWN torch implementation of the loss:
utils/loss_utils.py::ssim (lines 63-101) imported from /root/reference, CPU, with autograd for the
gradient.  Runs in the build container (the Python reference does not travel to the GPU box):

    python tests/golden/make_golden_ssim.py

The fixture pins oracle/ssim_oracle.py (tests/test_ssim_oracle_cpu.py) and is what the GPU kernel is compared
with in tests/test_ssim_gpu.py.
"""
import importlib.util
import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("B200GS_REFERENCE_ROOT", "/root/reference")


def main():
    spec = importlib.util.spec_from_file_location("ref_loss_utils", os.path.join(REF, "utils", "loss_utils.py"))
    lu = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(lu)
    out = {}
    for tag, (B, CH, H, W) in dict(a=(1, 3, 37, 53), b=(2, 1, 64, 64), c=(1, 3, 9, 70)).items():
        g = torch.Generator().manual_seed(hash(tag) % 1000 if False else {"a": 11, "b": 22, "c": 33}[tag])
        img2 = torch.rand((B, CH, H, W), generator=g)
        img1 = (img2 + 0.2 * torch.randn((B, CH, H, W), generator=g)).clamp(0, 1).requires_grad_(True)
        window = lu.create_window(11, CH)
        # size_average=True path of _ssim is map.mean(); rebuild the map with the reference's own pieces
        val = lu._ssim(img1, img2, window, 11, CH, True)
        val.backward()
        out[tag + "_img1"], out[tag + "_img2"] = img1.detach().numpy(), img2.numpy()
        out[tag + "_ssim"] = np.float64(val.item())
        out[tag + "_grad"] = img1.grad.numpy().copy()          # d mean(map) / d img1
        # the map itself (not returned by the reference function): same expressions, lines 81-94
        with torch.no_grad():
            F = torch.nn.functional
            mu1, mu2 = F.conv2d(img1, window, padding=5, groups=CH), F.conv2d(img2, window, padding=5, groups=CH)
            s1 = F.conv2d(img1 * img1, window, padding=5, groups=CH) - mu1.pow(2)
            s2 = F.conv2d(img2 * img2, window, padding=5, groups=CH) - mu2.pow(2)
            s12 = F.conv2d(img1 * img2, window, padding=5, groups=CH) - mu1 * mu2
            out[tag + "_map"] = (((2 * mu1 * mu2 + lu.C1) * (2 * s12 + lu.C2)) / ((mu1.pow(2) + mu2.pow(2) + lu.C1) * (s1 + s2 + lu.C2))).numpy()
        assert abs(float(out[tag + "_map"].mean()) - float(val)) < 1e-6
    np.savez_compressed(os.path.join(HERE, "ssim_ref.npz"), **out)
    print("wrote", os.path.join(HERE, "ssim_ref.npz"), {k: v.shape for k, v in out.items() if hasattr(v, "shape")})


if __name__ == "__main__":
    main()
