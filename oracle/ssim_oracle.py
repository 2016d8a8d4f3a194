"""CPU restatement of the SSIM image loss -- the ORACLE for csrc/ssim.cu ("next" row 3).

TEST INFRASTRUCTURE ONLY.  Nothing in the product package imports this file; only tests/ may.

What is restated: ssim() / _ssim() of the reference, utils/loss_utils.py:63-101 (11-tap Gaussian window of
sigma 1.5 built by gaussian() :47-49, zero padding of window_size // 2, C1 = 0.01^2, C2 = 0.03^2), which is the
map the fused CUDA kernel computes (submodules/fused-ssim/ssim.cu:187-286) -- written here as two separable
1-D passes in fp64-accumulated numpy instead of a grouped conv2d, plus the closed-form gradient with respect to
the first image (the three partial-derivative maps of ssim.cu:275-283 filtered with the same window).

Pinned: tests/golden/ssim_ref.npz holds the output of the reference's own torch ssim() (imported from
/root/reference/utils/loss_utils.py, CPU) and its autograd gradient on seeded images, written by
tests/golden/make_golden_ssim.py; tests/test_ssim_oracle_cpu.py compares this file's functions with it.
"""
from math import exp

import numpy as np

C1, C2 = 0.01 ** 2, 0.03 ** 2


def window(size=11, sigma=1.5):
    g = np.array([exp(-(x - size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(size)], dtype=np.float64)
    return (g / g.sum()).astype(np.float32).astype(np.float64)       # the reference holds it in fp32


def blur(x, g=None):
    """'same' zero-padded separable filter over the last two axes of x [..., H, W], fp64."""
    g = window() if g is None else g
    r = len(g) // 2
    x = np.asarray(x, dtype=np.float64)
    H, W = x.shape[-2:]
    xp = np.zeros(x.shape[:-1] + (W + 2 * r,))
    xp[..., r:r + W] = x
    hx = sum(g[i] * xp[..., i:i + W] for i in range(len(g)))
    yp = np.zeros(x.shape[:-2] + (H + 2 * r, W))
    yp[..., r:r + H, :] = hx
    return sum(g[i] * yp[..., i:i + H, :] for i in range(len(g)))


def ssim_terms(img1, img2):
    x1, x2 = np.asarray(img1, np.float64), np.asarray(img2, np.float64)
    mu1, mu2 = blur(x1), blur(x2)
    s1, s2, s12 = blur(x1 * x1) - mu1 * mu1, blur(x2 * x2) - mu2 * mu2, blur(x1 * x2) - mu1 * mu2
    A, B = mu1 * mu1 + mu2 * mu2 + C1, s1 + s2 + C2
    C, D = 2 * mu1 * mu2 + C1, 2 * s12 + C2
    return mu1, mu2, A, B, C, D


def ssim_map(img1, img2):
    """[B, CH, H, W] -> map of the same shape (utils/loss_utils.py:94)."""
    _, _, A, B, C, D = ssim_terms(img1, img2)
    return (C * D) / (A * B)


def ssim_value(img1, img2, padding="same"):
    m = ssim_map(img1, img2)
    if padding == "valid":
        m = m[..., 5:-5, 5:-5]
    return float(m.mean())


def derivative_maps(img1, img2):
    """dm/dmu1, dm/dsigma1_sq, dm/dsigma12 (ssim.cu:275-283)."""
    mu1, mu2, A, B, C, D = ssim_terms(img1, img2)
    dmu1 = (mu2 * 2 * D) / (A * B) - (mu2 * 2 * C) / (A * B) - (mu1 * 2 * C * D) / (A * A * B) + (mu1 * 2 * C * D) / (A * B * B)
    return dmu1, (-C * D) / (A * B * B), (2 * C) / (A * B)


def ssim_backward(img1, img2, dL_dmap):
    """dL/dimg1 for an upstream gradient on the map: the window is symmetric, so the adjoint of the filter is
    the filter (ssim.cu:288-365)."""
    dmu1, ds1, ds12 = derivative_maps(img1, img2)
    dL = np.asarray(dL_dmap, np.float64)
    x1, x2 = np.asarray(img1, np.float64), np.asarray(img2, np.float64)
    return blur(dL * dmu1) + 2 * x1 * blur(dL * ds1) + x2 * blur(dL * ds12)
