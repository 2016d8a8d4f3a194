"""GPU parity of csrc/ssim.cu through the C ABI and the drop-in `fused_ssim`
package: against the committed golden vectors of the reference's torch ssim() (tests/golden/ssim_ref.npz), the
CPU oracle at other sizes (ragged edges, tiles larger and smaller than the image), and the unmodified reference
fused-ssim extension (oracle/_ref/ref_ssim_C) where it was built.  Tolerances: map 5e-6 abs (values in
[-1, 1]), gradient 1e-5 of its largest entry."""
import os

import numpy as np
import pytest
import torch

import harness as hz

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ssim_ref.npz"))
MAP_ATOL, GRAD_RTOL = 5e-6, 1e-5
C1, C2 = 0.01 ** 2, 0.03 ** 2


def _pair(shape, seed):
    g = torch.Generator().manual_seed(seed)
    img2 = torch.rand(shape, generator=g)
    img1 = (img2 + 0.2 * torch.randn(shape, generator=g)).clamp(0, 1)
    return img1, img2


@pytest.mark.parametrize("tag", ["a", "b", "c"])
def test_against_reference_golden(tag):
    import fused_ssim as fs
    x1 = torch.from_numpy(GOLD[tag + "_img1"]).cuda().requires_grad_(True)
    x2 = torch.from_numpy(GOLD[tag + "_img2"]).cuda()
    m, *_ = fs.fusedssim(C1, C2, x1.detach(), x2, True)
    assert float((m.cpu() - torch.from_numpy(GOLD[tag + "_map"])).abs().max()) <= MAP_ATOL
    val = fs.fused_ssim(x1, x2)
    assert abs(float(val) - float(GOLD[tag + "_ssim"])) <= 1e-6
    val.backward()
    ref = GOLD[tag + "_grad"]
    assert np.abs(x1.grad.cpu().numpy() - ref).max() <= GRAD_RTOL * np.abs(ref).max()


@pytest.mark.parametrize("shape", [(1, 3, 1, 1), (1, 1, 5, 200), (2, 3, 33, 65), (1, 3, 128, 96), (1, 3, 545, 980)])
def test_against_cpu_oracle(shape):
    import fused_ssim as fs
    from oracle import ssim_oracle as so
    img1, img2 = _pair(shape, sum(shape))
    up = torch.randn(shape, generator=torch.Generator().manual_seed(5))
    m, d1, d2, d3 = fs.fusedssim(C1, C2, img1.cuda(), img2.cuda(), True)
    assert np.abs(m.cpu().numpy() - so.ssim_map(img1.numpy(), img2.numpy())).max() <= MAP_ATOL
    g = fs.fusedssim_backward(C1, C2, img1.cuda(), img2.cuda(), up.cuda(), d1, d2, d3)
    ref = so.ssim_backward(img1.numpy(), img2.numpy(), up.numpy())
    assert np.abs(g.cpu().numpy() - ref).max() <= GRAD_RTOL * np.abs(ref).max()


def test_padding_modes_inference_mode_and_errors():
    import b200gs
    import fused_ssim as fs
    from oracle import ssim_oracle as so
    img1, img2 = _pair((1, 3, 40, 50), 3)
    a, b = img1.cuda().requires_grad_(True), img2.cuda()
    v = fs.fused_ssim(a, b, padding="valid")
    assert abs(float(v) - so.ssim_value(img1.numpy(), img2.numpy(), padding="valid")) <= 1e-6
    v.backward()
    up = np.zeros(img1.shape)
    up[..., 5:-5, 5:-5] = 1.0 / (3 * 30 * 40)
    ref = so.ssim_backward(img1.numpy(), img2.numpy(), up)
    assert np.abs(a.grad.cpu().numpy() - ref).max() <= GRAD_RTOL * np.abs(ref).max()
    m, d1, d2, d3 = fs.fusedssim(C1, C2, a.detach(), b, False)          # train=False: no derivative maps
    assert d1.numel() == 0 and abs(float(m.mean()) - so.ssim_value(img1.numpy(), img2.numpy())) <= 1e-6
    with pytest.raises(RuntimeError):
        fs.fusedssim(C1, C2, img1, img2)                                   # host tensors: no CPU path
    with pytest.raises(RuntimeError):
        fs.fusedssim_backward(C1, C2, a.detach(), b, torch.ones_like(b), d1, d2, d3)
    assert b200gs.ssim_forward(C1, C2, 1, 1, 4, 4, None, None, None, None, None, None, None) != 0
    assert fs.fusedssim(C1, C2, a.detach()[:, :, :0], b[:, :, :0])[0].numel() == 0      # empty image


def test_against_reference_extension():
    ref = hz.ref_module("ref_ssim_C")
    if ref is None:
        pytest.skip("oracle/_ref/ref_ssim_C not built")
    import fused_ssim as fs
    img1, img2 = _pair((1, 3, 545, 980), 77)
    a, b = img1.cuda(), img2.cuda()
    up = torch.randn(a.shape, generator=torch.Generator().manual_seed(9)).cuda()
    mr, r1, r2, r3 = ref.fusedssim(C1, C2, a, b, True)
    mn, n1, n2, n3 = fs.fusedssim(C1, C2, a, b, True)
    assert float((mr - mn).abs().max()) <= MAP_ATOL
    for x, y in ((r1, n1), (r2, n2), (r3, n3)):
        assert float((x - y).abs().max()) <= 2e-5 * float(x.abs().max())
    gr = ref.fusedssim_backward(C1, C2, a, b, up, r1, r2, r3)
    gn = fs.fusedssim_backward(C1, C2, a, b, up, n1, n2, n3)
    assert float((gr - gn).abs().max()) <= GRAD_RTOL * float(gr.abs().max())
