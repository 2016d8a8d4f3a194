"""Drop-in for the reference's `fused_ssim` package (submodules/fused-ssim/fused_ssim/__init__.py:1-47),
on the sm_100a kernels of csrc/ssim.cu through the C ABI (include/b200gs.h: b200gs_ssim_forward /
b200gs_ssim_backward).  Same names, arguments and padding modes; train.py:168-169 calls
`fused_ssim(image.unsqueeze(0), gt_image.unsqueeze(0))` unchanged.  No CPU path."""
import torch

import b200gs as _b

allowed_padding = ["same", "valid"]


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _check_img(name, t):
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dim() == 4):
        raise RuntimeError(f"{name} must be a CUDA tensor [B, C, H, W] (there is no CPU path)")
    return t.contiguous().float()


def fusedssim(C1, C2, img1, img2, train=True):
    """-> (ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12) like fused_ssim_cuda.fusedssim (ssim.h:5-13);
    the derivative maps are empty tensors when train is False."""
    img1, img2 = _check_img("img1", img1), _check_img("img2", img2)
    if img1.shape != img2.shape:
        raise RuntimeError("img1 and img2 must have the same shape")
    B, CH, H, W = img1.shape
    ssim_map = torch.empty_like(img1)
    if train:
        d1, d2, d3 = torch.empty_like(img1), torch.empty_like(img1), torch.empty_like(img1)
    else:
        d1 = d2 = d3 = torch.empty((0,), dtype=torch.float32, device=img1.device)
    _b.check(_b.ssim_forward(float(C1), float(C2), B, CH, H, W, _b.ptr(img1), _b.ptr(img2), _b.ptr(ssim_map),
                             _b.ptr(d1), _b.ptr(d2), _b.ptr(d3), _stream()))
    return ssim_map, d1, d2, d3


def fusedssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12):
    """-> dL_dimg1 like fused_ssim_cuda.fusedssim_backward (ssim.h:15-26)."""
    img1, img2 = _check_img("img1", img1), _check_img("img2", img2)
    B, CH, H, W = img1.shape
    if dm_dmu1.numel() != img1.numel():
        raise RuntimeError("fusedssim_backward needs the derivative maps of a forward run with train=True")
    dL_dmap = _check_img("dL_dmap", dL_dmap)
    out = torch.empty_like(img1)
    _b.check(_b.ssim_backward(B, CH, H, W, _b.ptr(img1), _b.ptr(img2), _b.ptr(dL_dmap), _b.ptr(dm_dmu1.contiguous()),
                              _b.ptr(dm_dsigma1_sq.contiguous()), _b.ptr(dm_dsigma12.contiguous()), _b.ptr(out), _stream()))
    return out


class FusedSSIMMap(torch.autograd.Function):
    @staticmethod
    def forward(ctx, C1, C2, img1, img2, padding="same", train=True):
        ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12 = fusedssim(C1, C2, img1, img2, train)
        if padding == "valid":
            ssim_map = ssim_map[:, :, 5:-5, 5:-5]
        ctx.save_for_backward(img1.detach(), img2, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)
        ctx.C1, ctx.C2, ctx.padding = C1, C2, padding
        return ssim_map

    @staticmethod
    def backward(ctx, opt_grad):
        img1, img2, dm_dmu1, dm_dsigma1_sq, dm_dsigma12 = ctx.saved_tensors
        dL_dmap = opt_grad
        if ctx.padding == "valid":
            dL_dmap = torch.zeros_like(img1)
            dL_dmap[:, :, 5:-5, 5:-5] = opt_grad
        grad = fusedssim_backward(ctx.C1, ctx.C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)
        return None, None, grad, None, None, None


def fused_ssim(img1, img2, padding="same", train=True):
    C1 = 0.01 ** 2
    C2 = 0.03 ** 2
    assert padding in allowed_padding
    return FusedSSIMMap.apply(C1, C2, img1, img2, padding, train).mean()
