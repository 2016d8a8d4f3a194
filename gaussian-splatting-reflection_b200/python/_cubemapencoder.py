"""`_cubemapencoder` -- the two entry points of the reference's pybind module
(C/src/bindings.cpp:5-8, C/src/cubemapencoder.h:6-17) over the C ABI of libb200gs.so.

Both functions work IN PLACE on caller-allocated tensors and return None, like the
reference.  Checks mirror cubemapencoder.cu:435-448 / :719-738 (CUDA, contiguous,
floating).  The kernels are fp32 (the reference's backward is float-only, :746, and its
autograd wrapper casts the forward to fp32, cubemap_encoder.py:21); half / double
textures are converted around the call."""
import ctypes as C

import torch

import b200gs as _b

_FLOATING = (torch.float32, torch.float16, torch.float64)


def _check(name, t):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be a contiguous tensor")
    if t.dtype not in _FLOATING:
        raise RuntimeError(f"{name} must be a floating tensor")


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def cubemap_encode_forward(inputs, cubemap, fail_value, outputs, interp, seamless, B, Cn, L):
    """outputs[C,B] <- lookup(inputs[B,3], cubemap[6,C,L,L]); zero directions give fail_value."""
    for n, t in (("inputs", inputs), ("cubemap", cubemap), ("fail_value", fail_value), ("outputs", outputs)):
        _check(n, t)
    if B == 0:
        return
    f32 = torch.float32
    out = outputs if outputs.dtype == f32 else torch.empty_like(outputs, dtype=f32)
    i, c, f = (t if t.dtype == f32 else t.to(f32) for t in (inputs, cubemap, fail_value))
    with torch.cuda.device(cubemap.device):
        _b.check(_b.cubemap_forward(i.data_ptr(), c.data_ptr(), f.data_ptr(), out.data_ptr(), int(interp), int(seamless),
                                    int(B), int(Cn), int(L), _stream()))
    if out is not outputs:
        outputs.copy_(out)


def cubemap_encode_backward(grad_outputs, inputs, cubemap, grad_cubemap, grad_inputs, grad_fail, interp, seamless, B, Cn, L):
    """grad_cubemap / grad_fail (zero-filled by the caller) += ..., grad_inputs[B,3] <- ... ; float32 only."""
    for n, t in (("grad_outputs", grad_outputs), ("inputs", inputs), ("cubemap", cubemap), ("grad_cubemap", grad_cubemap),
                 ("grad_inputs", grad_inputs), ("grad_fail", grad_fail)):
        _check(n, t)
        if t.dtype != torch.float32:
            raise RuntimeError(f"expected scalar type Float but found {t.dtype}")   # .data<float>() in the reference
    if B == 0:
        return
    with torch.cuda.device(cubemap.device):
        _b.check(_b.cubemap_backward(grad_outputs.data_ptr(), inputs.data_ptr(), cubemap.data_ptr(), grad_cubemap.data_ptr(),
                                     grad_inputs.data_ptr(), grad_fail.data_ptr(), int(interp), int(seamless), int(B),
                                     int(Cn), int(L), _stream()))
