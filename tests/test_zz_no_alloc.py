"""b200gs_surfel_backward_static (include/b200gs.h, csrc/static_scratch.cu): the surfel backward with its per-Gaussian
accumulator in a static __device__ array inside the library, for callers that cannot allocate.  Called here straight
through the C ABI (ctypes), twice in a row, and compared with b200gs_surfel_backward on the same forward state.

The two entry points run the same kernels; what differs between runs is the order of the floating-point atomics of
the compositing backward, so the bar is the one the other tests use for run-to-run comparisons of this library
(rel L2 <= 1e-5, tests/test_surfel_gpu.py) -- far inside the north_star's 1e-4 relative for gradients.

Collected last on purpose (file name).
"""
import ctypes as C

import pytest
import torch

import harness as hz
import scenes

pytestmark = pytest.mark.gpu

RUN_TO_RUN = 1e-5
OUT_WIDTHS = dict(means2D=3, normal=3, opacity=1, colors=3, refl=1, means3D=3, transMat=9, sh=None, scales=2, rotations=4)


def _outputs(P, M, dev):
    """zero-filled gradient buffers of one backward call (the caller's allocation, made BEFORE the call)"""
    o = {}
    for k, w in OUT_WIDTHS.items():
        o[k] = torch.zeros((P, 3 * M if w is None else w), dtype=torch.float32, device=dev)
    return o


def _args(b, sc, cam, f, grads, out, accum):
    P, M = sc["means3D"].shape[0], sc["shs"].shape[1]
    H, W = cam["height"], cam["width"]
    return b.SurfelBwd(
        P=P, sh_degree=3, sh_coeffs=M, num_rendered=int(f["rendered"]), width=W, height=H,
        tan_fovx=float(cam["tanfovx"]), tan_fovy=float(cam["tanfovy"]), scale_modifier=1.0, debug=0,
        background=b.ptr(f["bg"]), means3D=b.ptr(sc["means3D"]), shs=b.ptr(sc["shs"]), colors_precomp=None,
        scales=b.ptr(sc["scales"]), rotations=b.ptr(sc["rotations"]), transMat_precomp=None,
        viewmatrix=b.ptr(cam["viewmatrix"]), projmatrix=b.ptr(cam["projmatrix"]), campos=b.ptr(cam["campos"]),
        radii=b.ptr(f["radii"]), geom=b.ptr(f["geom"]), binning=b.ptr(f["binning"]), img=b.ptr(f["img"]),
        dL_dout_color=b.ptr(grads["color"]), dL_dout_others=b.ptr(grads["others"]), dL_dout_refl=b.ptr(grads["refl"]),
        accum=(accum.data_ptr() if accum is not None else None), accum_bytes=(accum.numel() if accum is not None else 0),
        dL_dmeans2D=b.ptr(out["means2D"]), dL_dnormal=b.ptr(out["normal"]), dL_dopacity=b.ptr(out["opacity"]),
        dL_dcolors=b.ptr(out["colors"]), dL_drefl=b.ptr(out["refl"]), dL_dmeans3D=b.ptr(out["means3D"]),
        dL_dtransMat=b.ptr(out["transMat"]), dL_dsh=b.ptr(out["sh"]), dL_dscales=b.ptr(out["scales"]),
        dL_drotations=b.ptr(out["rotations"]), accumulate=0,
        geom_bytes=f["geom"].numel(), binning_bytes=f["binning"].numel(), img_bytes=f["img"].numel())


def _setup(name):
    import b200gs as b
    sc_np, cam_np = scenes.build(name)
    sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
    grads = hz.to_dev(scenes.upstream_grads(cam_np["width"], cam_np["height"]))
    grads = {k: v.contiguous() for k, v in grads.items()}
    f = hz.surfel_forward(hz.new_surfel_C(), sc, cam)
    f["radii"] = f["radii"].contiguous()
    return b, sc, cam, grads, f


@pytest.mark.parametrize("name", ["small", "c1"])
def test_static_scratch_backward_matches_the_allocating_one_twice(name):
    b, sc, cam, grads, f = _setup(name)
    dev = sc["means3D"].device
    P, M = sc["means3D"].shape[0], sc["shs"].shape[1]
    assert P <= b.STATIC_ACCUM_MAX_POINTS
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    want = _outputs(P, M, dev)
    accum = torch.empty((b.surfel_accum_bytes(P),), dtype=torch.uint8, device=dev)
    b.check(b.surfel_backward(C.byref(_args(b, sc, cam, f, grads, want, accum)), stream))
    torch.cuda.synchronize()
    assert float(want["sh"].abs().sum()) > 0 and float(want["means3D"].abs().sum()) > 0

    for call in range(2):                      # two calls in a row: the second finds the first one's sums in the array
        got = _outputs(P, M, dev)
        a = _args(b, sc, cam, f, grads, got, None)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        b.check(b.surfel_backward_static(C.byref(a), stream))
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before
        for k in want:
            assert torch.isfinite(got[k]).all(), (call, k)
            assert hz.rel_l2(got[k], want[k]) <= RUN_TO_RUN, (call, k, hz.rel_l2(got[k], want[k]))
            assert hz.grad_close_fraction(got[k], want[k], rtol=1e-4) >= 0.999, (call, k)


def test_static_scratch_rejects_more_points_than_the_array_holds():
    import b200gs as b
    too_many = b.STATIC_ACCUM_MAX_POINTS + 1
    a = b.SurfelBwd(P=too_many, width=16, height=16)
    rc = b.surfel_backward_static(C.byref(a), None)
    assert rc != 0
    msg = b.last_error().decode()
    assert "exceeds" in msg and str(b.STATIC_ACCUM_MAX_POINTS) in msg, msg
    with pytest.raises(RuntimeError):
        b.check(rc)
