"""GPU parity of the 3DGS (EWA) path against the UNMODIFIED reference extension
(oracle/_ref/ref_gauss_C) on identical synthetic scenes.  Bars as for the surfel path: integer
structures bit-exact, images <= 1e-5 abs (measured: bit-identical), gradients <= 1e-4 relative."""
import pytest
import torch

import harness as hz
import scenes

pytestmark = pytest.mark.gpu
IMG_ATOL, GRAD_RTOL = 1e-5, 1e-4
# exact lists: the reference's instance lists bit for bit; culled emission (default): only the
# instances whose tile the Gaussian can reach alpha >= 1/255 in (include/b200gs.h: b200gs_set_exact_lists)
MODES = [pytest.param(True, id="exact-lists"), pytest.param(False, id="culled-emission")]


@pytest.fixture(scope="module")
def refC():
    m = hz.ref_module("ref_gauss_C")
    if m is None:
        pytest.skip("oracle/_ref/ref_gauss_C not built")
    return m


def _setup(name):
    sc_np, cam_np = scenes.build(name)
    sc, cam = hz.gauss_scene(hz.to_dev(sc_np)), hz.to_dev(cam_np)
    grads = hz.to_dev(hz.gauss_grads(cam_np["width"], cam_np["height"]))
    return sc, cam, grads


@pytest.mark.parametrize("name,aa", [("tiny", False), ("small", True), ("c1", False)])
def test_forward_state_and_images(name, aa, refC):
    sc, cam, _ = _setup(name)
    P, W, H = sc["means3D"].shape[0], cam["width"], cam["height"]
    fr = hz.gauss_forward(refC, sc, cam, antialiasing=aa, bg=(0.2, 0.4, 0.6))
    with hz.exact_lists(True):
        fn = hz.gauss_forward(hz.new_gauss_C(), sc, cam, antialiasing=aa, bg=(0.2, 0.4, 0.6))
        assert fn["rendered"] == fr["rendered"]
        assert torch.equal(fn["radii"], fr["radii"])
        sn, sr = hz.new_gauss_state(fn, P, W, H), hz.ref_gauss_state(fr, P, W, H)
        for k in ("tiles_touched", "point_offsets", "keys", "point_list", "ranges", "n_contrib"):
            assert torch.equal(sn[k], sr[k]), k
        for k in ("color", "invdepth", "normal_map", "refl_map"):
            assert hz.max_abs(fn[k], fr[k]) <= IMG_ATOL, k
    with hz.exact_lists(False):
        fc = hz.gauss_forward(hz.new_gauss_C(), sc, cam, antialiasing=aa, bg=(0.2, 0.4, 0.6))
        assert 0 < fc["rendered"] <= fr["rendered"]
        assert torch.equal(fc["radii"], fr["radii"])
        for k in ("color", "invdepth", "normal_map", "refl_map"):   # same contributors, same order: same bits
            assert torch.equal(fc[k], fn[k]), k


@pytest.mark.parametrize("exact", MODES)
@pytest.mark.parametrize("name,aa", [("tiny", False), ("small", True), ("c1", False)])
def test_gradients(name, aa, exact, refC):
    sc, cam, grads = _setup(name)
    newC = hz.new_gauss_C()
    with hz.exact_lists(exact):
        fn = hz.gauss_forward(newC, sc, cam, antialiasing=aa)
        bn = hz.gauss_backward(newC, sc, cam, fn, grads)
    fr = hz.gauss_forward(refC, sc, cam, antialiasing=aa)
    br = hz.gauss_backward(refC, sc, cam, fr, grads)
    br2 = hz.gauss_backward(refC, sc, cam, fr, grads)
    for k in bn:
        assert bn[k].shape == br[k].shape, k
        noise = hz.rel_l2(br2[k], br[k])
        assert hz.rel_l2(bn[k], br[k]) <= max(GRAD_RTOL, 4 * noise), k
        assert hz.grad_close_fraction(bn[k], br[k], rtol=GRAD_RTOL) >= 0.999, k


@pytest.mark.parametrize("exact", MODES)
def test_variants(exact, refC):
    """precomputed colours, precomputed 3D covariance, lower SH degree, scale modifier."""
    with hz.exact_lists(exact):
        _variants(exact, refC)


def _variants(exact, refC):
    sc, cam, grads = _setup("tiny")
    newC = hz.new_gauss_C()
    P = sc["means3D"].shape[0]
    col = torch.rand((P, 3), device="cuda")
    cov = torch.rand((P, 6), device="cuda") * 1e-3
    cov[:, [0, 3, 5]] += 2e-3
    for kw in (dict(colors_precomp=col), dict(cov3D_precomp=cov), dict(sh_degree=1), dict(scale_modifier=0.6)):
        fn, fr = hz.gauss_forward(newC, sc, cam, **kw), hz.gauss_forward(refC, sc, cam, **kw)
        assert (fn["rendered"] == fr["rendered"] if exact else 0 < fn["rendered"] <= fr["rendered"]), kw
        assert torch.equal(fn["radii"], fr["radii"]), kw
        for k in ("color", "invdepth", "normal_map", "refl_map"):
            assert hz.max_abs(fn[k], fr[k]) <= IMG_ATOL, (k, kw)
        bn, br = hz.gauss_backward(newC, sc, cam, fn, grads), hz.gauss_backward(refC, sc, cam, fr, grads)
        br2 = hz.gauss_backward(refC, sc, cam, fr, grads)
        for k in bn:      # the contract's bar, or the reference's own run-to-run atomic noise where that is larger
            assert hz.rel_l2(bn[k], br[k]) <= max(1e-4, 4 * hz.rel_l2(br2[k], br[k])), (k, kw, hz.rel_l2(bn[k], br[k]))


def test_autograd_module_and_empty():
    import diff_gaussian_rasterization as d
    sc, cam, _ = _setup("tiny")
    rs = d.GaussianRasterizationSettings(
        image_height=cam["height"], image_width=cam["width"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
        bg=torch.zeros(3, device="cuda"), scale_modifier=1.0, viewmatrix=cam["viewmatrix"], projmatrix=cam["projmatrix"],
        sh_degree=3, campos=cam["campos"], prefiltered=False, debug=False, antialiasing=True)
    leaves = {k: sc[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales3", "rotations", "refl", "normals")}
    m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
    color, radii, invd, nmap, rmap = d.GaussianRasterizer(rs)(
        means3D=leaves["means3D"], means2D=m2d, opacities=leaves["opacities"], shs=leaves["shs"], normals=leaves["normals"],
        refl_strengths=leaves["refl"], scales=leaves["scales3"], rotations=leaves["rotations"])
    assert color.shape == (3, cam["height"], cam["width"]) and invd.shape[0] == 1 and nmap.shape[0] == 3 and rmap.shape[0] == 1
    (color.sum() + invd.sum() + nmap.sum() + rmap.sum()).backward()
    for k, v in leaves.items():
        assert v.grad is not None and torch.isfinite(v.grad).all(), k
    assert float(m2d.grad.abs().sum()) > 0
    vis = d.GaussianRasterizer(rs).markVisible(sc["means3D"])
    assert vis.dtype == torch.bool and int(vis.sum()) > 0
    empty = {k: (v[:0].contiguous() if isinstance(v, torch.Tensor) else v) for k, v in sc.items()}
    f = hz.gauss_forward(hz.new_gauss_C(), empty, cam)
    assert f["rendered"] == 0 and float(f["color"].abs().max()) == 0.0


def test_against_cpu_oracle():
    """sm_100a 3DGS path vs the CPU restatement (oracle/gauss_oracle.py, pinned to the reference's
    golden vectors by tests/test_gauss_oracle_cpu.py): integer structures equal, images to 1e-4."""
    import numpy as np
    from oracle import gauss_oracle as go
    sc_np, cam_np = scenes.build("tiny")
    sc, cam = hz.gauss_scene(hz.to_dev(sc_np)), hz.to_dev(cam_np)
    with hz.exact_lists(True):
        fn = hz.gauss_forward(hz.new_gauss_C(), sc, cam, bg=(0.1, 0.2, 0.3), antialiasing=True)
    o, _ = go.forward_backward(dict(sc_np, scales3=sc["scales3"].cpu().numpy(), normals=sc["normals"].cpu().numpy()),
                               cam_np, None, bg=(0.1, 0.2, 0.3), antialiasing=True)
    assert o["rendered"] == fn["rendered"]
    assert np.array_equal(o["radii"], fn["radii"].cpu().numpy())
    for k in ("color", "invdepth", "normal_map", "refl_map"):
        d = np.abs(o[k] - fn[k].cpu().numpy())
        assert (d > 1e-4).mean() <= 1e-3, (k, d.max())


@pytest.mark.parametrize("name", ["small", "c1", "c3"])
def test_depth_sliced_binning_same_bits(name, refC):
    """Depth slices on the 3DGS path (same bookkeeping as the surfel path's, tests/test_configs_gpu.py): the blending
    order inside a tile is untouched, so the four output planes and radii must be the SAME BITS as with one slice and
    the gradients equal up to the order of the atomic adds; the one-slice run itself is checked against the reference
    extension (images bit-identical, gradients at the reference's own noise level)."""
    sc, cam, grads = _setup(name)
    newC = hz.new_gauss_C()
    f1 = hz.gauss_forward(newC, sc, cam, depth_slices=1)
    b1 = hz.gauss_backward(newC, sc, cam, f1, grads)
    b1b = hz.gauss_backward(newC, sc, cam, f1, grads)
    fr = hz.gauss_forward(refC, sc, cam)
    br = hz.gauss_backward(refC, sc, cam, fr, grads)
    br2 = hz.gauss_backward(refC, sc, cam, fr, grads)
    for k in ("color", "invdepth", "normal_map", "refl_map"):
        assert hz.max_abs(f1[k], fr[k]) <= IMG_ATOL, k
    for kw in (dict(depth_slices=0), dict(depth_slices=4, slice_fractions=[0.002, 0.01, 0.2]),
               dict(depth_slices=8, slice_fractions=[0.01, 0.02, 0.04, 0.08, 0.16, 0.32, 0.64]),
               dict(depth_slices=2, slice_fractions=[0.5])):
        fs = hz.gauss_forward(newC, sc, cam, **kw)
        bs = hz.gauss_backward(newC, sc, cam, fs, grads)
        torch.cuda.synchronize()
        assert fs["rendered"] == f1["rendered"], kw
        for k in ("color", "invdepth", "normal_map", "refl_map", "radii"):
            assert torch.equal(fs[k], f1[k]), (name, kw, k)
        for k in bs:
            noise = hz.rel_l2(b1b[k], b1[k])
            # a different slicing changes which partial sums meet in which order in the float atomics: the yardstick is
            # the order noise (ours between two identical runs, the reference's between two of its own), not 0
            ref_noise = hz.rel_l2(br2[k], br[k])
            assert hz.rel_l2(bs[k], b1[k]) <= max(1e-5, 4 * noise, 4 * ref_noise), (name, kw, k, hz.rel_l2(bs[k], b1[k]), noise)
            assert hz.rel_l2(bs[k], br[k]) <= max(GRAD_RTOL, 4 * hz.rel_l2(br2[k], br[k])), (name, kw, k)
