"""CPU restatement of the 3DGS (EWA) differentiable rasterizer of the fork -- the ORACLE.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package imports this file; only tests/
may.  The reference's 3DGS extension is CUDA-only (submodules/diff-gaussian-rasterization),
ships no golden vectors or tests, so this restatement is pinned against
outputs of the reference extension itself, run on a B200 on the `tiny` scene and committed as
tests/golden/gauss_tiny_ref.npz by tests/golden/make_golden_gauss.py.

What is restated (paths relative to submodules/diff-gaussian-rasterization/):
  preprocess()  cuda_rasterizer/forward.cu:151-269 (preprocessCUDA) with computeCov3D :114-148
                (quaternion NOT normalised, :123), computeCov2D :74-109 (1.3 tan-fov clamp),
                the +0.3 dilation and optional antialiasing scale :217-226, eigenvalue radius
                :238-242, ndc2Pix auxiliary.h:40-43, getRect, in_frustum, computeColorFromSH :20-71
  binning       the same duplicateWithKeys / stable tile|depth sort / identifyTileRanges as the
                surfel path (cuda_rasterizer/rasterizer_impl.cu) -> surfel_oracle.binning
  composite()   cuda_rasterizer/forward.cu:274-411 (renderCUDA): colour, normal, reflection
                strength and inverse depth, alpha = min(0.99, opacity * exp(power))
  backward      torch autograd through the restated forward.  It reproduces the reference's
                hand-written backward (backward.cu) wherever that is the true derivative
                (colours/SH, normals, reflection strength, opacity, 3-D means, scales, rotations,
                precomputed covariances).  NOT reproduced: the gradient returned for means2D,
                which in this fork is the running-partial-sum quantity of backward.cu:614
                 and not a derivative -- the GPU tests check it
                against the reference extension directly.

Arithmetic is plain fp32/fp64 torch without the GPU's fma contraction, so agreement with the
CUDA paths is toleranced, never bit-exact.
"""
import math

import numpy as np
import torch

from oracle import surfel_oracle as so

TILE = 16


def _as(x, dtype):
    return torch.as_tensor(np.asarray(x), dtype=dtype)


class _AAScale(torch.autograd.Function):
    """h = sqrt(max(2.5e-5, det(cov) / det(cov + 0.3 I)))  (forward.cu:222-226) with the reference's
    hand-written derivative (backward.cu:209-246): the closed form there is written for the
    UN-dilated covariance entries but is evaluated with the already dilated c_xx, c_yy -- reproduced
    literally, it is what the reference's gradients contain."""

    @staticmethod
    def forward(ctx, cxx, cxy, cyy):
        det_cov = cxx * cyy - cxy * cxy
        det_h = (cxx + 0.3) * (cyy + 0.3) - cxy * cxy
        det_h = torch.where(det_h == 0, torch.ones_like(det_h), det_h)
        ratio = det_cov / det_h
        h = torch.sqrt(torch.clamp(ratio, min=0.000025))
        ctx.save_for_backward(cxx, cxy, cyy, ratio, h)
        return h

    @staticmethod
    def backward(ctx, gh):
        cxx, cxy, cyy, ratio, h = ctx.saved_tensors
        w = 0.3
        x, y, z = cxx + w, cyy + w, cxy
        d_root = torch.where(ratio <= 0.000025, torch.zeros_like(gh), gh / (2 * h))
        denom_f = d_root / (w * w + w * (x + y) + x * y - z * z) ** 2
        return (w * (w * y + y * y + z * z) * denom_f, -2.0 * w * z * (w + x + y) * denom_f,
                w * (w * x + x * x + z * z) * denom_f)


LEAF_KEYS = ("means3D", "scales3", "rotations", "opacities", "shs", "refl", "normals")


def leaves(scene, dtype=torch.float32, requires_grad=False):
    return {k: _as(scene[k], dtype).clone().requires_grad_(requires_grad) for k in LEAF_KEYS if k in scene}


def preprocess(lv, cam, sh_degree=3, scale_modifier=1.0, antialiasing=False, colors_precomp=None, cov3D_precomp=None):
    """forward.cu:151-269.  lv: dict of torch leaves (see leaves()).  Differentiable."""
    p = lv["means3D"]
    dt = p.dtype
    P = p.shape[0]
    W, H = int(cam["width"]), int(cam["height"])
    gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
    vm = _as(cam["viewmatrix"], dt)          # memory index 4*row+col, read column-major by the kernels
    pm = _as(cam["projmatrix"], dt)
    campos = _as(cam["campos"], dt)
    tanx, tany = float(cam["tanfovx"]), float(cam["tanfovy"])
    focal_x, focal_y = W / (2.0 * tanx), H / (2.0 * tany)      # rasterizer_impl.cu:216-217

    ones = torch.ones((P, 1), dtype=dt)
    ph = torch.cat([p, ones], 1)
    p_view = ph @ vm[:, :3]                                     # transformPoint4x3
    alive = p_view[:, 2].detach() > 0.2                         # in_frustum (auxiliary.h)
    p_hom = ph @ pm                                             # transformPoint4x4
    p_w = 1.0 / (p_hom[:, 3] + 1e-7)
    p_proj = p_hom[:, :3] * p_w[:, None]

    if cov3D_precomp is None:
        s = lv["scales3"] * scale_modifier
        q = lv["rotations"]                                     # raw, un-normalised (forward.cu:123)
        r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        R = torch.stack([
            1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
            2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
            2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).reshape(P, 3, 3)
        # glm::mat3(a..i) fills COLUMNS: R_glm[c][r]; as a math matrix Rm[r][c] = listed[c*3+r] -> transpose of the
        # row-listed array above.  M = S * R (glm) = Rm_math scaled by rows; Sigma = M^T M.
        Rm = R.transpose(1, 2)
        M = s[:, :, None] * Rm                                  # diag(s) @ Rm
        Sigma = M.transpose(1, 2) @ M
    else:
        c = _as(cov3D_precomp, dt) if not torch.is_tensor(cov3D_precomp) else cov3D_precomp
        Sigma = torch.stack([c[:, 0], c[:, 1], c[:, 2], c[:, 1], c[:, 3], c[:, 4], c[:, 2], c[:, 4], c[:, 5]], 1).reshape(P, 3, 3)
    cov3D = torch.stack([Sigma[:, 0, 0], Sigma[:, 0, 1], Sigma[:, 0, 2], Sigma[:, 1, 1], Sigma[:, 1, 2], Sigma[:, 2, 2]], 1)

    # computeCov2D (forward.cu:74-109)
    tz = p_view[:, 2]
    tz_safe = torch.where(alive, tz, torch.ones_like(tz))
    limx, limy = 1.3 * tanx, 1.3 * tany
    # forward.cu:82-87 clamps t.x/t.z; backward.cu:182-183 then treats a clamped t.x as a constant
    # (x_grad_mul = 0) and an unclamped one as t.x itself -- not as clamp(t.x/t.z)*t.z
    rx, ry = p_view[:, 0].detach() / tz_safe.detach(), p_view[:, 1].detach() / tz_safe.detach()
    tx = torch.where((rx < -limx) | (rx > limx), (torch.clamp(rx, -limx, limx) * tz_safe).detach(), p_view[:, 0])
    ty = torch.where((ry < -limy) | (ry > limy), (torch.clamp(ry, -limy, limy) * tz_safe).detach(), p_view[:, 1])
    zero = torch.zeros_like(tz)
    # math matrix J (rows): [[fx/tz, 0, -fx tx/tz^2], [0, fy/tz, -fy ty/tz^2], [0,0,0]]
    J = torch.stack([focal_x / tz_safe, zero, -(focal_x * tx) / (tz_safe * tz_safe),
                     zero, focal_y / tz_safe, -(focal_y * ty) / (tz_safe * tz_safe),
                     zero, zero, zero], 1).reshape(P, 3, 3)
    Wm = vm[:3, :3].t()                                         # world -> view rotation as a math matrix
    A = J @ Wm                                                  # cov = A Sigma A^T (EWA eq. 29/31)
    cov = A @ Sigma @ A.transpose(1, 2)
    cxx, cxy, cyy = cov[:, 0, 0], cov[:, 0, 1], cov[:, 1, 1]
    h = _AAScale.apply(cxx, cxy, cyy) if antialiasing else torch.ones_like(cxx)
    cxx = cxx + 0.3
    cyy = cyy + 0.3
    det = cxx * cyy - cxy * cxy
    alive = alive & (det.detach() != 0)
    det_safe = torch.where(alive, det, torch.ones_like(det))
    conic = torch.stack([cyy / det_safe, -cxy / det_safe, cxx / det_safe], 1)
    mid = 0.5 * (cxx + cyy)
    disc = torch.sqrt(torch.clamp(mid * mid - det, min=0.1))
    lam = torch.maximum(mid + disc, mid - disc)
    radius = torch.ceil(3.0 * torch.sqrt(lam.detach()))
    centre = torch.stack([((p_proj[:, 0] + 1.0) * W - 1.0) * 0.5, ((p_proj[:, 1] + 1.0) * H - 1.0) * 0.5], 1)

    cx, cy = centre[:, 0].detach(), centre[:, 1].detach()
    rad = torch.where(alive, radius, torch.zeros_like(radius))

    def clampi(v, hi):
        return torch.clamp(torch.trunc(v), 0, hi).to(torch.int64)
    x0, y0 = clampi((cx - rad) / TILE, gx), clampi((cy - rad) / TILE, gy)
    x1, y1 = clampi((cx + rad + TILE - 1) / TILE, gx), clampi((cy + rad + TILE - 1) / TILE, gy)
    tiles = (x1 - x0) * (y1 - y0)
    alive = alive & (tiles > 0)
    tiles = torch.where(alive, tiles, torch.zeros_like(tiles))
    radii = torch.where(alive, radius, torch.zeros_like(radius)).to(torch.int32)

    if colors_precomp is None:
        raw = so.sh_to_rgb(sh_degree, lv["shs"], p - campos) + 0.5
        clamped = raw.detach() < 0
        rgb = torch.clamp(raw, min=0.0)
    else:
        rgb = colors_precomp if torch.is_tensor(colors_precomp) else _as(colors_precomp, dt)
        clamped = torch.zeros((P, 3), dtype=torch.bool)
    opa = lv["opacities"].reshape(-1) * h
    return dict(alive=alive, depths=p_view[:, 2].detach(), radii=radii, tiles_touched=tiles.to(torch.int32),
                rect=torch.stack([x0, y0, x1, y1], 1), means2D=centre, conic=conic, opacity_h=opa, rgb=rgb,
                clamped=clamped, cov3D=cov3D, normals=lv["normals"],
                refl=lv["refl"].reshape(-1) if lv.get("refl") is not None else torch.zeros(P, dtype=dt),
                invdepth=1.0 / torch.where(alive, p_view[:, 2], torch.ones_like(tz)), grid=(gx, gy))


def composite(pre, point_list, ranges, cam, bg, tiles=None):
    """forward.cu:274-411 (renderCUDA), vectorised over the pixels of all selected tiles; the loop
    runs over the position j in each tile's depth-sorted list.  Differentiable."""
    W, H = int(cam["width"]), int(cam["height"])
    gx, gy = pre["grid"]
    dt = pre["conic"].dtype
    tiles = np.arange(gx * gy) if tiles is None else np.asarray(list(tiles), np.int64)
    S = tiles.size
    lens = (ranges[tiles, 1] - ranges[tiles, 0]).astype(np.int64)
    Lmax = int(lens.max()) if S else 0
    ids = np.full((S, max(Lmax, 1)), -1, np.int64)
    for k, tl in enumerate(tiles):
        ids[k, :lens[k]] = point_list[ranges[tl, 0]:ranges[tl, 1]]
    ids_t = torch.from_numpy(ids)
    ty, tx = np.divmod(tiles, gx)
    ly, lx = np.divmod(np.arange(TILE * TILE), TILE)
    px = torch.from_numpy((tx[:, None] * TILE + lx[None, :]).astype(np.int64))
    py = torch.from_numpy((ty[:, None] * TILE + ly[None, :]).astype(np.int64))
    inside = (px < W) & (py < H)
    fx, fy = px.to(dt), py.to(dt)

    z = lambda: torch.zeros((S, TILE * TILE), dtype=dt)  # noqa: E731
    T = torch.ones((S, TILE * TILE), dtype=dt)
    done = ~inside
    C, N = [z(), z(), z()], [z(), z(), z()]
    refl, invd = z(), z()
    last = torch.zeros((S, TILE * TILE), dtype=torch.int64)
    for j in range(Lmax):
        g = ids_t[:, j]
        valid = (g >= 0)[:, None] & ~done
        if not bool(valid.any()):
            if bool(done.all()):
                break
            continue
        gi = g.clamp(min=0)
        xy = pre["means2D"][gi]
        con = pre["conic"][gi]
        dx, dy = xy[:, None, 0] - fx, xy[:, None, 1] - fy
        power = -0.5 * (con[:, None, 0] * dx * dx + con[:, None, 2] * dy * dy) - con[:, None, 1] * dx * dy   # :355-358
        ok = valid & ~(power.detach() > 0)
        alpha = torch.clamp(pre["opacity_h"][gi][:, None] * torch.exp(torch.where(ok, power, torch.zeros_like(power))), max=0.99)
        ok = ok & (alpha.detach() >= 1.0 / 255.0)
        test_T = T * (1 - alpha)
        stop = ok & (test_T.detach() < 1e-4)
        ok = ok & ~stop
        w = torch.where(ok, alpha * T, torch.zeros_like(T))
        rgb, nrm = pre["rgb"][gi], pre["normals"][gi]
        for ch in range(3):
            C[ch] = C[ch] + rgb[:, None, ch] * w
            N[ch] = N[ch] + nrm[:, None, ch] * w
        refl = refl + pre["refl"][gi][:, None] * w
        invd = invd + pre["invdepth"][gi][:, None] * w
        T = torch.where(ok, test_T, T)
        last = torch.where(ok, torch.full_like(last, j + 1), last)
        done = done | stop
    bgt = _as(bg, dt)
    color = [C[ch] + T * bgt[ch] for ch in range(3)]
    return dict(px=px, py=py, inside=inside, color=color, normal=N, refl=refl, invdepth=invd, final_T=T, last=last)


def assemble(comp, W, H):
    idx = (comp["py"] * W + comp["px"])[comp["inside"]]

    def plane(v):
        out = torch.zeros(W * H, dtype=v.dtype)
        return out.index_put((idx,), v[comp["inside"]]).reshape(H, W)
    return (torch.stack([plane(c) for c in comp["color"]]), plane(comp["invdepth"])[None],
            torch.stack([plane(c) for c in comp["normal"]]), plane(comp["refl"])[None],
            plane(comp["final_T"]), plane(comp["last"]))


def forward_backward(scene, cam, grads=None, bg=(0.0, 0.0, 0.0), sh_degree=3, scale_modifier=1.0, antialiasing=False,
                     dtype=torch.float32):
    """scene: numpy dict with means3D, scales3, rotations, opacities, shs, refl, normals.
    grads (optional): numpy dict color/invdepth/normal/refl of upstream gradients.
    -> (outputs dict of numpy arrays, gradients dict of numpy arrays or None)."""
    lv = leaves(scene, dtype, requires_grad=grads is not None)
    W, H = int(cam["width"]), int(cam["height"])
    pre = preprocess(lv, cam, sh_degree, scale_modifier, antialiasing)
    keys, plist, ranges = so.binning(pre, W, H)
    comp = composite(pre, plist, ranges, cam, bg)
    color, invdepth, normal, refl, final_T, last = assemble(comp, W, H)
    d = lambda t: t.detach().numpy()  # noqa: E731
    out = dict(color=d(color), invdepth=d(invdepth), normal_map=d(normal), refl_map=d(refl), radii=d(pre["radii"]),
               tiles_touched=d(pre["tiles_touched"]), keys=keys, point_list=plist, ranges=ranges,
               n_contrib=d(last), rendered=int(plist.size))
    gr = None
    if grads is not None:
        loss = ((color * _as(grads["color"], dtype)).sum() + (invdepth * _as(grads["invdepth"], dtype)).sum()
                + (normal * _as(grads["normal"], dtype)).sum() + (refl * _as(grads["refl"], dtype)).sum())
        loss.backward()
        z = lambda t: torch.zeros_like(t) if t.grad is None else t.grad  # noqa: E731
        gr = dict(means3D=d(z(lv["means3D"])), sh=d(z(lv["shs"])), opacity=d(z(lv["opacities"])),
                  scales=d(z(lv["scales3"])), rotations=d(z(lv["rotations"])), refl=d(z(lv["refl"])),
                  normals=d(z(lv["normals"])))
    return out, gr
