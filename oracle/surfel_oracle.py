"""CPU restatement of the surfel (2DGS) differentiable rasterizer -- the ORACLE.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package imports this file;
only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference
legs may.  It is also the "naive per-pixel PyTorch-CPU rasterizer" baseline that
BASELINE.json's north_star asks for (the reference has no CPU path).

Parity status: the reference ships NO golden vectors or tests for this path
(SURVEY.md section 4), so this restatement is pinned against outputs of the
reference CUDA extension itself, run on a B200 on the `tiny` scene and committed
as tests/golden/surfel_tiny_ref.npz by tests/golden/make_golden.py.

What is restated (paths relative to submodules/diff-surfel-rasterization/):
  preprocess()  cuda_rasterizer/forward.cu:148-253 (preprocessCUDA) with
                compute_transmat :75-115, compute_aabb :119-145,
                computeColorFromSH :20-71, auxiliary.h:71-81 (getRect),
                :189-214 (in_frustum), :217-239 (quat_to_rotmat)
  binning()     cuda_rasterizer/rasterizer_impl.cu:70-111 (duplicateWithKeys),
                :308-313 (stable sort of tile|depth keys), :116-138 (ranges)
  composite()   cuda_rasterizer/forward.cu:258-489 (renderCUDA)
  backward      torch autograd through the same restated forward, with the
                reference's non-derivative choices reproduced explicitly
                (backward.cu:454-464 frozen intersection in the low-pass branch;
                backward.cu:657-659 densification proxy written to dL_dmeans2D;
                auxiliary.h:242-286 gradient w.r.t. the normalised quaternion).

Arithmetic is plain fp32 (or fp64 with dtype=torch.float64) WITHOUT the GPU's
fma contraction / MUFU approximations, so agreement with the CUDA paths is
toleranced, never bit-exact; integer structures agree except where a float lands
within an ulp of a decision boundary.
"""
import math

import numpy as np
import torch

NEAR_N, FAR_N = 0.2, 100.0          # auxiliary.h:41-42
FILTER_SIZE = 0.707106              # auxiliary.h:43
FILTER_INV_SQUARE = 2.0             # auxiliary.h:44
TILE = 16                           # config.h:15-17

SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
SH_C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
         -0.4570457994644658, 1.445305721320277, -0.5900435899266435)


def _as(x, dtype):
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().to(dtype)
    return torch.as_tensor(np.asarray(x), dtype=dtype)


def sh_to_rgb(deg, sh, dirs):
    """forward.cu:20-71.  sh [P,M,3], dirs [P,3] (un-normalised) -> rgb before +0.5/clamp."""
    d = dirs / dirs.norm(dim=1, keepdim=True)
    x, y, z = d[:, 0:1], d[:, 1:2], d[:, 2:3]
    r = SH_C0 * sh[:, 0]
    if deg > 0:
        r = r - SH_C1 * y * sh[:, 1] + SH_C1 * z * sh[:, 2] - SH_C1 * x * sh[:, 3]
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            r = (r + SH_C2[0] * xy * sh[:, 4] + SH_C2[1] * yz * sh[:, 5] + SH_C2[2] * (2.0 * zz - xx - yy) * sh[:, 6]
                 + SH_C2[3] * xz * sh[:, 7] + SH_C2[4] * (xx - yy) * sh[:, 8])
            if deg > 2:
                r = (r + SH_C3[0] * y * (3.0 * xx - yy) * sh[:, 9] + SH_C3[1] * xy * z * sh[:, 10]
                     + SH_C3[2] * y * (4.0 * zz - xx - yy) * sh[:, 11]
                     + SH_C3[3] * z * (2.0 * zz - 3.0 * xx - 3.0 * yy) * sh[:, 12]
                     + SH_C3[4] * x * (4.0 * zz - xx - yy) * sh[:, 13] + SH_C3[5] * z * (xx - yy) * sh[:, 14]
                     + SH_C3[6] * x * (xx - 3.0 * yy) * sh[:, 15])
    return r


def preprocess(leaves, cam, sh_degree=3, scale_modifier=1.0, colors_precomp=None):
    """forward.cu:148-253.  `leaves`: dict of torch tensors means3D[P,3], scales[P,2],
    rotations[P,4] (w,x,y,z), opacities[P,1], shs[P,M,3], refl[P,1].  Differentiable.
    Returns a dict of per-Gaussian state; integer fields are detached."""
    p = leaves["means3D"]
    dt = p.dtype
    W, H = int(cam["width"]), int(cam["height"])
    vm = _as(cam["viewmatrix"], dt)      # row-vector convention: p_view = [p 1] @ vm
    pm = _as(cam["projmatrix"], dt)
    campos = _as(cam["campos"], dt)
    P = p.shape[0]

    p_view = p @ vm[:3, :3] + vm[3, :3]                                  # auxiliary.h:83-91
    alive = p_view[:, 2].detach() > NEAR_N                               # in_frustum, auxiliary.h:204

    # quat_to_rotmat (auxiliary.h:217-239): gradient is taken w.r.t. the normalised
    # quaternion (quat_to_rotmat_vjp does not back-propagate through the rsqrt).
    q = leaves["rotations"]
    qn = q * torch.rsqrt((q * q).sum(1, keepdim=True)).detach()
    w, x, y, z = qn[:, 0], qn[:, 1], qn[:, 2], qn[:, 3]
    c0 = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y + w * z), 2 * (x * z - w * y)], 1)   # R[0] (column 0)
    c1 = torch.stack([2 * (x * y - w * z), 1 - 2 * (x * x + z * z), 2 * (y * z + w * x)], 1)   # R[1]
    c2 = torch.stack([2 * (x * z + w * y), 2 * (y * z - w * x), 1 - 2 * (x * x + y * y)], 1)   # R[2]
    s = leaves["scales"]
    L0 = (scale_modifier * s[:, 0:1]) * c0                                # forward.cu:88-90
    L1 = (scale_modifier * s[:, 1:2]) * c1
    # rows of transpose(splat2world) times world2ndc (forward.cu:93-112)
    Mu = L0 @ pm[:3, :]                                                   # [P,4]
    Mv = L1 @ pm[:3, :]
    Mc = p @ pm[:3, :] + pm[3, :]
    rows = (Mu, Mv, Mc)

    def build_T(rows, Wv, Hv):
        hw, hw1, hh, hh1 = Wv / 2.0, (Wv - 1) / 2.0, Hv / 2.0, (Hv - 1) / 2.0
        Tu = torch.stack([hw * m[:, 0] + hw1 * m[:, 3] for m in rows], 1)     # transMat[0..2]
        Tv = torch.stack([hh * m[:, 1] + hh1 * m[:, 3] for m in rows], 1)     # transMat[3..5]
        Tw = torch.stack([m[:, 3] for m in rows], 1)                          # transMat[6..8]
        return torch.cat([Tu, Tv, Tw], 1)
    transMat = build_T(rows, W, H)
    # backward.cu:637-638 (quirk S8): the backward rebuilds T with W' = int(focal_x*tan_fovx*2),
    # H' = int(focal_y*tan_fovy*2) evaluated in fp32 -- one less than W/H when the product rounds
    # down -- and with scale_modifier ignored (backward.cu:511, quirk S9).  Values below are the
    # forward's; gradients flow through the backward's T'.
    f32 = np.float32
    tx, ty = f32(cam["tanfovx"]), f32(cam["tanfovy"])
    Wb = int(f32(f32(f32(W) / f32(f32(2.0) * tx)) * tx) * f32(2.0))
    Hb = int(f32(f32(f32(H) / f32(f32(2.0) * ty)) * ty) * f32(2.0))
    if (Wb, Hb) != (W, H) or scale_modifier != 1.0:
        T_bwd = build_T(((s[:, 0:1] * c0) @ pm[:3, :], (s[:, 1:2] * c1) @ pm[:3, :], Mc), Wb, Hb)
    else:
        T_bwd = transMat
    bwd_size = (Wb, Hb)

    normal = c2 @ vm[:3, :3]                                              # transformVec4x3, forward.cu:113
    cos = -(p_view * normal).sum(1)                                       # forward.cu:211-216
    alive &= cos.detach() != 0
    normal = normal * torch.where(cos.detach() > 0, 1.0, -1.0).to(dt)[:, None]

    # Two aliases of T so that autograd separates the compositing gradient (what the
    # reference returns as dL_dtransMat) from the AABB-centre path (backward.cu:545-560).
    T_render = T_bwd * 1.0
    T_aabb = T_bwd * 1.0
    T_render_val = T_render + (transMat - T_bwd).detach()      # forward value, backward's Jacobian
    aTu, aTv, aTw = T_aabb[:, 0:3], T_aabb[:, 3:6], T_aabb[:, 6:9]
    t = torch.tensor([9.0, 9.0, -1.0], dtype=dt)                          # cutoff 3, forward.cu:119-145
    d = (t * aTw * aTw).sum(1)
    alive &= d.detach() != 0
    dsafe = torch.where(d == 0, torch.ones_like(d), d)
    f = t[None, :] / dsafe[:, None]
    centre = torch.stack([(f * aTu * aTw).sum(1), (f * aTv * aTw).sum(1)], 1)
    if T_bwd is not transMat:                                   # forward value of the centre uses the forward T
        with torch.no_grad():
            fTu, fTv, fTw = transMat[:, 0:3], transMat[:, 3:6], transMat[:, 6:9]
            fd = (t * fTw * fTw).sum(1)
            alive &= fd != 0
            ff = t[None, :] / torch.where(fd == 0, torch.ones_like(fd), fd)[:, None]
            cf = torch.stack([(ff * fTu * fTw).sum(1), (ff * fTv * fTw).sum(1)], 1)
            aTu, aTv, aTw, f = fTu, fTv, fTw, ff
        centre = centre + (cf - centre).detach()
    with torch.no_grad():
        h0 = centre * centre - torch.stack([(f * aTu * aTu).sum(1), (f * aTv * aTv).sum(1)], 1)
        ext = torch.sqrt(torch.clamp(h0, min=1e-4))
        radius = torch.ceil(torch.clamp(ext.max(1).values, min=3.0 * FILTER_SIZE))     # forward.cu:232
        radius = torch.where(alive, radius, torch.zeros_like(radius))
        gx, gy = (W + TILE - 1) // TILE, (H + TILE - 1) // TILE
        c = centre.to(torch.float32) if dt == torch.float32 else centre
        r = radius.to(c.dtype)
        # getRect (auxiliary.h:71-81): int() truncates toward zero
        x0 = torch.clamp(torch.trunc((c[:, 0] - r) / TILE), 0, gx).long()
        y0 = torch.clamp(torch.trunc((c[:, 1] - r) / TILE), 0, gy).long()
        x1 = torch.clamp(torch.trunc((c[:, 0] + r + TILE - 1) / TILE), 0, gx).long()
        y1 = torch.clamp(torch.trunc((c[:, 1] + r + TILE - 1) / TILE), 0, gy).long()
        tiles = (x1 - x0) * (y1 - y0)
        tiles = torch.where(alive, tiles, torch.zeros_like(tiles))
        alive &= tiles > 0
        radii = torch.where(alive, radius, torch.zeros_like(radius)).to(torch.int32)

    if colors_precomp is None:
        raw = sh_to_rgb(sh_degree, leaves["shs"], p - campos) + 0.5       # forward.cu:60-70
        clamped = raw.detach() < 0
        rgb = torch.clamp(raw, min=0.0)
    else:
        rgb = colors_precomp
        clamped = torch.zeros((P, 3), dtype=torch.bool)
    return dict(alive=alive, depths=p_view[:, 2].detach(), radii=radii, tiles_touched=tiles.to(torch.int32),
                rect=torch.stack([x0, y0, x1, y1], 1), means2D=centre, transMat=transMat, T_render=T_render,
                T_render_val=T_render_val, bwd_size=bwd_size,
                T_aabb=T_aabb, normal=normal, opacity=leaves["opacities"].reshape(-1), rgb=rgb, clamped=clamped,
                refl=leaves["refl"].reshape(-1) if leaves.get("refl") is not None else torch.zeros(P, dtype=dt),
                grid=(gx, gy))


def binning(pre, W, H, tiles=None):
    """duplicateWithKeys + stable sort + identifyTileRanges (rasterizer_impl.cu:70-138, :308-313).
    -> keys int64[R] (tile<<32 | fp32 depth bits), point_list int64[R], ranges int64[T,2].
    With `tiles` (bounded CPU-baseline samples) only Gaussians overlapping one of those tiles are
    emitted; the lists and ranges of exactly those tiles are then complete, the others are not."""
    gx, gy = pre["grid"]
    alive = pre["alive"].numpy()
    rect = pre["rect"].numpy()
    dbits = pre["depths"].to(torch.float32).numpy().view(np.uint32).astype(np.int64)
    if tiles is not None:
        hit = np.zeros_like(alive)
        for tl in tiles:
            ty, tx = divmod(int(tl), gx)
            hit |= (rect[:, 0] <= tx) & (tx < rect[:, 2]) & (rect[:, 1] <= ty) & (ty < rect[:, 3])
        alive = alive & hit
    ids = np.nonzero(alive)[0]          # emission order: ascending id, rows outer, columns inner
    x0, y0, x1 = rect[ids, 0], rect[ids, 1], rect[ids, 2]
    wdt = x1 - x0
    cnt = wdt * (rect[ids, 3] - y0)
    vals = np.repeat(ids, cnt).astype(np.int64)
    local = np.arange(vals.size, dtype=np.int64) - np.repeat(np.cumsum(cnt) - cnt, cnt)
    wrep = np.repeat(wdt, cnt)
    tile = (np.repeat(y0, cnt) + local // np.maximum(wrep, 1)) * gx + np.repeat(x0, cnt) + local % np.maximum(wrep, 1)
    keys = (tile.astype(np.int64) << 32) | dbits[vals]
    order = np.argsort(keys, kind="stable")
    keys, vals = keys[order], vals[order]
    ranges = np.zeros((gx * gy, 2), np.int64)
    if keys.size:
        tiles = keys >> 32
        starts = np.nonzero(np.diff(tiles, prepend=-1))[0]
        ends = np.append(starts[1:], keys.size)
        ranges[tiles[starts], 0] = starts
        ranges[tiles[starts], 1] = ends
    return keys, vals, ranges


def composite(pre, point_list, ranges, cam, bg, mask=None, tiles=None, collect_weights=True):
    """forward.cu:258-489 (renderCUDA), vectorised over the pixels of all selected tiles;
    the loop runs over the position j in each tile's depth-sorted list.  Differentiable.
    `tiles`: iterable of tile ids to render (default all).  Returns per-pixel planes for the
    selected tiles as [S,256] tensors plus pixel coordinates, see assemble()."""
    W, H = int(cam["width"]), int(cam["height"])
    gx, gy = pre["grid"]
    dt = pre["transMat"].dtype
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
    px = torch.from_numpy((tx[:, None] * TILE + lx[None, :]).astype(np.int64))      # [S,256]
    py = torch.from_numpy((ty[:, None] * TILE + ly[None, :]).astype(np.int64))
    inside = (px < W) & (py < H)
    fx, fy = px.to(dt), py.to(dt)

    T_r = pre["T_render_val"]
    z = lambda: torch.zeros((S, TILE * TILE), dtype=dt)  # noqa: E731
    T = torch.ones((S, TILE * TILE), dtype=dt)
    done = ~inside
    C = [z(), z(), z()]
    N = [z(), z(), z()]
    refl, D, M1, M2, dist, median = z(), z(), z(), z(), z(), z()
    maskp = torch.zeros((S, TILE * TILE), dtype=torch.bool)
    last = torch.zeros((S, TILE * TILE), dtype=torch.int64)
    median_c = torch.zeros((S, TILE * TILE), dtype=torch.int64)   # float -1 -> u32 0 when none
    P = pre["transMat"].shape[0]
    weights = torch.zeros(P, dtype=dt)
    # forward.cu:444 indexes the mask FLAT by Gaussian id, whatever its shape ([P] or [P,3])
    mask_g = torch.ones(P, dtype=torch.bool) if mask is None else torch.as_tensor(np.asarray(mask)).reshape(-1)[:P].bool()
    far_scale = FAR_N / (FAR_N - NEAR_N)

    for j in range(Lmax):
        g = ids_t[:, j]
        valid = (g >= 0)[:, None] & ~done                    # loop condition `!done && j < ...`
        if not bool(valid.any()):
            if bool(done.all()):
                break
            continue
        gi = g.clamp(min=0)
        Tm = T_r[gi]                                          # [S,9]
        Tu, Tv, Tw = Tm[:, None, 0:3], Tm[:, None, 3:6], Tm[:, None, 6:9]
        k = fx[..., None] * Tw - Tu                           # forward.cu:367-368
        l = fy[..., None] * Tw - Tv
        pz = k[..., 0] * l[..., 1] - k[..., 1] * l[..., 0]    # cross(k,l)
        pxc = k[..., 1] * l[..., 2] - k[..., 2] * l[..., 1]
        pyc = k[..., 2] * l[..., 0] - k[..., 0] * l[..., 2]
        unstable = pz.detach().abs() < 1e-4                   # forward.cu:373
        pz_safe = torch.where(unstable, torch.ones_like(pz), pz)
        sx = torch.where(unstable, torch.zeros_like(pz), pxc / pz_safe)
        sy = torch.where(unstable, torch.zeros_like(pz), pyc / pz_safe)
        rho3d = torch.where(unstable, torch.full_like(pz, 1e8), sx * sx + sy * sy)
        xy = pre["means2D"][gi]
        dx, dy = xy[:, None, 0] - fx, xy[:, None, 1] - fy
        rho2d = FILTER_INV_SQUARE * (dx * dx + dy * dy)
        use3d = rho3d.detach() <= rho2d.detach()
        rho = torch.where(use3d, rho3d, rho2d)
        # backward.cu:454-464: in the low-pass branch the reference differentiates depth w.r.t. Tw only
        sxd = torch.where(use3d, sx, sx.detach())
        syd = torch.where(use3d, sy, sy.detach())
        depth = (sxd * Tw[..., 0] + syd * Tw[..., 1]) + Tw[..., 2]
        ok = valid & (depth.detach() >= NEAR_N)
        power = -0.5 * rho
        ok = ok & ~(power.detach() > 0)
        opa = pre["opacity"][gi][:, None]
        alpha = torch.clamp(opa * torch.exp(torch.where(ok, power, torch.zeros_like(power))), max=0.99)
        ok = ok & (alpha.detach() >= 1.0 / 255.0)
        test_T = T * (1 - alpha)
        stop = ok & (test_T.detach() < 1e-4)
        ok = ok & ~stop
        w = torch.where(ok, alpha * T, torch.zeros_like(T))
        depth_s = torch.where(ok, depth, torch.ones_like(depth))
        A = 1 - T
        m = far_scale * (1 - NEAR_N / depth_s)
        dist = dist + (m * m * A + M2 - 2 * m * M1) * w
        D = D + depth_s * w
        M1 = M1 + m * w
        M2 = M2 + m * m * w
        med = ok & (T.detach() > 0.5)
        median = torch.where(med, depth, median)
        median_c = torch.where(med, torch.full_like(median_c, j + 1), median_c)
        nrm = pre["normal"][gi]
        rgb = pre["rgb"][gi]
        for ch in range(3):
            N[ch] = N[ch] + nrm[:, None, ch] * w
            C[ch] = C[ch] + rgb[:, None, ch] * w
        refl = refl + pre["refl"][gi][:, None] * w
        maskp = maskp | (ok & mask_g[gi][:, None])
        T = torch.where(ok, test_T, T)
        last = torch.where(ok, torch.full_like(last, j + 1), last)
        done = done | stop
        if collect_weights:
            with torch.no_grad():                              # exact max (the reference's is racy)
                wm = w.max(1).values
                sel = g >= 0
                weights.scatter_reduce_(0, gi[sel], wm[sel], "amax", include_self=True)
    bgt = _as(bg, dt)
    color = [C[ch] + T * bgt[ch] for ch in range(3)]
    others = [D, 1 - T, N[0], N[1], N[2], median, dist, maskp.to(dt)]
    return dict(px=px, py=py, inside=inside, color=color, others=others, refl=refl, final_T=T, M1=M1, M2=M2,
                last=last, median_c=median_c, weights=weights)


def assemble(comp, W, H, dtype):
    """Scatter the [S,256] tile planes of composite() into full [C,H,W] images (zeros elsewhere)."""
    idx = (comp["py"] * W + comp["px"])[comp["inside"]]

    def plane(v):
        out = torch.zeros(W * H, dtype=v.dtype)
        return out.index_put((idx,), v[comp["inside"]]).reshape(H, W)
    color = torch.stack([plane(c) for c in comp["color"]])
    others = torch.stack([plane(c) for c in comp["others"]])
    refl = plane(comp["refl"])[None]
    accum = torch.stack([plane(comp["final_T"]), plane(comp["M1"]), plane(comp["M2"])])
    n_contrib = torch.stack([plane(comp["last"]), plane(comp["median_c"])])
    return color, others, refl, accum, n_contrib


LEAF_KEYS = ("means3D", "scales", "rotations", "opacities", "shs", "refl")


def _leaves(scene, dtype, requires_grad):
    return {k: _as(scene[k], dtype).clone().requires_grad_(requires_grad) for k in LEAF_KEYS if k in scene}


def forward(scene, cam, bg=(0.0, 0.0, 0.0), sh_degree=3, scale_modifier=1.0, mask=None, dtype=torch.float32,
            tiles=None, requires_grad=False, _keep=None):
    """Full forward on the CPU.  scene/cam: dicts of numpy arrays (tests/scenes.py).
    -> dict of numpy arrays: color[3,H,W], others[8,H,W], refl_map[1,H,W], radii[P] i32,
    tiles_touched[P], keys[R], point_list[R], ranges[T,2], accum[3,H,W], n_contrib[2,H,W],
    weights[P], plus depths/means2D/transMat/normal_opacity/rgb/clamped per Gaussian."""
    W, H = int(cam["width"]), int(cam["height"])
    leaves = _leaves(scene, dtype, requires_grad)
    with torch.set_grad_enabled(requires_grad):
        pre = preprocess(leaves, cam, sh_degree, scale_modifier)
        keys, plist, ranges = binning(pre, W, H)
        comp = composite(pre, plist, ranges, cam, bg, mask=mask, tiles=tiles)
        color, others, refl, accum, ncon = assemble(comp, W, H, dtype)
    if _keep is not None:
        _keep.update(leaves=leaves, pre=pre, color=color, others=others, refl=refl)
    d = lambda t: t.detach().numpy()  # noqa: E731
    no = torch.cat([pre["normal"], pre["opacity"][:, None]], 1)
    return dict(color=d(color), others=d(others), refl_map=d(refl), radii=d(pre["radii"]),
                tiles_touched=d(pre["tiles_touched"]), rendered=int(keys.size), keys=keys, point_list=plist,
                ranges=ranges, accum=d(accum), n_contrib=d(ncon), weights=d(comp["weights"]), depths=d(pre["depths"]),
                means2D=d(pre["means2D"]), transMat=d(pre["transMat"]), normal_opacity=d(no), rgb=d(pre["rgb"]),
                clamped=d(pre["clamped"]), alive=d(pre["alive"]))


def forward_backward(scene, cam, grads, bg=(0.0, 0.0, 0.0), sh_degree=3, scale_modifier=1.0, mask=None,
                     dtype=torch.float32, tiles=None):
    """Forward + backward.  `grads`: dict color[3,H,W], others[8,H,W], refl[1,H,W] upstream
    gradients.  -> (forward dict, gradient dict) with the reference's output names
    (S/rasterize_points.cu:153-267): means2D (densification proxy, backward.cu:657-659),
    colors, refl, opacity, means3D, transMat, sh, scales, rotations."""
    W, H = int(cam["width"]), int(cam["height"])
    keep = {}
    out = forward(scene, cam, bg, sh_degree, scale_modifier, mask, dtype, tiles, requires_grad=True, _keep=keep)
    pre, leaves = keep["pre"], keep["leaves"]
    for k in ("T_render", "rgb"):
        pre[k].retain_grad()
    go = _as(grads["others"], dtype).clone()
    go[7] = 0                                            # the mask plane has no gradient path
    loss = ((keep["color"] * _as(grads["color"], dtype)).sum() + (keep["others"] * go).sum()
            + (keep["refl"] * _as(grads["refl"], dtype)).sum())
    loss.backward()
    P = leaves["means3D"].shape[0]
    g = lambda t: (t.grad if t.grad is not None else torch.zeros_like(t)).detach()  # noqa: E731
    vis = pre["radii"] > 0
    dT = torch.where(vis[:, None], g(pre["T_render"]), torch.zeros((P, 9), dtype=dtype))
    # rgb.grad is d/d(clamped colour) == dL_dcolors (the clamp mask is applied after it, backward.cu:36-39)
    dcol = torch.where(vis[:, None], g(pre["rgb"]), torch.zeros((P, 3), dtype=dtype))
    Tm = pre["transMat"].detach()
    m2d = torch.zeros((P, 3), dtype=dtype)
    Wb, Hb = pre["bwd_size"]
    m2d[:, 0] = dT[:, 2] * Tm[:, 8] * 0.5 * Wb           # backward.cu:657-659 (W', H' of :637-638)
    m2d[:, 1] = dT[:, 5] * Tm[:, 8] * 0.5 * Hb
    n = lambda t: t.numpy()  # noqa: E731
    gr = dict(means2D=n(m2d), colors=n(dcol), refl=n(g(leaves["refl"])), opacity=n(g(leaves["opacities"])),
              means3D=n(g(leaves["means3D"])), transMat=n(dT), sh=n(g(leaves["shs"])), scales=n(g(leaves["scales"])),
              rotations=n(g(leaves["rotations"])))
    return out, gr
