"""CPU restatement (numpy) of the cubemap environment lookup -- the ORACLE for K6.

TEST INFRASTRUCTURE ONLY (see oracle/surfel_oracle.py's header for the rules).

Follows submodules/cubemapencoder/src/cubemapencoder.cu of the reference:
  face_uv()  Compute_Cubemap_UV :147-187          taps()  Compute_Seamless_Index :189-263 +
  EDGE       EdgeTable :66-106 (LEFT_TOP_AS_ORIGIN)        the in-face / nearest kernels :336-427
  forward()  :297-427                              backward()  :509-710, Compute_Cubemap_UV_Backward :265-292
Pinned against outputs of the reference extension on a B200: tests/golden/cubemap_ref.npz
(tests/golden/make_golden_cubemap.py).  The reference is built with -use_fast_math, so agreement is
toleranced (~1e-6), and a ray within an ulp of a texel boundary may pick the neighbouring texel.
"""
import numpy as np

# (new face, x source, y source) per face and edge (0: u<0, 1: u>=L, 2: v<0, 3: v>=L);
# sources: 0 -> 0, 1 -> L-1, 2 -> x, 3 -> L-1-x, 4 -> y, 5 -> L-1-y
EDGE = np.array([
    [[4, 1, 4], [5, 0, 4], [3, 1, 2], [2, 1, 2]],
    [[5, 1, 4], [4, 0, 4], [3, 0, 3], [2, 0, 3]],
    [[1, 5, 1], [0, 4, 1], [4, 2, 1], [5, 3, 1]],
    [[1, 5, 0], [0, 4, 0], [4, 2, 0], [5, 3, 0]],
    [[1, 1, 4], [0, 0, 4], [3, 2, 0], [2, 2, 0]],
    [[0, 1, 4], [1, 0, 4], [3, 3, 1], [2, 3, 1]],
], np.int64)


def face_uv(d):
    x, y, z = d[:, 0], d[:, 1], d[:, 2]
    ax, ay, az = np.abs(x), np.abs(y), np.abs(z)
    dim = np.zeros(len(d), np.int64)
    mv = ax.copy()
    s = ay > mv; dim[s] = 1; mv[s] = ay[s]
    s = az > mv; dim[s] = 2
    with np.errstate(divide="ignore", invalid="ignore"):
        u = np.select([dim == 0, dim == 1], [z / x, x / y], x / z)
        v = np.select([dim == 0, dim == 1], [y / x, z / y], y / z)
    major = np.select([dim == 0, dim == 1], [x, y], z)
    face = 2 * dim + (major < 0)
    su = np.select([face == 0, face == 1, face == 3], [-1, -1, -1], 1).astype(d.dtype)
    sv = np.select([face == 0, face == 3, face == 4], [-1, -1, -1], 1).astype(d.dtype)
    return face, u * su, v * sv


def _src(code, x, y, L):
    return np.select([code == 0, code == 1, code == 2, code == 3, code == 4], [0 * x, 0 * x + L - 1, x, L - 1 - x, y], L - 1 - y)


def _neigh(face, edge, x, y, L):
    r = EDGE[face, edge]
    return r[:, 0], _src(r[:, 1], x, y, L), _src(r[:, 2], x, y, L)


def taps(d, L, interp=1, seamless=1):
    """-> dict face[4,B], x[4,B], y[4,B], kx, ky, flag, vertex (tap 3 unused when vertex)."""
    dt = d.dtype
    face, u, v = face_uv(d)
    fu = (u * dt.type(0.5) + dt.type(0.5)) * dt.type(L)
    fv = (-v * dt.type(0.5) + dt.type(0.5)) * dt.type(L)
    B = len(d)
    F = np.tile(face, (4, 1)); X = np.zeros((4, B), np.int64); Y = np.zeros((4, B), np.int64)
    flag = np.zeros(B, np.int64); vertex = np.zeros(B, bool)
    if interp == 0:
        ux = np.clip(np.trunc(fu).astype(np.int64), 0, L - 1); uy = np.clip(np.trunc(fv).astype(np.int64), 0, L - 1)
        X[:] = ux; Y[:] = uy
        return dict(face=F, x=X, y=Y, kx=np.zeros(B, dt), ky=np.zeros(B, dt), flag=flag, vertex=vertex, fdir=face)
    x0 = np.floor(fu - dt.type(0.5)).astype(np.int64); y0 = np.floor(fv - dt.type(0.5)).astype(np.int64)
    kx = fu - x0.astype(dt) - dt.type(0.5); ky = fv - y0.astype(dt) - dt.type(0.5)
    x1 = np.clip(x0 + 1, 0, L - 1); y1 = np.clip(y0 + 1, 0, L - 1)
    x0 = np.clip(x0, 0, L - 1); y0 = np.clip(y0, 0, L - 1)
    X[0], Y[0] = x0, y0; X[1], Y[1] = x1, y0; X[2], Y[2] = x0, y1; X[3], Y[3] = x1, y1
    if seamless:
        lo_u, hi_u = fu < 0.5, fu >= dt.type(L) - dt.type(0.5)
        lo_v, hi_v = fv < 0.5, fv >= dt.type(L) - dt.type(0.5)
        flag = lo_u * 1 + (~lo_u & hi_u) * 2 + lo_v * 4 + (~lo_v & hi_v) * 8
        kx = np.where(lo_u, dt.type(0.5) - fu, kx); ky = np.where(lo_v, dt.type(0.5) - fv, ky)
        eu = np.where(lo_u, 0, 1); ev = np.where(lo_v, 2, 3)
        has_u, has_v = (flag & 3) != 0, (flag & 12) != 0
        vertex = has_u & has_v
        nu0 = _neigh(face, eu, x0, y0, L); nu1 = _neigh(face, eu, x0, y1, L)
        nv0 = _neigh(face, ev, x0, y0, L); nv1 = _neigh(face, ev, x1, y0, L)
        only_u, only_v = has_u & ~has_v, has_v & ~has_u
        for k, (cond, src) in enumerate([(None, None), (only_u | vertex, nu0), (None, None), (None, None)]):
            if cond is not None:
                F[k] = np.where(cond, src[0], F[k]); X[k] = np.where(cond, src[1], X[k]); Y[k] = np.where(cond, src[2], Y[k])
        # edge-u: taps (x0,y0), N(x0,y0), (x0,y1), N(x0,y1)
        X[2] = np.where(only_u, x0, X[2]); Y[2] = np.where(only_u, y1, Y[2])
        F[3] = np.where(only_u, nu1[0], F[3]); X[3] = np.where(only_u, nu1[1], X[3]); Y[3] = np.where(only_u, nu1[2], Y[3])
        # edge-v: (x0,y0), (x1,y0), N(x0,y0), N(x1,y0)
        F[2] = np.where(only_v | vertex, nv0[0], F[2]); X[2] = np.where(only_v | vertex, nv0[1], X[2]); Y[2] = np.where(only_v | vertex, nv0[2], Y[2])
        F[3] = np.where(only_v, nv1[0], F[3]); X[3] = np.where(only_v, nv1[1], X[3]); Y[3] = np.where(only_v, nv1[2], Y[3])
    return dict(face=F, x=X, y=Y, kx=kx.astype(dt), ky=ky.astype(dt), flag=flag, vertex=vertex, fdir=face)


def forward(inputs, cubemap, fail_value, interp=1, seamless=1):
    """inputs[B,3], cubemap[6,C,L,L], fail_value[C] -> outputs[C,B] (channel-major like the reference)."""
    d = np.asarray(inputs); tex = np.asarray(cubemap); dt = d.dtype
    C, L = tex.shape[1], tex.shape[2]
    zero = (d == 0).all(1)
    ds = np.where(zero[:, None], np.array([1, 0, 0], dt), d)
    t = taps(ds, L, interp, seamless)
    val = tex[t["face"][:, None, :], np.arange(C)[None, :, None], t["y"][:, None, :], t["x"][:, None, :]]   # [4,C,B]
    if interp == 0:
        out = val[0]
    else:
        v00, v01, v10 = val[0], val[1], val[2]
        v11 = np.where(t["vertex"], (v00 + v01 + v10) / dt.type(3), val[3])
        kx, ky = t["kx"], t["ky"]
        out = (1 - ky) * ((1 - kx) * v00 + kx * v01) + ky * ((1 - kx) * v10 + kx * v11)
    return np.where(zero[None, :], np.asarray(fail_value)[:, None], out).astype(dt)


def backward(grad_outputs, inputs, cubemap, interp=1, seamless=1):
    """-> (grad_inputs[B,3], grad_cubemap[6,C,L,L], grad_fail[C])."""
    d = np.asarray(inputs); tex = np.asarray(cubemap); g = np.asarray(grad_outputs); dt = d.dtype
    C, L = tex.shape[1], tex.shape[2]
    zero = (d == 0).all(1)
    live = ~zero
    ds = np.where(zero[:, None], np.array([1, 0, 0], dt), d)
    t = taps(ds, L, interp, seamless)
    kx, ky, vert = t["kx"], t["ky"], t["vertex"]
    gcm = np.zeros(tex.shape, np.float64)
    gfail = (g * zero[None, :]).sum(1)
    if interp == 0:
        w = [np.ones_like(kx)]
    else:
        extra = np.where(vert, ky * kx / dt.type(3), 0)
        w = [(1 - ky) * (1 - kx) + extra, (1 - ky) * kx + extra, ky * (1 - kx) + extra, np.where(vert, 0, ky * kx)]
    for k, wk in enumerate(w):
        for c in range(C):
            np.add.at(gcm, (t["face"][k], c, t["y"][k], t["x"][k]), np.where(live, wk * g[c], 0).astype(np.float64))
    gin = np.zeros_like(d)
    if interp != 0:
        val = tex[t["face"][:, None, :], np.arange(C)[None, :, None], t["y"][:, None, :], t["x"][:, None, :]]
        v00, v01, v10 = val[0], val[1], val[2]
        v11 = np.where(vert, (v00 + v01 + v10) / dt.type(3), val[3])
        gu = ((1 - ky) * (v01 - v00) + ky * (v11 - v10)) * (dt.type(0.5) * L * g)
        gv = ((1 - kx) * (v10 - v00) + kx * (v11 - v01)) * (dt.type(0.5) * L * g)
        gu = np.where(t["flag"] & 1, -gu, gu)
        gv = -np.where(t["flag"] & 4, -gv, gv)
        face = t["fdir"]
        x, y, z = ds[:, 0], ds[:, 1], ds[:, 2]
        su = np.select([face == 0, face == 1, face == 3], [-1, -1, -1], 1)
        sv = np.select([face == 0, face == 3, face == 4], [-1, -1, -1], 1)
        gu, gv = gu * su, gv * sv
        with np.errstate(divide="ignore", invalid="ignore"):
            a0 = np.stack([-(z * gu + y * gv) / (x * x), gv / x, gu / x], -1)
            a1 = np.stack([gu / y, -(x * gu + z * gv) / (y * y), gv / y], -1)
            a2 = np.stack([gu / z, gv / z, -(x * gu + y * gv) / (z * z)], -1)
        pair = (face // 2)[None, :, None]
        gd = np.where(pair == 0, a0, np.where(pair == 1, a1, a2)).sum(0)     # sum over channels -> [B,3]
        gin = np.where(live[:, None], gd, 0).astype(dt)
    return gin, gcm.astype(dt), gfail.astype(dt)
