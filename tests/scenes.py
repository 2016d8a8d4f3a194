"""Deterministic synthetic scenes and cameras for parity tests, smoke() and bench.py.

Scene recipes follow SURVEY.md section 8(d).  Camera matrices follow the
reference's own construction (utils/graphics_utils.py:38-71 getWorld2View2 /
getProjectionMatrix, scene/cameras.py:56-69): viewmatrix = W2C^T, projmatrix =
viewmatrix @ P^T, campos = inverse(viewmatrix)[3, :3]  (Inria "row vector times
matrix" convention).  Everything is generated on the CPU as float32 numpy arrays
so every implementation under test sees identical bits.
"""
import math

import numpy as np


def look_at(eye, target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0)):
    """World-to-camera 4x4 (COLMAP axes: x right, y down, z forward)."""
    eye = np.asarray(eye, np.float64)
    f = np.asarray(target, np.float64) - eye
    f /= np.linalg.norm(f)
    r = np.cross(-np.asarray(up, np.float64), f)
    r /= np.linalg.norm(r)
    d = np.cross(f, r)
    R = np.stack([r, d, f])  # rows = camera axes in world coordinates
    Rt = np.eye(4)
    Rt[:3, :3] = R
    Rt[:3, 3] = -R @ eye
    return np.float32(Rt)


def projection(znear, zfar, fovx, fovy):
    """utils/graphics_utils.py:51-71 (getProjectionMatrix), float32 like torch.zeros(4,4)."""
    ty, tx = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = ty * znear, tx * znear
    P = np.zeros((4, 4), np.float32)
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_camera(eye, W, H, fovx=None, focal=None, target=(0.0, 0.0, 0.0), up=(0.0, 0.0, 1.0)):
    if fovx is None:
        fovx = 2 * math.atan(W / (2 * focal))
    tanx = math.tan(fovx / 2)
    tany = tanx * H / W
    fovy = 2 * math.atan(tany)
    w2c = look_at(eye, target, up)
    view = np.ascontiguousarray(w2c.T)                       # world_view_transform
    proj = np.ascontiguousarray(projection(0.01, 100.0, fovx, fovy).T)
    full = (view.astype(np.float32) @ proj.astype(np.float32)).astype(np.float32)  # full_proj_transform
    campos = np.linalg.inv(view.astype(np.float64))[3, :3].astype(np.float32)
    return dict(width=W, height=H, tanfovx=float(tanx), tanfovy=float(tany), fovx=fovx, fovy=fovy,
                viewmatrix=view, projmatrix=np.ascontiguousarray(full), campos=campos, w2c=w2c)


def _quats(rng, P):
    q = rng.uniform(0.0, 1.0, (P, 4)).astype(np.float32)     # gaussian_model.py:172 init, :129 normalise
    return (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(np.float32)


def _appearance(rng, P, sh_coeffs=16):
    sh = np.empty((P, sh_coeffs, 3), np.float32)
    sh[:, 0] = rng.normal(0.0, 0.5, (P, 3))
    sh[:, 1:] = rng.normal(0.0, 0.1, (P, sh_coeffs - 1, 3))
    opacity = rng.uniform(0.05, 0.95, (P, 1)).astype(np.float32)
    refl = rng.uniform(0.0, 1.0, (P, 1)).astype(np.float32)
    return sh, opacity, refl


def box_scene(P, seed=0, scale_dims=2):
    """c1/c2 recipe: xyz ~ U[-1.3,1.3]^3, scales = exp(N(log s0, 0.3^2)), s0 = 0.75 (17.576/P)^(1/3)."""
    rng = np.random.default_rng(20261104 + seed)
    xyz = rng.uniform(-1.3, 1.3, (P, 3)).astype(np.float32)
    s0 = 0.75 * (17.576 / P) ** (1.0 / 3.0)
    scales = np.exp(rng.normal(math.log(s0), 0.3, (P, scale_dims))).astype(np.float32)
    sh, opacity, refl = _appearance(rng, P)
    return dict(means3D=xyz, scales=scales, rotations=_quats(rng, P), opacities=opacity, shs=sh, refl=refl)


def room_scene(P, seed=0, scale_dims=2):
    """c3/c4/c5 recipe: 70% N(0,1.5^2) clipped to r<4, 30% on shells r in [4,40] with s0 ~ r."""
    rng = np.random.default_rng(20261104 + 1000 + seed)
    n_in = int(0.7 * P)
    inner = rng.normal(0.0, 1.5, (n_in, 3))
    r = np.linalg.norm(inner, axis=1, keepdims=True)
    inner = np.where(r > 4.0, inner * (4.0 * rng.uniform(0.2, 1.0, r.shape) / np.maximum(r, 1e-9)), inner)
    n_out = P - n_in
    d = rng.normal(0.0, 1.0, (n_out, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rad = rng.uniform(4.0, 40.0, (n_out, 1))
    outer = d * rad
    xyz = np.concatenate([inner, outer]).astype(np.float32)
    s_in = 0.75 * (4.0 / 3.0 * math.pi * 4.0 ** 3 / max(n_in, 1)) ** (1.0 / 3.0)
    base = np.concatenate([np.full((n_in, 1), s_in), s_in * rad / 4.0])
    scales = (base * np.exp(rng.normal(0.0, 0.3, (P, scale_dims)))).astype(np.float32)
    sh, opacity, refl = _appearance(rng, P)
    perm = rng.permutation(P)  # no spatial order in the id space
    return dict(means3D=xyz[perm], scales=scales[perm], rotations=_quats(rng, P), opacities=opacity, shs=sh, refl=refl)


def ring_camera(k, n, W, H, focal=None, fovx=None, radius=4.0, height=1.0, jitter_seed=None):
    ang = 2 * math.pi * k / max(n, 1)
    if jitter_seed is not None:
        rng = np.random.default_rng(jitter_seed + k)
        ang += rng.uniform(-0.05, 0.05)
    eye = (radius * math.cos(ang), radius * math.sin(ang), height)
    return make_camera(eye, W, H, fovx=fovx, focal=focal)


CONFIGS = {
    # name: (scene fn, P, W, H, camera kwargs)
    "tiny": dict(scene="box", P=2000, W=64, H=48, cam=dict(eye=(4.0, 0.3, 0.5), fovx=0.6911)),
    "small": dict(scene="box", P=20000, W=160, H=112, cam=dict(eye=(4.0, 0.3, 0.5), fovx=0.6911)),
    "c1": dict(scene="box", P=100_000, W=256, H=256, cam=dict(eye=(4.0, 0.0, 0.0), fovx=0.6911)),
    "c2": dict(scene="box", P=300_000, W=800, H=800, cam=dict(eye=(4.0, 0.0, 0.0), fovx=0.6911)),
    "c3": dict(scene="room", P=3_000_000, W=1237, H=822, ring=dict(focal=1160.0)),
    "c4": dict(scene="room", P=1_000_000, W=1920, H=1080, ring=dict(focal=1160.0)),
    "c5": dict(scene="room", P=6_000_000, W=1920, H=1080, ring=dict(focal=1160.0)),
    "bench": dict(scene="room", P=3_000_000, W=1920, H=1080, ring=dict(focal=1160.0)),
}


def build(name, P=None, seed=0, view=0, n_views=32):
    """-> (scene dict of numpy arrays, camera dict)."""
    cfg = CONFIGS[name]
    P = P or cfg["P"]
    scene = (box_scene if cfg["scene"] == "box" else room_scene)(P, seed)
    if "cam" in cfg:
        c = dict(cfg["cam"])
        cam = make_camera(c.pop("eye"), cfg["W"], cfg["H"], **c)
    else:
        cam = ring_camera(view, n_views, cfg["W"], cfg["H"], **cfg["ring"])
    return scene, cam


def camera(name, view=0, n_views=32):
    """Camera of build(name, view=view, n_views=n_views) without rebuilding the scene."""
    cfg = CONFIGS[name]
    if "cam" in cfg:
        c = dict(cfg["cam"])
        return make_camera(c.pop("eye"), cfg["W"], cfg["H"], **c)
    return ring_camera(view, n_views, cfg["W"], cfg["H"], **cfg["ring"])


def upstream_grads(W, H, seed=0):
    """N(0,1) upstream gradients for every differentiable output plane (fixed seed)."""
    rng = np.random.default_rng(777 + seed)
    return dict(color=rng.normal(0, 1, (3, H, W)).astype(np.float32),
                others=rng.normal(0, 1, (8, H, W)).astype(np.float32),
                refl=rng.normal(0, 1, (1, H, W)).astype(np.float32))
