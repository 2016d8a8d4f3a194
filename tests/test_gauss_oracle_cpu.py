"""CPU-only: the 3DGS oracle (oracle/gauss_oracle.py) against golden vectors produced by the
UNMODIFIED reference 3DGS extension on a B200 (tests/golden/make_golden_gauss.py).  The
reference has no tests or fixtures for this path; these vectors pin the
oracle.  Tolerances as for the surfel oracle: plain fp32 without fma contraction / MUFU, so
floats agree to ~1e-6 and an isolated discrete decision may flip on a pixel."""
import os

import numpy as np
import pytest
import torch

import scenes
from oracle import gauss_oracle as go

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gauss_tiny_ref.npz")


def _rel(a, b):
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-30))


@pytest.fixture(scope="module", params=[False, True], ids=["plain", "antialiasing"])
def run(request):
    torch.set_num_threads(1)
    aa = request.param
    G = np.load(GOLD)
    sc, cam = scenes.build("tiny")
    sc = dict(sc, scales3=G["scales3"], normals=G["normals"])
    rng = np.random.default_rng(991)          # tests/harness.py:gauss_grads
    W, H = cam["width"], cam["height"]
    g = dict(color=rng.normal(0, 1, (3, H, W)).astype(np.float32), invdepth=rng.normal(0, 1, (1, H, W)).astype(np.float32),
             normal=rng.normal(0, 1, (3, H, W)).astype(np.float32), refl=rng.normal(0, 1, (1, H, W)).astype(np.float32))
    o, gr = go.forward_backward(sc, cam, g, bg=tuple(G["bg"]), antialiasing=aa)
    return G, ("aa_" if aa else ""), o, gr


def test_integer_structures(run):
    G, t, o, _ = run
    assert o["rendered"] == int(G[t + "rendered"])
    assert np.array_equal(o["radii"], G[t + "radii"])
    assert np.array_equal(o["tiles_touched"], G[t + "state_tiles_touched"])
    assert np.array_equal(o["point_list"], G[t + "state_point_list"])
    assert np.array_equal(o["ranges"], G[t + "state_ranges"])
    assert np.array_equal(o["keys"] >> 32, G[t + "state_keys"] >> 32)
    assert np.abs((o["keys"] & 0xffffffff) - (G[t + "state_keys"] & 0xffffffff)).max() <= 1
    assert (o["n_contrib"].reshape(-1) != G[t + "state_n_contrib"].reshape(-1)).mean() < 1e-3


def test_images(run):
    G, t, o, _ = run
    for k in ("color", "invdepth", "normal_map", "refl_map"):
        d = np.abs(o[k] - G[t + k])
        assert np.median(d) <= 1e-6, k
        assert (d > 1e-5).mean() <= 1e-3, (k, d.max())


def test_gradients(run):
    """Autograd through the restated forward == the reference's hand-written backward, for every
    tensor whose reference gradient is a true derivative (all but means2D, see the oracle header)."""
    G, t, _, gr = run
    for ko, kg in (("means3D", "means3D"), ("sh", "sh"), ("opacity", "opacity"), ("scales", "scales"),
                   ("rotations", "rotations"), ("refl", "refl"), ("normals", "normals")):
        r = G[t + "grad_" + kg].reshape(gr[ko].shape)
        assert _rel(gr[ko], r) <= 2e-4, (ko, _rel(gr[ko], r))
