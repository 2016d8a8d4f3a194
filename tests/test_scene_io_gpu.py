"""GPU parity of the point-cloud loader / writer (csrc/scene_io.cu through b200gs.scene_io)
against the oracle's restatement of GaussianModel.load_ply / save_ply (oracle/scene_io_oracle.py).  Byte work:
every comparison is bit-exact (compared as uint32 so that NaN payloads and signed zeros count)."""
import os

import numpy as np
import pytest
import torch

from oracle import scene_io_oracle as so

pytestmark = pytest.mark.gpu
KEYS = ("xyz", "features_dc", "features_rest", "opacity", "refl_strength", "scaling", "rotation")


def _same_bits(t, a):
    t = t.detach().cpu().contiguous().numpy()
    return t.shape == a.shape and np.array_equal(t.view(np.uint32), np.ascontiguousarray(a).view(np.uint32))


@pytest.mark.parametrize("P,deg,ns", [(0, 3, 2), (1, 3, 2), (63, 3, 2), (64, 0, 3), (65, 1, 2), (1000, 2, 3), (70001, 3, 2)])
def test_load_matches_oracle(tmp_path, P, deg, ns):
    from b200gs.scene_io import load_ply
    m = so.random_model(P, deg, ns, seed=P)
    if P:
        m["xyz"][0, 0], m["opacity"][-1, 0] = np.float32("nan"), np.float32(-0.0)
    path = str(tmp_path / "point_cloud.ply")
    so.write_model(path, m)
    ref = so.read_ply(path, deg)
    got = load_ply(path, deg)
    for k in KEYS:
        assert _same_bits(got[k], ref[k]), k
    assert got["features"].shape == (P, (deg + 1) ** 2, 3) and got["features"].is_contiguous()
    assert got["features_dc"].data_ptr() == got["features"].data_ptr()          # views of the packed SH block
    assert got["active_sh_degree"] == deg and "env_map" not in got


def test_load_goes_by_property_name(tmp_path):
    """Shuffled declaration order and properties the model does not know: the reader looks columns up by name."""
    from b200gs.scene_io import load_ply
    P, deg = 333, 3
    m = so.random_model(P, deg, 2, seed=5)
    names = so.attributes(3, 45, 2, 4) + ["confidence", "label"]
    np.random.default_rng(0).shuffle(names)
    path = str(tmp_path / "shuffled.ply")
    so.write_model(path, m, names)
    ref, got = so.read_ply(path, deg), load_ply(path, deg)
    for k in KEYS:
        assert _same_bits(got[k], ref[k]) and _same_bits(got[k], m[k]), k


@pytest.mark.parametrize("P,deg,ns", [(0, 3, 2), (1, 0, 2), (257, 3, 2), (4099, 1, 3)])
def test_save_writes_the_oracle_file(tmp_path, P, deg, ns):
    from b200gs.scene_io import load_ply, save_ply
    m = so.random_model(P, deg, ns, seed=P + 1)
    want, path = str(tmp_path / "want.ply"), str(tmp_path / "out" / "iteration_7" / "point_cloud.ply")
    so.write_model(want, m)
    save_ply(path, {k: torch.from_numpy(v).cuda() for k, v in m.items()})
    assert open(path, "rb").read() == open(want, "rb").read()
    # ... and from the loader's packed views (strided rows), without copies changing anything
    again = str(tmp_path / "again.ply")
    save_ply(again, load_ply(path, deg))
    assert open(again, "rb").read() == open(want, "rb").read()


def test_env_map_companion_round_trip(tmp_path):
    from b200gs.scene_io import load_ply, save_ply
    from cubemapencoder import CubemapEncoder
    m = so.random_model(50, 3, 2, seed=2)
    env = CubemapEncoder(output_dim=3, resolution=16).cuda()
    path = str(tmp_path / "point_cloud.ply")
    save_ply(path, {k: torch.from_numpy(v).cuda() for k, v in m.items()}, env_map=env)
    assert os.path.exists(str(tmp_path / "point_cloud.map"))
    data = torch.load(str(tmp_path / "point_cloud.map"))
    assert sorted(data) == ["params.Cubemap_failv", "params.Cubemap_texture"]    # what load_ply :330-333 reads
    got = load_ply(path, 3)
    assert got["env_map"].resolution == 16
    assert torch.equal(got["env_map"].params["Cubemap_texture"], env.params["Cubemap_texture"])


def test_errors(tmp_path):
    import b200gs
    from b200gs.scene_io import load_ply
    m = so.random_model(10, 3, 2, seed=3)
    path = str(tmp_path / "a.ply")
    so.write_model(path, m)
    with pytest.raises(AssertionError):                               # scene/gaussian_model.py:311
        load_ply(path, 2)
    raw = open(path, "rb").read()
    open(str(tmp_path / "cut.ply"), "wb").write(raw[:-8])
    with pytest.raises(ValueError):
        load_ply(str(tmp_path / "cut.ply"), 3)
    so.write_model(str(tmp_path / "norefl.ply"), m, [n for n in so.attributes(3, 45, 2, 4) if n != "refl"])
    with pytest.raises(ValueError):
        load_ply(str(tmp_path / "norefl.ply"), 3)
    L = b200gs.PlyLayout()
    L.stride, L.n_fields = 4, 2
    x = torch.zeros(8, device="cuda")
    L.field[0], L.field[1] = b200gs.PlyField(x.data_ptr(), 1, 1, 0), b200gs.PlyField(x.data_ptr(), 1, 1, 1)
    L.map[0], L.map[1] = 2, 2                                          # one record column feeding two arrays
    import ctypes as C
    with pytest.raises(b200gs.B200gsError):
        b200gs.check(b200gs.ply_unpack(2, x.data_ptr(), C.byref(L), None))
    L.map[1] = 9                                                       # outside the record
    with pytest.raises(b200gs.B200gsError):
        b200gs.check(b200gs.ply_unpack(2, x.data_ptr(), C.byref(L), None))


def test_full_size_round_trip_on_the_device():
    """BASELINE-size model (3 M Gaussians, degree 3): pack(unpack(records)) gives the records back with the unmapped
    normal columns zeroed; unpack(pack(arrays)) gives the arrays back.  No files, no oracle: size-independent property."""
    import ctypes as C
    import b200gs
    from b200gs import scene_io as sio
    P, K = 3_000_000, 16
    names = sio.ply_attributes(3, 45, 2, 4)
    cols = sio._columns(names, 3)
    rec = torch.randn((P, len(names)), device="cuda")
    feat = torch.empty((P, K, 3), device="cuda")
    arrs = {"xyz": torch.empty((P, 3), device="cuda"), "opacity": torch.empty((P, 1), device="cuda"),
            "refl_strength": torch.empty((P, 1), device="cuda"), "scaling": torch.empty((P, 2), device="cuda"),
            "rotation": torch.empty((P, 4), device="cuda")}
    fields = [(arrs[k].data_ptr(), arrs[k].shape[1], arrs[k].shape[1], cols[k]) for k in arrs]
    fields += [(feat.data_ptr(), 3, 48, cols["features_dc"]), (feat.data_ptr() + 12, 45, 48, cols["features_rest"])]
    L = sio._layout(len(names), fields)
    b200gs.check(b200gs.ply_unpack(P, rec.data_ptr(), C.byref(L), None))
    assert torch.equal(arrs["xyz"], rec[:, :3]) and torch.equal(arrs["rotation"], rec[:, -4:])
    assert torch.equal(feat[:, 1:, :], rec[:, 9:54].reshape(P, 3, 15).transpose(1, 2))
    back = torch.full_like(rec, 7.0)
    b200gs.check(b200gs.ply_pack(P, C.byref(L), back.data_ptr(), None))
    want = rec.clone()
    want[:, 3:6] = 0
    assert torch.equal(back, want)
