"""CPU checks of the point-cloud IO: the oracle (oracle/scene_io_oracle.py) against the
reference's documented layout (scene/gaussian_model.py:225-257, :296-373), and the host logic of
python/b200gs/scene_io.py (header parser, column map, error behaviour).  No kernel runs here."""
import io

import numpy as np
import pytest

from oracle import scene_io_oracle as so


@pytest.mark.parametrize("P,deg,ns", [(0, 3, 2), (1, 3, 2), (77, 0, 3), (300, 1, 2), (129, 3, 3)])
def test_oracle_round_trip_is_the_identity(tmp_path, P, deg, ns):
    m = so.random_model(P, deg, ns, seed=P + deg)
    path = str(tmp_path / "pc" / "point_cloud.ply")
    so.write_model(path, m)
    back = so.read_ply(path, deg)
    for k, v in m.items():
        assert back[k].shape == v.shape and back[k].dtype == np.float32
        assert np.array_equal(back[k].view(np.uint32), v.view(np.uint32)), k


def test_oracle_file_layout(tmp_path):
    """Attribute order of construct_list_of_attributes, channel-major f_rest on disk, zero normals."""
    P, deg = 5, 2
    m = so.random_model(P, deg, 2, seed=3)
    path = str(tmp_path / "a.ply")
    so.write_model(path, m)
    raw = open(path, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode().split("\n")
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", "element vertex 5"]
    names = [l.split()[2] for l in lines[3:] if l]
    assert names == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"] + [f"f_rest_{i}" for i in range(24)] + \
        ["opacity", "refl", "scale_0", "scale_1", "rot_0", "rot_1", "rot_2", "rot_3"]
    rec = np.frombuffer(body, "<f4").reshape(P, len(names))
    assert np.all(rec[:, 3:6] == 0)
    K1 = 8
    for ch in range(3):
        for j in range(K1):
            assert np.array_equal(rec[:, 9 + ch * K1 + j], m["features_rest"][:, j, ch])
    assert np.array_equal(rec[:, 6:9], m["features_dc"][:, 0, :])
    assert np.array_equal(rec[:, names.index("refl")], m["refl_strength"][:, 0])


def test_host_header_and_column_map_agree_with_the_oracle(tmp_path):
    from b200gs import scene_io as sio
    m = so.random_model(9, 3, 2, seed=1)
    path = str(tmp_path / "b.ply")
    so.write_model(path, m)
    raw = open(path, "rb").read()
    names = so.attributes(3, 45, 2, 4)
    assert raw.startswith(sio.ply_header(9, sio.ply_attributes(3, 45, 2, 4)))
    with open(path, "rb") as f:
        P, props, off = sio.read_ply_header(f)
    assert (P, props) == (9, names) and raw[off - 11:off] == b"end_header\n"
    cols = sio._columns(props, 3)
    rec = np.frombuffer(raw[off:], "<f4").reshape(9, len(names))
    ref = so.read_ply(path, 3)
    for key in ("xyz", "features_dc", "features_rest", "opacity", "refl_strength", "scaling", "rotation"):
        assert np.array_equal(rec[:, cols[key]].reshape(ref[key].shape), ref[key]), key


def test_host_header_errors():
    from b200gs import scene_io as sio
    ok = b"ply\nformat binary_little_endian 1.0\ncomment made by hand\nelement vertex 2\nproperty float x\nproperty float32 y\n" \
         b"element face 0\nproperty list uchar int vertex_indices\nend_header\n"
    assert sio.read_ply_header(io.BytesIO(ok))[:2] == (2, ["x", "y"])
    with pytest.raises(ValueError):
        sio.read_ply_header(io.BytesIO(b"plx\n"))
    with pytest.raises(ValueError):
        sio.read_ply_header(io.BytesIO(b"ply\nformat ascii 1.0\nelement vertex 1\nproperty float x\nend_header\n"))
    with pytest.raises(ValueError):
        sio.read_ply_header(io.BytesIO(b"ply\nformat binary_little_endian 1.0\nelement vertex 1\nproperty double x\nend_header\n"))
    with pytest.raises(ValueError):
        sio.read_ply_header(io.BytesIO(b"ply\nformat binary_little_endian 1.0\nelement vertex 1\nproperty float x\n"))
    names = so.attributes(3, 45, 2, 4)
    with pytest.raises(AssertionError):                               # scene/gaussian_model.py:311
        sio._columns(names, 2)
    with pytest.raises(ValueError):                                   # plydata.elements[0]["refl"] on a file without it
        sio._columns([n for n in names if n != "refl"], 3)
