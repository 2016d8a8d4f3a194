"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of the reference's point-cloud IO, the checker for
csrc/scene_io.cu + python/b200gs/scene_io.py.  Never imported by the product path.

Follows GaussianModel.save_ply (scene/gaussian_model.py:239-257) and load_ply (:296-373) line by line; the file
format itself comes from `plyfile` (a pip dependency of the reference, environment.yml; NOT vendored and NOT
installed in this image), whose binary writer for PlyData([PlyElement.describe(structured f4 array, 'vertex')])
emits

    ply\\nformat binary_little_endian 1.0\\nelement vertex <P>\\nproperty float <name>\\n ... end_header\\n

followed by the structured array's bytes (P rows of little-endian fp32, declaration order).

PARITY UNPINNED for the header text: with plyfile absent there is no reference-written file to compare with, and
the reference ships no .ply fixture.  What IS checked against the reference's lines: the attribute order
(construct_list_of_attributes :225-238), the f_dc / f_rest flattening (transpose(1, 2).flatten, :244-245), the
name-sorted look-up and the reshape((P, 3, K-1)) + transpose(1, 2) of the reader (:308-316, :366-367), zeros for
nx ny nz (:243).  Values are copied, never computed: parity is bit-exact."""
import os

import numpy as np


def attributes(n_dc, n_rest, n_scale, n_rot):                # construct_list_of_attributes, :225-238
    l = ["x", "y", "z", "nx", "ny", "nz"]
    for i in range(n_dc):
        l.append("f_dc_{}".format(i))
    for i in range(n_rest):
        l.append("f_rest_{}".format(i))
    l.append("opacity")
    l.append("refl")
    for i in range(n_scale):
        l.append("scale_{}".format(i))
    for i in range(n_rot):
        l.append("rot_{}".format(i))
    return l


def write_ply(path, xyz, features_dc, features_rest, opacity, scaling, rotation, refl, names=None):
    """save_ply :239-257.  features_dc [P,1,3], features_rest [P,K-1,3] as the model holds them.
    `names`: optional permutation / superset of the attribute list (to exercise the reader's by-name look-up);
    attributes not in the reference's list are filled with their column index."""
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    P = xyz.shape[0]
    normals = np.zeros_like(xyz)
    f_dc = np.ascontiguousarray(features_dc.transpose(0, 2, 1).reshape(P, features_dc.shape[1] * features_dc.shape[2]))
    f_rest = np.ascontiguousarray(features_rest.transpose(0, 2, 1).reshape(P, features_rest.shape[1] * features_rest.shape[2]))
    std = attributes(f_dc.shape[1], f_rest.shape[1], scaling.shape[1], rotation.shape[1])
    table = np.concatenate((xyz, normals, f_dc, f_rest, opacity, refl, scaling, rotation), axis=1).astype(np.float32)
    names = std if names is None else names
    elements = np.empty(P, dtype=[(a, "<f4") for a in names])
    for k, a in enumerate(names):
        elements[a] = table[:, std.index(a)] if a in std else np.float32(k)
    header = "\n".join(["ply", "format binary_little_endian 1.0", "element vertex {}".format(P)]
                       + ["property float {}".format(a) for a in names] + ["end_header"]) + "\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(elements.tobytes())


def read_ply(path, max_sh_degree):
    """load_ply :296-373 (without the .map branch).  -> dict of fp32 arrays shaped as the nn.Parameters."""
    with open(path, "rb") as f:
        assert f.readline().strip() == b"ply"
        names, P = [], None
        while True:
            tok = f.readline().decode("ascii").split()
            if tok[0] == "end_header":
                break
            if tok[0] == "element" and P is None:
                P = int(tok[2])
            elif tok[0] == "property":
                names.append(tok[2])
        el = np.frombuffer(f.read(P * len(names) * 4), dtype=[(a, "<f4") for a in names])
    xyz = np.stack((np.asarray(el["x"]), np.asarray(el["y"]), np.asarray(el["z"])), axis=1)
    opacities = np.asarray(el["opacity"])[..., np.newaxis]
    refls = np.asarray(el["refl"])[..., np.newaxis]
    features_dc = np.zeros((P, 3, 1))
    for c in range(3):
        features_dc[:, c, 0] = np.asarray(el["f_dc_{}".format(c)])
    extra = sorted([n for n in names if n.startswith("f_rest_")], key=lambda x: int(x.split("_")[-1]))
    assert len(extra) == 3 * (max_sh_degree + 1) ** 2 - 3
    features_extra = np.zeros((P, len(extra)))
    for idx, a in enumerate(extra):
        features_extra[:, idx] = np.asarray(el[a])
    features_extra = features_extra.reshape((P, 3, (max_sh_degree + 1) ** 2 - 1))
    sn = sorted([n for n in names if n.startswith("scale_")], key=lambda x: int(x.split("_")[-1]))
    scales = np.zeros((P, len(sn)))
    for idx, a in enumerate(sn):
        scales[:, idx] = np.asarray(el[a])
    rn = sorted([n for n in names if n.startswith("rot")], key=lambda x: int(x.split("_")[-1]))
    rots = np.zeros((P, len(rn)))
    for idx, a in enumerate(rn):
        rots[:, idx] = np.asarray(el[a])
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)      # torch.tensor(..., dtype=torch.float)
    return {"xyz": f32(xyz), "features_dc": f32(features_dc.transpose(0, 2, 1)), "features_rest": f32(features_extra.transpose(0, 2, 1)),
            "opacity": f32(opacities), "scaling": f32(scales), "rotation": f32(rots), "refl_strength": f32(refls)}


def random_model(P, max_sh_degree=3, n_scale=2, seed=0):
    g = np.random.default_rng(seed)
    K = (max_sh_degree + 1) ** 2
    r = lambda *s: g.standard_normal(s).astype(np.float32)
    return {"xyz": r(P, 3), "features_dc": r(P, 1, 3), "features_rest": r(P, K - 1, 3), "opacity": r(P, 1),
            "scaling": r(P, n_scale), "rotation": r(P, 4), "refl_strength": r(P, 1)}


def write_model(path, m, names=None):
    write_ply(path, m["xyz"], m["features_dc"], m["features_rest"], m["opacity"], m["scaling"], m["rotation"], m["refl_strength"], names)
