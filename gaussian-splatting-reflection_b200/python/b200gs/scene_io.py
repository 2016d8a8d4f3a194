"""Scene IO next to the path ("next" row 4): point_cloud.ply and its `.map` companion,
straight into the arrays the preprocess kernels read.

Mirrors GaussianModel.load_ply / save_ply (scene/gaussian_model.py:296-373, :225-262):

  load_ply(path, max_sh_degree)  -> dict with the reference's parameter tensors on the device
      xyz [P,3], features_dc [P,1,3], features_rest [P,K-1,3], opacity [P,1], scaling [P,S], rotation [P,4],
      refl_strength [P,1]  (+ "features" [P,K,3]: ONE packed block whose column ranges features_dc and
      features_rest are -- what `get_features` concatenates every view in the reference is already laid out),
      "env_map": a CubemapEncoder when `<path>.map` exists (torch.load of its state_dict, as the reference).
  save_ply(path, tensors, env_map=None)

The reference reads the file with plyfile and walks it property by property through float64 numpy columns
(62 strided passes over the records for the degree-3 model), transposes f_rest on the device and uploads
seven tensors.  Here the header is parsed on the host (a few hundred bytes), the vertex block goes from the
file into pinned memory and to the device in ONE copy, and csrc/scene_io.cu scatters it (b200gs_ply_unpack);
save_ply runs the same kernel the other way (b200gs_ply_pack).  The values are never touched: every tensor is
the same bits the reference builds (fp32 -> fp64 -> fp32 is the identity).

Only what the reference's own writer produces is read: binary_little_endian, all-fp32 vertex properties,
the vertex element first.  Anything else raises ValueError (the reference would go through plyfile's general
reader; that path is out of scope and says so rather than guessing)."""
import ctypes as C
import os

import numpy as np
import torch

import b200gs

_FLOAT_NAMES = ("float", "float32")


def read_ply_header(f):
    """Parse the header of an open binary file.  -> (P, [property names of the first element], data offset)."""
    if f.readline().strip() != b"ply":
        raise ValueError("not a PLY file")
    fmt, count, props, first_element_done = None, None, [], False
    while True:
        raw = f.readline()
        if not raw:
            raise ValueError("PLY header ends before end_header")
        line = raw.decode("ascii").strip()
        if line == "end_header":
            break
        tok = line.split()
        if not tok or tok[0] in ("comment", "obj_info"):
            continue
        if tok[0] == "format":
            fmt = tok[1]
        elif tok[0] == "element":
            if count is None:
                count = int(tok[2])
            else:
                first_element_done = True
        elif tok[0] == "property" and not first_element_done:
            if count is None:
                raise ValueError("PLY property before any element")
            if tok[1] == "list" or tok[1] not in _FLOAT_NAMES:
                raise ValueError(f"unsupported PLY property type '{tok[1]}' (the writer of scene/gaussian_model.py:239-257 emits fp32 only)")
            props.append(tok[2])
    if fmt != "binary_little_endian":
        raise ValueError(f"unsupported PLY format '{fmt}' (binary_little_endian only)")
    if count is None:
        raise ValueError("PLY file without elements")
    return count, props, f.tell()


def _column(props, name):
    try:
        return props.index(name)
    except ValueError:
        raise ValueError(f"no field of name {name}") from None      # numpy's message for plydata.elements[0][name]


def _sorted_prefixed(props, prefix):
    names = [p for p in props if p.startswith(prefix)]
    return sorted(names, key=lambda x: int(x.split("_")[-1]))         # scene/gaussian_model.py:310, :319, :325


def _layout(stride, fields):
    """fields: [(device pointer, cols, row_stride, [record columns])]"""
    L = b200gs.PlyLayout()
    L.stride, L.n_fields = stride, len(fields)
    first = 0
    for k, (data_ptr, cols, row_stride, columns) in enumerate(fields):
        assert len(columns) == cols
        L.field[k] = b200gs.PlyField(data_ptr, cols, row_stride, first)
        for c, col in enumerate(columns):
            L.map[first + c] = col
        first += cols
    return L


def _columns(props, max_sh_degree):
    """Record columns of every parameter, by property NAME as load_ply looks them up (:299-328)."""
    K = (max_sh_degree + 1) ** 2
    rest = _sorted_prefixed(props, "f_rest_")
    assert len(rest) == 3 * K - 3                                     # scene/gaussian_model.py:311
    rest_cols = [_column(props, n) for n in rest]
    # features_extra.reshape(P, 3, K-1) then transpose(1, 2): features_rest[:, j, ch] = f_rest_{ch*(K-1)+j}
    rest_t = [rest_cols[ch * (K - 1) + j] for j in range(K - 1) for ch in range(3)]
    return {
        "xyz": [_column(props, n) for n in ("x", "y", "z")],
        "features_dc": [_column(props, f"f_dc_{c}") for c in range(3)],
        "features_rest": rest_t,
        "opacity": [_column(props, "opacity")],
        "refl_strength": [_column(props, "refl")],
        "scaling": [_column(props, n) for n in _sorted_prefixed(props, "scale_")],
        "rotation": [_column(props, n) for n in _sorted_prefixed(props, "rot")],
    }


def load_ply(path, max_sh_degree=3, device="cuda", stream=None):
    with open(path, "rb") as f:
        P, props, _ = read_ply_header(f)
        stride = len(props)
        if stride > b200gs.PLY_MAX_COLUMNS:
            raise ValueError(f"{stride} vertex properties (at most {b200gs.PLY_MAX_COLUMNS})")
        cols = _columns(props, max_sh_degree)
        host = torch.empty((P, stride), dtype=torch.float32, pin_memory=True)
        want = P * stride * 4
        got = f.readinto(memoryview(host.numpy()).cast("B")) if want else 0
        if got != want:
            raise ValueError(f"PLY vertex block truncated: {got} of {want} bytes")
    K = (max_sh_degree + 1) ** 2
    with torch.cuda.device(device):
        s = stream or torch.cuda.current_stream()
        with torch.cuda.stream(s):
            records = host.to(device, non_blocking=True)
            out = {"xyz": torch.empty((P, 3), device=device), "features": torch.empty((P, K, 3), device=device),
                   "opacity": torch.empty((P, 1), device=device), "refl_strength": torch.empty((P, 1), device=device),
                   "scaling": torch.empty((P, len(cols["scaling"])), device=device),
                   "rotation": torch.empty((P, len(cols["rotation"])), device=device)}
            feat = out["features"]
            fields = [(out["xyz"].data_ptr(), 3, 3, cols["xyz"]),
                      (feat.data_ptr(), 3 * K, 3 * K, cols["features_dc"] + cols["features_rest"]),   # whole SH rows
                      (out["opacity"].data_ptr(), 1, 1, cols["opacity"]),
                      (out["refl_strength"].data_ptr(), 1, 1, cols["refl_strength"]),
                      (out["scaling"].data_ptr(), out["scaling"].shape[1], out["scaling"].shape[1], cols["scaling"]),
                      (out["rotation"].data_ptr(), out["rotation"].shape[1], out["rotation"].shape[1], cols["rotation"])]
            fields = [fl for fl in fields if fl[1] > 0]
            b200gs.check(b200gs.ply_unpack(P, b200gs.ptr(records), C.byref(_layout(stride, fields)), s.cuda_stream))
    out["features_dc"], out["features_rest"] = feat[:, :1, :], feat[:, 1:, :]
    out["active_sh_degree"] = max_sh_degree                           # scene/gaussian_model.py:373
    map_path = path.replace(".ply", ".map")                           # :328-333
    if os.path.exists(map_path):
        out["env_map"] = load_env_map(map_path, device)
    return out


def load_env_map(map_path, device="cuda"):
    from cubemapencoder import CubemapEncoder
    data = torch.load(map_path, map_location=device)
    env = CubemapEncoder(output_dim=3, resolution=data["params.Cubemap_texture"].shape[2]).to(device)
    env.load_state_dict(data)
    return env


def ply_attributes(n_dc, n_rest, n_scale, n_rot):
    """construct_list_of_attributes (scene/gaussian_model.py:225-238)."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += [f"f_dc_{i}" for i in range(n_dc)] + [f"f_rest_{i}" for i in range(n_rest)]
    names += ["opacity", "refl"]
    names += [f"scale_{i}" for i in range(n_scale)] + [f"rot_{i}" for i in range(n_rot)]
    return names


def ply_header(P, names):
    """The header plyfile writes for PlyData([PlyElement.describe(elements, 'vertex')]) with f4 fields."""
    lines = ["ply", "format binary_little_endian 1.0", f"element vertex {P}"]
    lines += [f"property float {n}" for n in names] + ["end_header"]
    return ("\n".join(lines) + "\n").encode("ascii")


def save_ply(path, tensors, env_map=None, stream=None):
    """tensors: xyz, features_dc [P,1,3], features_rest [P,K-1,3] (any strides: views of one packed block are
    read in place), opacity, scaling, rotation, refl_strength -- all on one device."""
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)                                 # mkdir_p, :240
    t = {k: tensors[k].detach() for k in ("xyz", "features_dc", "features_rest", "opacity", "refl_strength", "scaling", "rotation")}
    P = t["xyz"].shape[0]
    dev = t["xyz"].device
    n_rest = t["features_rest"].shape[1] * t["features_rest"].shape[2]
    names = ply_attributes(t["features_dc"].shape[1] * t["features_dc"].shape[2], n_rest, t["scaling"].shape[1], t["rotation"].shape[1])
    stride = len(names)
    cols = _columns(names, int(round((n_rest // 3 + 1) ** 0.5)) - 1)

    def rows(x, inner):
        """(pointer, row stride) of a [P, ...] tensor whose rows are `inner` contiguous floats; copies otherwise."""
        x = x.float()
        if P > 1 and x[0].is_contiguous() and x.stride(0) >= inner:
            rs = x.stride(0)
        else:
            x, rs = x.contiguous(), inner
        keep.append(x)
        return x.data_ptr(), rs

    keep = []
    with torch.cuda.device(dev):
        s = stream or torch.cuda.current_stream()
        with torch.cuda.stream(s):
            fields = []
            for key in ("xyz", "features_dc", "features_rest", "opacity", "refl_strength", "scaling", "rotation"):
                inner = t[key][0].numel() if P else int(np.prod(t[key].shape[1:]))
                if inner == 0:
                    continue
                p, rs = rows(t[key], inner)
                if fields and fields[-1][2] == rs and p == fields[-1][0] + 4 * fields[-1][1] and fields[-1][1] + inner <= rs:
                    q = fields.pop()                                  # features_rest right behind features_dc: one row
                    fields.append((q[0], q[1] + inner, rs, q[3] + cols[key]))
                else:
                    fields.append((p, inner, rs, cols[key]))
            records = torch.empty((P, stride), device=dev)
            b200gs.check(b200gs.ply_pack(P, C.byref(_layout(stride, fields)), b200gs.ptr(records), s.cuda_stream))
            host = torch.empty((P, stride), dtype=torch.float32, pin_memory=True)
            host.copy_(records, non_blocking=True)
            s.synchronize()
    with open(path, "wb") as f:
        f.write(ply_header(P, names))
        if P:
            f.write(memoryview(host.numpy()).cast("B"))
    if env_map is not None:                                           # :259-261
        torch.save(env_map.state_dict(), path.replace(".ply", ".map"))
