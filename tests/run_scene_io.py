"""GPU harness: the point-cloud loader / writer (csrc/scene_io.cu).

  python tests/run_scene_io.py [P_kernel] [P_file] > gpurun_out/scene_io.jsonl

Line 1 -- the kernels alone at P_kernel Gaussians (default 3 M, degree 3, surfel model: 62 properties on disk,
59 mapped): ms per launch from the library's stage events, algorithmic bytes 4*(62+59) = 484 B per Gaussian each way
against the measured HBM copy peak, and next to them the torch ops the reference runs on the device for the same
result once its numpy columns are uploaded (transpose(1, 2).contiguous() of f_dc / f_rest + get_features' cat).
Line 2 -- load and save end to end on a file of P_file Gaussians (default 1 M = 248 MB, written under gpurun_out/):
this repo's load_ply / save_ply (file -> pinned -> device -> kernel) next to the oracle's restatement of the
reference's (per-property float64 numpy walk, upload of seven tensors; the reference's element-wise tuple list in
save_ply is NOT restated -- the oracle writes the structured array directly, which flatters the reference), with
the bit-exact parity gate of the same run."""
import ctypes as C
import json
import os
import shutil
import sys
import tempfile
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import harness as hz  # noqa: E402,F401
from run_update import batch_timed, hbm_peak  # noqa: E402


def kernels(P):
    import b200gs
    from b200gs import scene_io as sio
    names = sio.ply_attributes(3, 45, 2, 4)
    cols = sio._columns(names, 3)
    rec = torch.randn((P, len(names)), device="cuda")
    feat = torch.empty((P, 16, 3), device="cuda")
    arrs = {"xyz": torch.empty((P, 3), device="cuda"), "opacity": torch.empty((P, 1), device="cuda"),
            "refl_strength": torch.empty((P, 1), device="cuda"), "scaling": torch.empty((P, 2), device="cuda"),
            "rotation": torch.empty((P, 4), device="cuda")}
    fields = [(arrs[k].data_ptr(), arrs[k].shape[1], arrs[k].shape[1], cols[k]) for k in arrs]
    fields += [(feat.data_ptr(), 48, 48, cols["features_dc"] + cols["features_rest"])]
    L = sio._layout(len(names), fields)
    out = {"what": "ply kernels", "P": P, "properties": len(names), "mapped": 59}
    peak, _ = hbm_peak()
    for name, fn in (("unpack", lambda: b200gs.check(b200gs.ply_unpack(P, rec.data_ptr(), C.byref(L), None))),
                     ("pack", lambda: b200gs.check(b200gs.ply_pack(P, C.byref(L), rec.data_ptr(), None)))):
        for _ in range(3):
            fn()
        b200gs.timing_enable(1)
        b200gs.timing_read()
        for _ in range(20):
            fn()
        torch.cuda.synchronize()
        st = b200gs.timing_read()["scene_io"]
        b200gs.timing_enable(0)
        ms = st[0] / st[1]
        out[name + "_ms"], out[name + "_gbps"] = ms, 484.0 * P / ms / 1e6
        out[name + "_hbm_frac"] = out[name + "_gbps"] / peak
    out["hbm_peak_gbps"] = peak
    # what the reference runs on the device after uploading its columns: two transposes + the per-view cat
    dc, rest = torch.randn((P, 3, 1), device="cuda"), torch.randn((P, 3, 15), device="cuda")

    def ref_device_side():
        a, b = dc.transpose(1, 2).contiguous(), rest.transpose(1, 2).contiguous()
        return torch.cat((a, b), dim=1)
    out["reference_device_reshuffle_ms"] = batch_timed(ref_device_side)
    return out


def files(P):
    from b200gs.scene_io import load_ply, save_ply
    from oracle import scene_io_oracle as so
    os.makedirs("gpurun_out", exist_ok=True)
    d = tempfile.mkdtemp(dir="gpurun_out")
    try:
        m = so.random_model(P, 3, 2, seed=1)
        path = os.path.join(d, "point_cloud.ply")
        so.write_model(path, m)
        out = {"what": "ply files", "P": P, "file_MB": os.path.getsize(path) / 1e6}

        def ref_load():
            r = so.read_ply(path, 3)
            t = {k: torch.tensor(v, dtype=torch.float, device="cuda") for k, v in r.items()}
            torch.cuda.synchronize()
            return t

        def new_load():
            t = load_ply(path, 3)
            torch.cuda.synchronize()
            return t
        ref, new = ref_load(), new_load()
        out["bit_exact"] = all(torch.equal(ref[k].view(torch.int32), new[k].contiguous().view(torch.int32)) for k in ref)
        for name, fn in (("ref_load_s", ref_load), ("new_load_s", new_load)):
            ts = []
            for _ in range(3):
                t0 = time.perf_counter()
                fn()
                ts.append(time.perf_counter() - t0)
            out[name] = min(ts)
        out["load_speedup"] = out["ref_load_s"] / out["new_load_s"]
        p2, p3 = os.path.join(d, "new.ply"), os.path.join(d, "ref.ply")

        def ref_save():
            so.write_model(p3, {k: v.detach().cpu().numpy() for k, v in ref.items()})

        def new_save():
            save_ply(p2, new)
        for name, fn in (("ref_save_s", ref_save), ("new_save_s", new_save)):
            ts = []
            for _ in range(3):
                t0 = time.perf_counter()
                fn()
                ts.append(time.perf_counter() - t0)
            out[name] = min(ts)
        out["save_speedup"] = out["ref_save_s"] / out["new_save_s"]
        out["saved_files_identical"] = open(p2, "rb").read() == open(p3, "rb").read() == open(path, "rb").read()
        return out
    finally:
        shutil.rmtree(d, ignore_errors=True)


if __name__ == "__main__":
    a = [int(v) for v in sys.argv[1:]]
    print(json.dumps(kernels(a[0] if a else 3_000_000)), flush=True)
    torch.cuda.empty_cache()
    print(json.dumps(files(a[1] if len(a) > 1 else 1_000_000)), flush=True)
