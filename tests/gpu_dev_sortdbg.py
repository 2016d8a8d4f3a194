"""Dev harness (GPU): compare binning state new vs reference for the gauss cov3D_precomp variant."""
import torch
import harness as hz
import scenes

sc_np, cam_np = scenes.build("tiny")
sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
sc["scales"] = torch.cat([sc["scales"], sc["scales"][:, :1]], 1).contiguous() if sc["scales"].shape[1] == 2 else sc["scales"]
import test_gauss_gpu as tg
sc, cam, grads = tg._setup("tiny")
newC, refC = hz.new_gauss_C(), hz.ref_module("ref_gauss_C")
P = sc["means3D"].shape[0]
torch.manual_seed(0)
col = torch.rand((P, 3), device="cuda")
nbad = 0
for trial in range(60):
    cov = torch.rand((P, 6), device="cuda") * 1e-3
    cov[:, [0, 3, 5]] += 2e-3
    if trial % 3 == 0:
        hz.gauss_forward(newC, sc, cam, colors_precomp=col)
    fn, fr = hz.gauss_forward(newC, sc, cam, cov3D_precomp=cov), hz.gauss_forward(refC, sc, cam, cov3D_precomp=cov)
    W, H = cam["width"], cam["height"]
    err = hz.max_abs(fn["color"], fr["color"])
    if err <= 1e-5:
        continue
    nbad += 1
    sn, sr = hz.new_gauss_state(fn, P, W, H), hz.ref_gauss_state(fr, P, W, H)
    print("trial", trial, "R", fn["rendered"], fr["rendered"], "img", err)
    for k in ("tiles_touched", "point_offsets", "keys", "point_list", "ranges", "n_contrib"):
        eq = torch.equal(sn[k], sr[k])
        print("  ", k, eq)
        if not eq and k in ("keys", "point_list"):
            bad = (sn[k] != sr[k]).nonzero().flatten()
            print("     first bad", bad[:5].tolist(), "n bad", bad.numel())
            i = int(bad[0])
            print("     new", sn[k][max(0, i - 2):i + 4].tolist(), "ref", sr[k][max(0, i - 2):i + 4].tolist())
    bad = ((fn["color"] - fr["color"]).abs().amax(0) > 1e-5).nonzero()
    print("   bad pixels", bad.shape[0], bad.tolist())
    nn, nr = sn["n_contrib"].reshape(H, W), sr["n_contrib"].reshape(H, W)
    print("   n_contrib new", nn[bad[:, 0], bad[:, 1]].tolist())
    print("   n_contrib ref", nr[bad[:, 0], bad[:, 1]].tolist())
    print("   ranges[0:2]", sn["ranges"][:2].tolist())
    if nbad >= 2:
        break
    fn2 = hz.gauss_forward(newC, sc, cam, cov3D_precomp=cov)
    print("   rerun new vs ref", hz.max_abs(fn2["color"], fr["color"]), "new vs new", hz.max_abs(fn2["color"], fn["color"]))
print("bad trials", nbad)
