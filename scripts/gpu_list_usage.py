"""GPU diagnostic: how much of the sorted per-tile instance lists does the compositing actually use?
For a config / view: per tile the list length (culled-emission mode) and the position of the last contributor of any
of its pixels; then a simulation of depth-sliced binning (slice boundaries = global depth quantiles of the visible
Gaussians; a tile stops receiving instances after the slice in which all its pixels saturated)."""
import json, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import torch
import harness as hz
import scenes

for name in sys.argv[1:] or ["bench", "c3", "c4", "c2"]:
    sc_np, cam_np = scenes.build(name)
    sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
    P, W, H = sc["means3D"].shape[0], cam_np["width"], cam_np["height"]
    C = hz.new_surfel_C()
    f = hz.surfel_forward(C, sc, cam, exact=False)
    st = hz.new_surfel_state(f, P, W, H)
    ranges = st["ranges"].long()
    T = ranges.shape[0]
    gx = (W + 15) // 16
    lens = ranges[:, 1] - ranges[:, 0]
    last = st["n_contrib"][0].long().reshape(H, W)
    # per tile: max last contributor over its pixels, and whether every pixel is saturated (final T < 1e-4 reached => done)
    Hp, Wp = (H + 15) // 16 * 16, gx * 16
    pad = torch.zeros((Hp, Wp), dtype=torch.long, device="cuda"); pad[:H, :W] = last
    tl_last = pad.reshape(Hp // 16, 16, gx, 16).permute(0, 2, 1, 3).reshape(T, 256).max(1).values
    Tfin = st["accum"][0].reshape(H, W)
    padT = torch.zeros((Hp, Wp), device="cuda"); padT[:H, :W] = Tfin
    # a pixel is `done` when the NEXT blend would push T below 1e-4; proxy: final T < 2e-4 ... use tile_last < len as "stopped early"
    early = tl_last < lens
    R = int(lens.sum()); used = int(tl_last.sum())
    out = {"config": name, "R_emitted": R, "tiles": T, "mean_len": R / T, "used_entries": used, "used_frac": used / max(R, 1),
           "tiles_stopped_early_frac": float(early.float().mean())}
    # depth of each tile's last used entry
    keys = st["keys"]; depth_bits = (keys & 0xffffffff)
    tile_of = torch.repeat_interleave(torch.arange(T, device="cuda"), lens)
    idx_last = (ranges[:, 0] + (tl_last - 1).clamp(min=0))
    d_last = torch.where(tl_last > 0, depth_bits[idx_last.clamp(max=max(R - 1, 0))], torch.zeros_like(idx_last))
    d_last = torch.where(early, d_last, torch.full_like(d_last, 0x7fffffff))      # never saturated: needs everything
    vis_depth = depth_bits.unique() if R < 5_000_000 else None
    alld = depth_bits
    for fr in ([0.25, 1.0], [0.125, 0.25, 1.0], [0.1, 0.2, 0.4, 1.0], [0.06, 0.12, 0.25, 0.5, 1.0], [0.05, 0.1, 0.15, 0.2, 0.3, 0.5, 1.0]):
        # boundaries = quantiles of the INSTANCE depth distribution (equal-ish instance counts up to the boundary)
        q = torch.tensor(fr[:-1], device="cuda", dtype=torch.float64)
        srt = alld.sort().values
        bounds = srt[(q * (R - 1)).long()]
        bounds = torch.cat([bounds, torch.tensor([0x7fffffff], device="cuda")])
        k = torch.searchsorted(bounds, d_last.clamp(max=0x7ffffffe), right=False)        # slice in which the tile finishes
        stop_depth = bounds[k]
        processed = int((alld <= stop_depth[tile_of]).sum())
        out[f"sliced_{fr}"] = processed / max(R, 1)
    print(json.dumps(out), flush=True)
    del f, st
    torch.cuda.empty_cache()
