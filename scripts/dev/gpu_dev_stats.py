"""Dev harness (GPU): workload statistics of a scene -- list lengths, early-termination depth,
hit-mask density -- to see where the compositing kernels spend their pair evaluations."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import sys
import torch
import harness as hz
import scenes


def al(x, a=128):
    return (x + a - 1) // a * a


def main(name="bench"):
    newC = hz.new_surfel_C()
    sc_np, cam_np = scenes.build(name)
    sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
    W, H = cam_np["width"], cam_np["height"]
    f = hz.surfel_forward(newC, sc, cam)
    P = sc["means3D"].shape[0]
    R = f["rendered"]
    st = hz.new_surfel_state(f, P, W, H)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    ranges = st["ranges"].long()
    n = (ranges[:, 1] - ranges[:, 0])
    last = st["n_contrib"][0].long().reshape(H, W)
    pad = torch.zeros((gy * 16, gx * 16), dtype=torch.long, device="cuda")
    pad[:H, :W] = last
    t = pad.reshape(gy, 16, gx, 16).permute(0, 2, 1, 3)            # [gy,gx,16,16]
    tile_max = t.reshape(gy * gx, -1).max(1).values                 # where the tile's loop can stop
    # warp blocks: 8x4 pixels; warp w: col w&1, row w>>1
    wb = t.reshape(gy, gx, 4, 4, 2, 8).permute(0, 1, 2, 4, 3, 5).reshape(gy * gx, 8, 32)
    warp_max = wb.max(2).values                                     # [tiles, 8]
    print(f"{name}: P={P} R={R} tiles={gx*gy} mean list={n.float().mean():.0f} max list={int(n.max())}")
    print(f"  sum over tiles of list length          = {int(n.sum())}")
    print(f"  sum over tiles of max last_contributor = {int(tile_max.sum())}  ({tile_max.sum().item()/max(n.sum().item(),1):.3f} of lists)")
    print(f"  (warp,entry) pairs up to warp's last   = {int(warp_max.sum())}  ({warp_max.sum().item()/max(8*n.sum().item(),1):.3f} of 8*lists)")
    print(f"  mean last_contributor per pixel        = {last.float().mean():.1f}")
    import b200gs
    b200gs.cull_stats(reset=True)
    b200gs.debug_cull_audit(1)
    hz.surfel_forward(newC, sc, cam)
    b200gs.debug_cull_audit(0)
    s = b200gs.cull_stats()
    print(f"  audit: hit pairs visited={s[0]} with>=1 lane={s[1]} ({s[1]/max(s[0],1):.3f}) lane contributions={s[2]} "
          f"({s[2]/max(32*s[0],1):.4f} of lanes) sum busiest-lane per (warp,batch)={s[3]} warp-batches={s[4]} "
          f"-> per-lane-iteration model: {s[3]/max(s[0],1):.3f} of visited pairs")


if __name__ == "__main__":
    main(*(sys.argv[1:2]))
