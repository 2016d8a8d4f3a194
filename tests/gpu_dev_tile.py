"""Dev harness (GPU): brute-force one tile's first entries with torch to see why hit pairs do not contribute."""
import sys
import math
import torch
import harness as hz
import scenes


def main(name="bench", tx=60, ty=34, K=600):
    newC = hz.new_surfel_C()
    sc_np, cam_np = scenes.build(name)
    sc, cam = hz.to_dev(sc_np), hz.to_dev(cam_np)
    W, H = cam_np["width"], cam_np["height"]
    f = hz.surfel_forward(newC, sc, cam)
    P = sc["means3D"].shape[0]
    st = hz.new_surfel_state(f, P, W, H)
    gx = (W + 15) // 16
    tile = ty * gx + tx
    r0, r1 = [int(x) for x in st["ranges"][tile]]
    ids = st["point_list"][r0:min(r1, r0 + K)].long()
    T = st["transMat"][ids].double()        # [K,9] Tu Tv Tw
    no = st["normal_opacity"][ids].double()
    xy = st["means2D"][ids].double()
    opa = no[:, 3]
    px = (tx * 16 + torch.arange(16, device="cuda")).double()
    py = (ty * 16 + torch.arange(16, device="cuda")).double()
    PX, PY = torch.meshgrid(px, py, indexing="xy")        # [16,16] (row=y)
    PX, PY = PX.reshape(1, -1), PY.reshape(1, -1)
    Tu, Tv, Tw = T[:, 0:3], T[:, 3:6], T[:, 6:9]
    k = PX.unsqueeze(-1) * Tw.unsqueeze(1) - Tu.unsqueeze(1)    # [K,256,3]
    l = PY.unsqueeze(-1) * Tw.unsqueeze(1) - Tv.unsqueeze(1)
    p = torch.cross(k, l, dim=-1)
    sx, sy = p[..., 0] / p[..., 2], p[..., 1] / p[..., 2]
    rho3d = sx * sx + sy * sy
    rho2d = 2.0 * ((xy[:, 0:1] - PX) ** 2 + (xy[:, 1:2] - PY) ** 2)
    rho = torch.minimum(rho3d, rho2d)
    depth = sx * Tw[:, 0:1] + sy * Tw[:, 1:2] + Tw[:, 2:3]
    alpha = torch.clamp(opa[:, None] * torch.exp(-0.5 * rho), max=0.99)
    ok = (alpha >= 1 / 255) & (depth >= 0.2)
    n_ok = ok.sum(1)
    n_alpha = (alpha >= 1 / 255).sum(1)
    print(f"tile ({tx},{ty}) list={r1-r0}; first {len(ids)} entries: with>=1 contributing px: {(n_ok>0).sum().item()}, "
          f"alpha-ok but depth<0.2 only: {((n_alpha>0)&(n_ok==0)).sum().item()}")
    c2 = 2 * torch.log(255 * opa)
    for j in range(0, min(len(ids), 40)):
        print(f"  j={j} id={int(ids[j])} opa={opa[j]:.3f} c2={c2[j]:.2f} xy=({xy[j,0]:.1f},{xy[j,1]:.1f}) Tw=({Tw[j,0]:.4f},{Tw[j,1]:.4f},{Tw[j,2]:.3f}) "
              f"min rho3d={rho3d[j].min():.2f} min rho2d={rho2d[j].min():.1f} n_alpha={int(n_alpha[j])} n_ok={int(n_ok[j])} mindepth={depth[j].min():.3f}")


if __name__ == "__main__":
    main(*(sys.argv[1:2]))
