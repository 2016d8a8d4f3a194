"""Multi-GPU harness (torchrun, one rank per GPU): the gradient arena's own all-reduce kernel (csrc/peer_allreduce.cu:
NVLS multimem.ld_reduce / multimem.st over the multicast mapping, or peer-to-peer loads / stores) against ncclAllReduce
on the same values, and their times.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tests/run_dp_check.py [P]
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import harness as hz  # noqa: E402,F401  (import paths)
from b200gs.dp import GradArena  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    P = int(sys.argv[1]) if len(sys.argv) > 1 else 3_000_000
    out = {"world": world, "P": P, "bytes": P * 59 * 4}
    arena = GradArena(P, 16, "cuda", accumulate=True, symmetric=True)
    out["algo"] = arena.algo
    out["multicast"] = bool(arena._mc)
    g = torch.Generator(device="cuda").manual_seed(100 + rank)
    vals = torch.randn(arena.numel, device="cuda", generator=g)
    arena.flat.copy_(vals)
    ref = vals.clone()
    dist.all_reduce(ref)
    arena.all_reduce()
    torch.cuda.synchronize()
    err = float((arena.flat - ref).abs().max())
    scale = float(ref.abs().max())
    out["max_abs_err"], out["max_abs_ref"] = err, scale
    ok = err <= 1e-5 * scale * world
    # every rank must hold the same bits
    chk = arena.flat.double().sum().reshape(1)
    all_chk = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(all_chk, chk)
    out["ranks_agree"] = bool(all(float(c) == float(all_chk[0]) for c in all_chk))

    def timed(fn, n=20, w=5):
        for _ in range(w):
            fn()
        ts = []
        for _ in range(n):
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        t = torch.tensor(sorted(ts)[len(ts) // 2], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)
    out["own_kernel_ms"] = timed(arena.all_reduce)
    out["sweep_ctas_ms"] = {}
    for ctas in (8, 16, 32, 64, 148, 296, 1184):
        arena.ctas = ctas
        out["sweep_ctas_ms"][ctas] = round(timed(arena.all_reduce, n=8, w=2), 4)
    arena.ctas = 0
    if arena._mc:                       # the same kernel family without the multicast object: peer-to-peer loads / stores
        mc, arena._mc = arena._mc, 0
        arena.flat.copy_(vals)
        arena.all_reduce()
        torch.cuda.synchronize()
        out["p2p_max_abs_err"] = float((arena.flat - ref).abs().max())
        out["p2p_ms"] = timed(arena.all_reduce, n=8, w=2)
        arena._mc = mc
    plain = torch.zeros(arena.numel, device="cuda")
    out["nccl_ms"] = timed(lambda: dist.all_reduce(plain))
    out["nccl_on_symmetric_ms"] = timed(lambda: dist.all_reduce(arena.flat))
    for k in ("own_kernel_ms", "nccl_ms"):
        out[k.replace("_ms", "_bus_gbs")] = 2 * (world - 1) / world * out["bytes"] / (out[k] * 1e-3) / 1e9
    out["ok"] = bool(ok and out["ranks_agree"])
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.destroy_process_group()
    if not out["ok"]:
        sys.exit(1)


if __name__ == "__main__":
    main()
