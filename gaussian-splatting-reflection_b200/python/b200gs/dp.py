"""View-parallel data parallelism for the rendering path.

The path shards across camera views only: Gaussians are replicated on every GPU, rank r
renders views r, r+n, ... of a batch, and the per-Gaussian gradients are summed with ONE
all-reduce.  To make that all-reduce zero-copy, the backward writes its per-parameter
gradient tensors straight into one flat fp32 arena (`GradArena`), laid out as contiguous
blocks  [means3D 3P | sh 3MP | opacity P | scales 2P | rotations 4P | refl P]  (59 floats per
Gaussian at SH degree 3: the surfel model's parameter set, scene/gaussian_model.py:196-204),
so NCCL reduces the very memory the kernels wrote.  No collective exists in the reference
(it is single-GPU); this module is the whole multi-GPU surface.
"""
import torch

BLOCKS = (("means3D", 3), ("sh", None), ("opacity", 1), ("scales", 2), ("rotations", 4), ("refl", 1))


def shard_views(n_views, rank, world):
    """Views rendered by `rank`: r, r+world, ...  (independent given replicated Gaussians)."""
    return list(range(rank, n_views, world))


class GradArena:
    """One flat fp32 buffer holding the six parameter-gradient blocks of P Gaussians."""

    def __init__(self, P, sh_coeffs, device, accumulate=False):
        """accumulate=False: every backward overwrites the arena (one view per step).
        accumulate=True: backwards ADD to it -- zero_() once per batch, render the rank's views,
        then all_reduce() (a 32-view batch sharded over the ranks)."""
        self.P, self.M = int(P), int(sh_coeffs)
        self.accumulate = bool(accumulate)
        self.widths = {k: (w if w is not None else 3 * self.M) for k, w in BLOCKS}
        self.offsets, off = {}, 0
        for k, _ in BLOCKS:
            self.offsets[k] = off
            off += self.widths[k] * self.P
        self.flat = torch.zeros(off, dtype=torch.float32, device=device)

    @property
    def floats_per_gaussian(self):
        return sum(self.widths.values())

    def view(self, name):
        w, o = self.widths[name], self.offsets[name]
        t = self.flat[o:o + w * self.P]
        return t.view(self.P, self.M, 3) if name == "sh" else t.view(self.P, w)

    def views(self):
        return {k: self.view(k) for k, _ in BLOCKS}

    def zero_(self):
        self.flat.zero_()
        return self

    def all_reduce(self, group=None, async_op=False):
        """Sum the arena over all ranks in place (NCCL on GPUs, gloo in the CPU tests)."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
