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

    def __init__(self, P, sh_coeffs, device, accumulate=False, symmetric=False, group=None):
        """accumulate=False: every backward overwrites the arena (one view per step).
        accumulate=True: backwards ADD to it -- zero_() once per batch, render the rank's views,
        then all_reduce() (a 32-view batch sharded over the ranks).
        symmetric=True (collective: every rank of `group` must construct its arena at the same point): the buffer
        is allocated in symmetric memory and all_reduce() runs this library's own NVLS / peer-to-peer kernel
        (csrc/peer_allreduce.cu) on it instead of ncclAllReduce."""
        self.P, self.M = int(P), int(sh_coeffs)
        self.accumulate = bool(accumulate)
        self.widths = {k: (w if w is not None else 3 * self.M) for k, w in BLOCKS}
        self.offsets, off = {}, 0
        for k, _ in BLOCKS:
            self.offsets[k] = off
            off += self.widths[k] * self.P
        self.numel = off
        self.symmetric, self.algo, self._hdl = False, "ncclAllReduce", None
        if symmetric:
            try:
                self._alloc_symmetric(off, device, group)
            except Exception as e:  # noqa: BLE001 -- no symmetric memory on this system / torch build (every rank fails alike)
                import warnings
                warnings.warn(f"b200gs.dp.GradArena: symmetric memory unavailable ({type(e).__name__}: {e}); "
                              "the arena is an ordinary tensor and is reduced with ncclAllReduce")
                self.symmetric, self._hdl = False, None
                self.algo = "ncclAllReduce (symmetric memory unavailable)"
                self.flat = torch.zeros(off, dtype=torch.float32, device=device)
        else:
            self.flat = torch.zeros(off, dtype=torch.float32, device=device)

    def _alloc_symmetric(self, off, device, group):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        group = group if group is not None else dist.group.WORLD
        try:
            symm.enable_symm_mem_for_group(group.group_name)      # needed by older torch, a deprecated no-op in newer
        except Exception:  # noqa: BLE001
            pass
        padded = (off + 3) // 4 * 4                                # the kernel moves 16-B vectors
        buf = symm.empty(padded, dtype=torch.float32, device=torch.device(device))
        buf.zero_()
        self._hdl = symm.rendezvous(buf, group)
        self._buf = buf
        self.flat = buf[:off]
        self.symmetric = True
        self._world, self._rank = self._hdl.world_size, self._hdl.rank
        # where this tensor sits inside the handle's buffer (the multicast mapping covers the same range)
        self._offset = buf.data_ptr() - int(self._hdl.buffer_ptrs[self._rank])
        self._mc = int(self._hdl.multicast_ptr) + self._offset if int(self._hdl.multicast_ptr) else 0
        # Measured on 8 x B200 / NVSwitch, 708 MB (tests/run_dp_check.py, profiles/r2i_dp_check_n*.json): two ranks --
        # peer-to-peer loads / stores 1.09 ms, ncclAllReduce 1.31, multimem 1.75 (nothing for the switch to reduce);
        # eight ranks -- multimem 1.60 ms with 16-64 CTAs (more CTAs only contend in the switch: 1.70-1.75),
        # ncclAllReduce 1.76, peer-to-peer 2.04.
        if self._world == 2 and self._world in (2, 4, 8):
            self._mc = 0
        self.ctas = 32 if self._mc else 0
        self.algo = ("own kernel: NVLS multimem.ld_reduce + multimem.st over the multicast mapping" if self._mc
                     else "own kernel: peer-to-peer loads / stores over NVLink" if self._world in (2, 4, 8)
                     else "ncclAllReduce on the symmetric buffer (no multicast object; peer-to-peer kernel built for 2 / 4 / 8 ranks)")

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

    # ---- landing the arena in the model's .grad (drop-in autograd path) -------------------------------
    # With `GaussianRasterizer(..., grad_arena=arena)` the backward of every view ADDS its gradients with
    # respect to the rasterizer INPUTS (world-space means, activated opacity / scales / rotations / refl, the
    # concatenated SH block) into the arena and hands autograd nothing for them.  Two ways on from there:

    def bind_grads(self, tensors):
        """Leaf inputs (the caller feeds parameters straight into the rasterizer): make `.grad` of each leaf BE
        its arena block, once.  tensors: {block name: leaf tensor}.  After the batch (and all_reduce()) the
        optimizer reads .grad as usual; zero the batch with arena.zero_() (or zero_grad(set_to_none=False))."""
        for k, t in tensors.items():
            v = self.view(k)
            if tuple(t.shape) != tuple(v.shape) and t.numel() == v.numel():
                v = v.view(t.shape)
            if tuple(t.shape) != tuple(v.shape) or t.device != v.device or t.dtype != v.dtype:
                raise ValueError(f"{k}: tensor {tuple(t.shape)} does not match arena block {tuple(v.shape)}")
            t.grad = v
        return self

    def backward_through(self, inputs):
        """Activated inputs (the reference's model: opacity = sigmoid(_opacity), scales = exp(_scaling),
        rotations = normalize(_rotation), shs = cat(_features_dc, _features_rest), scene/gaussian_model.py:34-52,
        :124-147): the activations are the same for every view of a batch, so the chain rule through them is applied
        ONCE to the batch sum -- after all_reduce(), call this with {block name: activated tensor} (e.g. the
        model's get_xyz / get_features / get_opacity / get_scaling / get_rotation / get_refl, evaluated with
        grad enabled) and the raw parameters receive their .grad through autograd."""
        ts, gs = [], []
        for k, t in inputs.items():
            if t.requires_grad:
                ts.append(t)
                gs.append(self.view(k).view(t.shape))
        if ts:
            torch.autograd.backward(ts, gs)
        return self

    def all_reduce(self, group=None, async_op=False):
        """Sum the arena over all ranks in place.  Symmetric arenas: this library's NVLS / peer-to-peer kernel on
        the current stream, bracketed by the handle's device-side barriers (every rank's backward has finished
        writing before any slice is reduced; every slice has landed before anyone reads).  Otherwise
        ncclAllReduce (gloo in the CPU tests) on the same memory."""
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return None
        if self.symmetric and not self._mc and self._world not in (2, 4, 8):
            # no multicast object and a rank count the peer-to-peer kernel is not built for: NCCL on the same memory
            return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
        if self.symmetric:
            import ctypes as C
            import b200gs
            h = self._hdl
            peers = (C.c_void_p * self._world)(*[int(p) + self._offset for p in h.buffer_ptrs])
            h.barrier(channel=0)
            b200gs.check(b200gs.peer_allreduce_f32(self._mc or None, peers, self._rank, self._world, self._buf.numel(),
                                                   int(getattr(self, "ctas", 0)), torch.cuda.current_stream().cuda_stream))
            h.barrier(channel=1)
            return None
        return dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
