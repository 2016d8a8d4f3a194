"""Per-Gaussian gradient consumer ("next" row 1): what the reference does
right after the rendering path's backward, on the sm_100a kernels of csrc/update.cu.

  GaussianAdam      <- gaussians.optimizer = torch.optim.Adam(l, lr=0.0, eps=1e-15) over the groups
                       xyz / f_dc / f_rest / opacity / scaling / rotation / refl
                       (scene/gaussian_model.py:196-207) and its .step() (train.py:306)
  DensifyStats      <- max_radii2D update (train.py:242-243) + add_densification_stats
                       (scene/gaussian_model.py:579-584)

GaussianAdam IS a torch.optim.Optimizer (same constructor call, `param_groups` with "name"/"lr", `state`
keyed by the parameter tensor with step / exp_avg / exp_avg_sq, state_dict / load_state_dict), so
`update_learning_rate` (scene/gaussian_model.py:216-222), the densify / prune / reset-opacity state surgery
(:383-478) and capture / restore (:55-94) work on it unchanged.  Gradients are read from .grad, or -- for
callers that feed the parameters straight into the rasterizer -- from the blocks of a b200gs.dp.GradArena
after its all-reduce; f_dc and f_rest then both read the [P, M, 3] dL_dsh block (columns [0,3) and [3,3M)),
no split/copy.

There is no CPU path: tensors must live on a CUDA device.
"""
import torch

import b200gs

# group name -> (gradient block it reads, first column, columns) ; None = "the rest of the row"
_SH_SPLIT = {"f_dc": (0, 3), "f_rest": (3, None)}


class GaussianAdam(torch.optim.Optimizer):
    """Adam with torch.optim.Adam's surface -- `param_groups`, `state` KEYED BY THE PARAMETER TENSOR with
    `step` / `exp_avg` / `exp_avg_sq` entries, `state_dict()` / `load_state_dict()`, `zero_grad()` -- so the
    reference's optimizer surgery (replace_tensor_to_optimizer, _prune_optimizer, cat_tensors_to_optimizer,
    construct_list_of_attributes, capture / restore; scene/gaussian_model.py:55-94, :383-478) works on it
    unchanged, and with ONE kernel launch per step() over all parameter tensors instead of torch's ~10
    multi-tensor passes.  Construct it exactly like the reference does:
    `GaussianAdam(l, lr=0.0, eps=1e-15)` (scene/gaussian_model.py:207); groups may hold several tensors (the
    reference's "env" group holds the cube-map texture and its fail value).

    State entries are created lazily at a parameter's first step, like torch's; `step` is a CPU fp32 scalar
    tensor like torch's default (non-capturable) Adam, so state dicts are interchangeable with
    torch.optim.Adam's.  A parameter whose gradient is None is skipped and its step does not advance."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        if not 0.0 <= eps:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not (0.0 <= betas[0] < 1.0 and 0.0 <= betas[1] < 1.0):
            raise ValueError(f"Invalid beta parameters: {betas}")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    @staticmethod
    def _check(p, what, name):
        if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous()):
            raise ValueError(f"group {name}: {what} must be a contiguous fp32 CUDA tensor")

    def _state_of(self, p):
        st = self.state[p]
        if len(st) == 0:
            st["step"] = torch.tensor(0.0, dtype=torch.float32)
            st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    @torch.no_grad()
    def step(self, grads=None, closure=None):
        """grads: optional {group name: tensor} read INSTEAD of .grad -- the blocks of an all-reduced
        b200gs.dp.GradArena; "sh" may stand in for both f_dc and f_rest ([P, M, 3], the rasterizer's dL_dsh,
        read as column slices, no split / copy).  Only meaningful when the rasterizer inputs ARE the
        parameters; a model with activations in between (the reference's) propagates the arena with
        GradArena.backward_through() and then calls step() with no argument, which reads every .grad."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        batches = {}          # (beta1, beta2, eps) -> list of kernel group entries
        keep = []
        for g in self.param_groups:
            name = g.get("name")
            for p in g["params"]:
                if p.numel() == 0:
                    continue
                row = p[0].numel() if p.dim() > 1 else 1
                grad, grow, gcol = None, row, 0
                if grads is not None and name in grads:
                    grad = grads[name]
                elif grads is not None and name in _SH_SPLIT and "sh" in grads:
                    grad = grads["sh"]
                    grow = grad[0].numel()
                    gcol = _SH_SPLIT[name][0]
                    if gcol + row > grow:
                        raise ValueError(f"{name}: parameter row {row} does not fit the sh gradient row {grow}")
                elif grads is None:
                    grad = p.grad
                if grad is None:
                    continue
                if grad.is_sparse:
                    raise RuntimeError("GaussianAdam does not support sparse gradients")
                self._check(p, "parameters", name)
                self._check(grad, "gradients", name)
                if grad.numel() != (p.numel() // row) * grow:
                    raise ValueError(f"group {name}: gradient has {grad.numel()} floats, expected {(p.numel() // row) * grow}")
                st = self._state_of(p)
                m, v = st["exp_avg"], st["exp_avg_sq"]
                if m.shape != p.shape or v.shape != p.shape:
                    # the reference's densify / prune surgery resizes parameter and moments together; a mismatch
                    # here would make the kernel run over the moments' old extent
                    raise RuntimeError(f"group {name}: optimizer state {tuple(m.shape)} / {tuple(v.shape)} does not "
                                       f"match the parameter {tuple(p.shape)}")
                self._check(m, "exp_avg", name)
                self._check(v, "exp_avg_sq", name)
                st["step"] += 1
                e = b200gs.AdamGroup()
                e.param, e.grad, e.exp_avg, e.exp_avg_sq = p.data_ptr(), grad.data_ptr(), m.data_ptr(), v.data_ptr()
                e.n, e.param_row, e.grad_row, e.grad_col = p.numel(), row, grow, gcol
                e.lr, e.step = float(g["lr"]), int(st["step"])
                batches.setdefault((float(g["betas"][0]), float(g["betas"][1]), float(g["eps"])), []).append(e)
                keep.append(grad)
        stream = torch.cuda.current_stream().cuda_stream if batches else None
        for (b1, b2, eps), entries in batches.items():
            for i in range(0, len(entries), b200gs.ADAM_MAX_GROUPS):
                a = b200gs.Adam()
                a.beta1, a.beta2, a.eps, a.step = b1, b2, eps, 1
                chunk = entries[i:i + b200gs.ADAM_MAX_GROUPS]
                for k, e in enumerate(chunk):
                    a.group[k] = e
                a.groups = len(chunk)
                b200gs.check(b200gs.adam_step(a, stream))
        return loss


def surfel_param_groups(xyz, f_dc, f_rest, opacity, scaling, rotation, refl, lrs):
    """The seven per-Gaussian groups of scene/gaussian_model.py:196-204, `lrs` = {name: lr}."""
    t = dict(xyz=xyz, f_dc=f_dc, f_rest=f_rest, opacity=opacity, scaling=scaling, rotation=rotation, refl=refl)
    return [{"params": [t[k]], "lr": lrs[k], "name": k} for k in t]


# rasterizer / arena gradient names -> optimizer group names
GRAD_TO_GROUP = {"means3D": "xyz", "sh": "sh", "opacity": "opacity", "scales": "scaling", "rotations": "rotation", "refl": "refl"}


def arena_grads(arena):
    """{group name: gradient block} of a b200gs.dp.GradArena (after arena.all_reduce()) for GaussianAdam.step().
    The blocks are gradients with respect to the rasterizer's INPUTS: valid as parameter gradients only when
    the parameters are fed to the rasterizer as they are.  The reference's model puts sigmoid / exp / normalize
    between them (scene/gaussian_model.py:124-147): there, use arena.backward_through({...activated tensors...})
    and a plain optimizer.step()."""
    return {GRAD_TO_GROUP[k]: v for k, v in arena.views().items()}


class DensifyStats:
    """xyz_gradient_accum / denom / accum_w / denom_w / max_radii2D of the reference model
    (scene/gaussian_model.py:190-194, :183), held as the five rows of one [5, P] fp32 buffer so a
    view-parallel run combines them with two collectives (rows 0-3: sum, row 4: max)."""

    ROWS = ("xyz_gradient_accum", "denom", "accum_w", "denom_w", "max_radii2D")

    def __init__(self, P, device="cuda"):
        self.P = int(P)
        self.flat = torch.zeros((5, self.P), dtype=torch.float32, device=device)

    def __getattr__(self, name):
        if name in DensifyStats.ROWS:
            return self.flat[DensifyStats.ROWS.index(name)]
        raise AttributeError(name)

    def add(self, viewspace_grad, radii, render_weight=None):
        """One rendered view: viewspace_grad [P,3] (dL_dmeans2D), radii [P] int32, render_weight [P]
        (the rasterizer's gaussian_weights) or None."""
        if not (viewspace_grad.is_cuda and viewspace_grad.is_contiguous() and viewspace_grad.dtype == torch.float32
                and tuple(viewspace_grad.shape) == (self.P, 3)):
            raise ValueError("viewspace_grad must be a contiguous fp32 CUDA tensor [P, 3]")
        if not (radii.is_cuda and radii.dtype == torch.int32 and radii.is_contiguous() and radii.numel() == self.P):
            raise ValueError("radii must be a contiguous int32 CUDA tensor [P]")
        if render_weight is not None and not (render_weight.is_cuda and render_weight.dtype == torch.float32
                                              and render_weight.is_contiguous() and render_weight.numel() == self.P):
            raise ValueError("render_weight must be a contiguous fp32 CUDA tensor [P]")
        f = self.flat
        b200gs.check(b200gs.densify_stats(self.P, radii.data_ptr(), viewspace_grad.data_ptr(), b200gs.ptr(render_weight),
                                          f[4].data_ptr(), f[0].data_ptr(), f[1].data_ptr(), f[2].data_ptr(), f[3].data_ptr(),
                                          torch.cuda.current_stream().cuda_stream))

    def all_reduce(self, group=None):
        """Combine the per-rank statistics of a view-parallel run (sum, sum, sum, sum, max)."""
        return all_reduce_stats(self.flat, group)


def all_reduce_stats(flat, group=None):
    """flat [5, P]: rows 0-3 summed, row 4 maximised over the ranks (NCCL on GPUs, gloo in the CPU tests)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return flat
    dist.all_reduce(flat[:4], op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(flat[4], op=dist.ReduceOp.MAX, group=group)
    return flat
