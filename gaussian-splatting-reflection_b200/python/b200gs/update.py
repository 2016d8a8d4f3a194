"""Per-Gaussian gradient consumer ("next" row 1): what the reference does
right after the rendering path's backward, on the sm_100a kernels of csrc/update.cu.

  GaussianAdam      <- gaussians.optimizer = torch.optim.Adam(l, lr=0.0, eps=1e-15) over the groups
                       xyz / f_dc / f_rest / opacity / scaling / rotation / refl
                       (scene/gaussian_model.py:196-207) and its .step() (train.py:306)
  DensifyStats      <- max_radii2D update (train.py:242-243) + add_densification_stats
                       (scene/gaussian_model.py:579-584)

Same names, argument meaning and state layout (`param_groups` with "name"/"lr", `state[...]` with
exp_avg / exp_avg_sq) as the torch optimizer the reference builds, so `update_learning_rate`
(scene/gaussian_model.py:216-222) works unchanged.  Gradients are read where the backward left
them: separate tensors, or the blocks of a b200gs.dp.GradArena after its all-reduce; f_dc and
f_rest both read the [P, M, 3] dL_dsh block (columns [0,3) and [3,3M)), no split/copy.

There is no CPU path: tensors must live on a CUDA device.
"""
import torch

import b200gs

# group name -> (gradient block it reads, first column, columns) ; None = "the rest of the row"
_SH_SPLIT = {"f_dc": (0, 3), "f_rest": (3, None)}


class GaussianAdam:
    """Adam over named parameter groups, all groups in ONE kernel launch per step."""

    def __init__(self, groups, betas=(0.9, 0.999), eps=1e-15):
        """groups: list of {"params": [tensor], "lr": float, "name": str} as in training_setup()."""
        if len(groups) > b200gs.ADAM_MAX_GROUPS:
            raise ValueError(f"at most {b200gs.ADAM_MAX_GROUPS} groups per launch")
        self.param_groups, self.state = [], {}
        for g in groups:
            ps = list(g["params"])
            if len(ps) != 1:
                raise ValueError("one tensor per group (the reference's Gaussian groups hold one each)")
            p = ps[0]
            if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous()):
                raise ValueError(f"group {g.get('name')}: parameters must be contiguous fp32 CUDA tensors")
            self.param_groups.append({"params": [p], "lr": float(g["lr"]), "name": g["name"]})
            self.state[g["name"]] = {"step": 0, "exp_avg": torch.zeros_like(p), "exp_avg_sq": torch.zeros_like(p)}
        self.betas, self.eps = (float(betas[0]), float(betas[1])), float(eps)
        self.step_count = 0

    def step(self, grads=None):
        """grads: {group name: tensor}; "sh" may stand in for both f_dc and f_rest ([P, M, 3], the
        rasterizer's dL_dsh).  Without `grads` every parameter's .grad is used (like torch's step())."""
        a = b200gs.Adam()
        a.beta1, a.beta2, a.eps = self.betas[0], self.betas[1], self.eps
        self.step_count += 1
        a.step = self.step_count
        k, keep = 0, []
        for g in self.param_groups:
            p, name = g["params"][0], g["name"]
            if p.numel() == 0:
                continue
            row = p[0].numel() if p.dim() > 1 else 1
            grad, grow, gcol = None, row, 0
            if grads is not None:
                if name in grads:
                    grad = grads[name]
                elif name in _SH_SPLIT and "sh" in grads:
                    grad = grads["sh"]
                    grow = grad[0].numel()
                    gcol = _SH_SPLIT[name][0]
                    if gcol + row > grow:
                        raise ValueError(f"{name}: parameter row {row} does not fit the sh gradient row {grow}")
            else:
                grad = p.grad
            if grad is None:
                continue
            if not (grad.is_cuda and grad.dtype == torch.float32 and grad.is_contiguous()):
                raise ValueError(f"group {name}: gradients must be contiguous fp32 CUDA tensors")
            if grad.numel() != (p.numel() // row) * grow:
                raise ValueError(f"group {name}: gradient has {grad.numel()} floats, expected {(p.numel() // row) * grow}")
            st = self.state[name]
            st["step"] = self.step_count
            e = a.group[k]
            e.param, e.grad, e.exp_avg, e.exp_avg_sq = p.data_ptr(), grad.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
            e.n, e.param_row, e.grad_row, e.grad_col, e.lr = p.numel(), row, grow, gcol, g["lr"]
            keep.append(grad)
            k += 1
        a.groups = k
        if k:
            b200gs.check(b200gs.adam_step(a, torch.cuda.current_stream().cuda_stream))

    def zero_grad(self, set_to_none=True):
        for g in self.param_groups:
            p = g["params"][0]
            if p.grad is not None:
                if set_to_none:
                    p.grad = None
                else:
                    p.grad.zero_()


def surfel_param_groups(xyz, f_dc, f_rest, opacity, scaling, rotation, refl, lrs):
    """The seven per-Gaussian groups of scene/gaussian_model.py:196-204, `lrs` = {name: lr}."""
    t = dict(xyz=xyz, f_dc=f_dc, f_rest=f_rest, opacity=opacity, scaling=scaling, rotation=rotation, refl=refl)
    return [{"params": [t[k]], "lr": lrs[k], "name": k} for k in t]


# rasterizer / arena gradient names -> optimizer group names
GRAD_TO_GROUP = {"means3D": "xyz", "sh": "sh", "opacity": "opacity", "scales": "scaling", "rotations": "rotation", "refl": "refl"}


def arena_grads(arena):
    """{group name: gradient block} of a b200gs.dp.GradArena (after arena.all_reduce())."""
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
