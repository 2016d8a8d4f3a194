"""CPU restatement of the step that consumes the rendering path's gradients -- the ORACLE for
csrc/update.cu ("next" row 1).

TEST INFRASTRUCTURE ONLY.  Nothing in the product package imports this file; only tests/ may.

What is restated:
  adam_step()      torch.optim.Adam as the reference builds it: Adam(l, lr=0.0, eps=1e-15) with
                   one lr per named group (scene/gaussian_model.py:196-207), stepped once per
                   iteration (train.py:306).  The algorithm lives in a dependency that is not
                   vendored in the reference: torch (the reference pins no version; this image has
                   torch 2.11.0), torch/optim/adam.py `_single_tensor_adam` with amsgrad=False,
                   weight_decay=0, maximize=False:
                       exp_avg.lerp_(grad, 1 - beta1)
                       exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                       step_size = lr / (1 - beta1**step)
                       denom = (exp_avg_sq.sqrt() / sqrt(1 - beta2**step)).add_(eps)
                       param.addcdiv_(exp_avg, denom, value=-step_size)
  densify_stats()  train.py:242-243 (max_radii2D) + add_densification_stats
                   (scene/gaussian_model.py:579-584).

Pinned: tests/test_update_oracle_cpu.py runs torch.optim.Adam itself (CPU, the reference's exact
constructor call and group layout) and the reference's statistics lines as literal torch, and
compares them with these numpy functions on the same seeded inputs.  Plain numpy fp32, one
rounding per operation (no fma), so agreement with torch and with the CUDA kernel is toleranced.
"""
import numpy as np

f32 = np.float32


def adam_step(param, grad, exp_avg, exp_avg_sq, lr, step, beta1=0.9, beta2=0.999, eps=1e-15):
    """One Adam step on fp32 arrays; returns new (param, exp_avg, exp_avg_sq)."""
    p, g, m, v = (np.asarray(x, dtype=f32) for x in (param, grad, exp_avg, exp_avg_sq))
    m = m + f32(1.0 - beta1) * (g - m)
    v = v * f32(beta2) + f32(1.0 - beta2) * g * g
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    step_size = f32(lr / bc1)
    denom = np.sqrt(v) / f32(bc2 ** 0.5) + f32(eps)
    p = p - step_size * (m / denom)
    return p.astype(f32), m.astype(f32), v.astype(f32)


def split_sh_grad(dL_dsh):
    """[P, M, 3] gradient of the concatenated SH features -> (f_dc [P,1,3], f_rest [P,M-1,3]) gradients:
    get_features = cat(_features_dc, _features_rest) (scene/gaussian_model.py:116-119)."""
    g = np.asarray(dL_dsh, dtype=f32)
    return g[:, :1, :].copy(), g[:, 1:, :].copy()


def densify_stats(stats, viewspace_grad, radii, render_weight=None):
    """stats: dict of fp32 [P] arrays xyz_gradient_accum / denom / accum_w / denom_w / max_radii2D,
    updated in place for one rendered view."""
    vis = np.asarray(radii) > 0
    g = np.asarray(viewspace_grad, dtype=f32)
    stats["max_radii2D"][vis] = np.maximum(stats["max_radii2D"][vis], np.asarray(radii)[vis].astype(f32))
    stats["xyz_gradient_accum"][vis] += np.sqrt((g[vis] * g[vis]).sum(-1, dtype=f32)).astype(f32)
    stats["denom"][vis] += f32(1.0)
    if render_weight is not None:
        w = np.asarray(render_weight, dtype=f32)
        m = w > 0
        stats["accum_w"][m] += w[m]
        stats["denom_w"][m] += f32(1.0)
    return stats
