// update.cu -- the per-Gaussian gradient consumer that follows the rendering path
// ("next" row 1): Adam over the model's parameter groups and the
// densification statistics, straight on the gradient blocks the backward wrote (the flat
// arena of python/b200gs/dp.py in view-parallel runs).
//
// What it replaces in the reference:
//   * gaussians.optimizer.step()  (train.py:306) -- torch.optim.Adam(l, lr=0.0, eps=1e-15) over the
//     groups xyz / f_dc / f_rest / opacity / scaling / rotation / refl (scene/gaussian_model.py:196-207),
//     which torch runs as ~10 multi-tensor passes over every parameter (lerp, mul, addcmul, sqrt, div,
//     add, addcdiv); here ONE pass: 4 reads + 3 writes of 4 B per parameter float.
//   * gaussians.max_radii2D[...] = max(...)  (train.py:242-243) and add_densification_stats
//     (scene/gaussian_model.py:579-584): five masked gather/scatter torch ops -> one kernel.
//
// Arithmetic follows torch's Adam (torch/optim/adam.py, _single_tensor_adam / _multi_tensor_adam,
// torch 2.11: the dependency the reference calls, not vendored in it):
//   m = m + (1 - beta1) (g - m);  v = beta2 v + (1 - beta2) g g;
//   p = p - (lr / (1 - beta1^t)) * m / (sqrt(v) / sqrt(1 - beta2^t) + eps)
// with the scalar factors formed on the host in double and rounded to fp32 as torch does.
//
// HBM-bound streaming kernel: 28 algorithmic bytes per parameter float; 59 floats per Gaussian of
// the surfel model -> 1652 B per Gaussian.
#include "common.cuh"
#include "../../include/b200gs.h"

namespace b200gs {

constexpr int UPD_THREADS = 256;
constexpr int UPD_ITEMS = 8;                           // floats per thread
constexpr int UPD_CHUNK = UPD_THREADS * UPD_ITEMS;     // floats per CTA
constexpr int UPD_MAX_GROUPS = B200GS_ADAM_MAX_GROUPS;

struct AdamGroupDev {
  float* p; const float* g; float* m; float* v;
  long long n;             // floats in the parameter tensor
  long long chunk0;        // first CTA of this group
  int p_row, g_row, g_col; // parameter row length, gradient row stride and first column
  float step_size;         // lr / (1 - beta1^t)
  float bc2_sqrt;          // sqrt(1 - beta2^t), t = the group's step
  int vec;                 // p, m, v are 16-B aligned: 128-bit accesses
  int gvec;                // ... and so is the (unstrided) gradient
  unsigned magic;          // ceil(2^32 / p_row): q / p_row == umulhi(q, magic) for q * p_row < 2^32
};
struct AdamArgs {
  AdamGroupDev grp[UPD_MAX_GROUPS];
  int groups;
  float one_minus_beta1, beta2, one_minus_beta2, eps;
};

// Exact zeros (Gaussians no view has reached yet: g = m = v = 0) are stepped around the IEEE division
// and square root, whose slow paths they would otherwise take; the results are the same bits
// (sqrt(0) / c + eps == eps, and p - s * (+-0 / d) == p).
__device__ __forceinline__ void adam_update(const AdamArgs& a, float step_size, float bc2_sqrt, float g, float& p, float& m, float& v) {
  m = __fmaf_rn(a.one_minus_beta1, __fsub_rn(g, m), m);                              // lerp_(grad, 1 - beta1)
  v = __fmaf_rn(__fmul_rn(g, g), a.one_minus_beta2, __fmul_rn(v, a.beta2));          // mul_(beta2).addcmul_(g, g, 1 - beta2)
  if (m == 0.0f && (v > 0.0f || a.eps > 0.0f)) return;
  float denom = a.eps;
  if (v != 0.0f) denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), bc2_sqrt), a.eps);
  p = __fmaf_rn(-step_size, __fdiv_rn(m, denom), p);                                 // addcdiv_(m, denom, -step_size)
}

__global__ void __launch_bounds__(UPD_THREADS) adam_kernel(const __grid_constant__ AdamArgs a) {
  // which group does this CTA belong to (<= 16 entries)
  int gi = 0;
#pragma unroll
  for (int k = 1; k < UPD_MAX_GROUPS; ++k)
    if (k < a.groups && (long long)blockIdx.x >= a.grp[k].chunk0) gi = k;
  const AdamGroupDev& G = a.grp[gi];
  const long long base = ((long long)blockIdx.x - G.chunk0) * UPD_CHUNK;
  const bool strided = (G.g_row != G.p_row) || (G.g_col != 0);
  // gradient index of parameter float base + o, o < UPD_CHUNK: one 64-bit division per thread, 32-bit ones per element
  const long long row0 = strided ? base / G.p_row : 0;
  const unsigned rem0 = strided ? (unsigned)(base - row0 * G.p_row) : 0u;
  const float* gbase = strided ? G.g + row0 * G.g_row + G.g_col : G.g + base;
  auto gidx = [&](unsigned o) -> unsigned {
    if (!strided) return o;
    const unsigned q = rem0 + o, r = __umulhi(q, G.magic);
    return r * (unsigned)G.g_row + (q - r * (unsigned)G.p_row);
  };

  if (G.vec && base + UPD_CHUNK <= G.n) {              // full chunk, 128-bit accesses
    constexpr int V = UPD_ITEMS / 4;
    float4 pv[V], mv[V], vv[V], gv[V];
#pragma unroll
    for (int k = 0; k < V; ++k) {                      // all loads first
      const unsigned o = (k * UPD_THREADS + threadIdx.x) * 4u;
      pv[k] = *reinterpret_cast<const float4*>(G.p + base + o);
      mv[k] = *reinterpret_cast<const float4*>(G.m + base + o);
      vv[k] = *reinterpret_cast<const float4*>(G.v + base + o);
      if (G.gvec) gv[k] = __ldcs(reinterpret_cast<const float4*>(gbase + o));   // gradients are read once
      else gv[k] = make_float4(__ldcs(gbase + gidx(o)), __ldcs(gbase + gidx(o + 1)), __ldcs(gbase + gidx(o + 2)), __ldcs(gbase + gidx(o + 3)));
    }
#pragma unroll
    for (int k = 0; k < V; ++k) {
      const unsigned o = (k * UPD_THREADS + threadIdx.x) * 4u;
      adam_update(a, G.step_size, G.bc2_sqrt, gv[k].x, pv[k].x, mv[k].x, vv[k].x);
      adam_update(a, G.step_size, G.bc2_sqrt, gv[k].y, pv[k].y, mv[k].y, vv[k].y);
      adam_update(a, G.step_size, G.bc2_sqrt, gv[k].z, pv[k].z, mv[k].z, vv[k].z);
      adam_update(a, G.step_size, G.bc2_sqrt, gv[k].w, pv[k].w, mv[k].w, vv[k].w);
      *reinterpret_cast<float4*>(G.p + base + o) = pv[k];
      *reinterpret_cast<float4*>(G.m + base + o) = mv[k];
      *reinterpret_cast<float4*>(G.v + base + o) = vv[k];
    }
    return;
  }

  float pv[UPD_ITEMS], gv[UPD_ITEMS], mv[UPD_ITEMS], vv[UPD_ITEMS];
#pragma unroll
  for (int k = 0; k < UPD_ITEMS; ++k) {                // tail chunk / unaligned tensors: 32-bit accesses
    const unsigned o = k * UPD_THREADS + threadIdx.x;
    if (base + o < G.n) {
      pv[k] = G.p[base + o]; mv[k] = G.m[base + o]; vv[k] = G.v[base + o];
      gv[k] = __ldcs(gbase + gidx(o));
    }
  }
#pragma unroll
  for (int k = 0; k < UPD_ITEMS; ++k) {
    const unsigned o = k * UPD_THREADS + threadIdx.x;
    if (base + o < G.n) {
      adam_update(a, G.step_size, G.bc2_sqrt, gv[k], pv[k], mv[k], vv[k]);
      G.p[base + o] = pv[k]; G.m[base + o] = mv[k]; G.v[base + o] = vv[k];
    }
  }
}

// train.py:242-243 + scene/gaussian_model.py:579-584, one thread per Gaussian.
__global__ void __launch_bounds__(256) densify_stats_kernel(int P, const int* __restrict__ radii, const float* __restrict__ dmeans2D,
                                                            const float* __restrict__ weights, float* max_radii2D,
                                                            float* xyz_gradient_accum, float* denom, float* accum_w, float* denom_w) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= P) return;
  const int r = radii[i];
  if (r > 0) {                                         // visibility_filter
    max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
    const float x = dmeans2D[3 * i], y = dmeans2D[3 * i + 1], z = dmeans2D[3 * i + 2];
    xyz_gradient_accum[i] += sqrtf(x * x + y * y + z * z);   // torch.norm(grad, dim=-1) over all three components
    denom[i] += 1.0f;
  }
  if (weights) {
    const float w = weights[i];
    if (w > 0.0f) { accum_w[i] += w; denom_w[i] += 1.0f; }
  }
}

}  // namespace b200gs

using namespace b200gs;

extern "C" {

int b200gs_adam_step(const b200gs_adam* a, void* stream) {
  if (!a) { set_error("adam_step: null argument"); return B200GS_EINVAL; }
  if (a->groups < 0 || a->groups > UPD_MAX_GROUPS) { set_error("adam_step: %d groups (max %d)", a->groups, UPD_MAX_GROUPS); return B200GS_EINVAL; }
  if (a->step < 1) { set_error("adam_step: step must be >= 1"); return B200GS_EINVAL; }
  if (!(a->beta1 >= 0.0 && a->beta1 < 1.0) || !(a->beta2 >= 0.0 && a->beta2 < 1.0) || !(a->eps >= 0.0)) {
    set_error("adam_step: bad beta / eps"); return B200GS_EINVAL; }
  AdamArgs k;
  k.groups = 0;
  long long chunks = 0;
  for (int i = 0; i < a->groups; ++i) {
    const b200gs_adam_group& g = a->group[i];
    if (g.n == 0) continue;
    if (g.n < 0 || !g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq) { set_error("adam_step: group %d: bad argument", i); return B200GS_EINVAL; }
    const int p_row = g.param_row > 0 ? g.param_row : 1;
    const int g_row = g.grad_row > 0 ? g.grad_row : p_row;
    if (g.grad_col < 0 || g.grad_col + p_row > g_row || g.n % p_row != 0) { set_error("adam_step: group %d: bad row layout", i); return B200GS_EINVAL; }
    const long long step = g.step > 0 ? g.step : a->step;
    if (step < 1) { set_error("adam_step: group %d: step must be >= 1", i); return B200GS_EINVAL; }
    const double bc1 = 1.0 - pow(a->beta1, (double)step);
    const double bc2 = 1.0 - pow(a->beta2, (double)step);
    AdamGroupDev& d = k.grp[k.groups++];
    d.p = g.param; d.g = g.grad; d.m = g.exp_avg; d.v = g.exp_avg_sq;
    d.n = g.n; d.chunk0 = chunks;
    d.p_row = p_row; d.g_row = g_row; d.g_col = g.grad_col;
    d.step_size = (float)(g.lr / bc1);
    d.bc2_sqrt = (float)sqrt(bc2);
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15u) == 0; };
    d.vec = (al16(g.param) && al16(g.exp_avg) && al16(g.exp_avg_sq)) ? 1 : 0;
    d.gvec = (d.vec && g_row == p_row && g.grad_col == 0 && al16(g.grad)) ? 1 : 0;
    d.magic = (unsigned)((0x100000000ull + (unsigned)p_row - 1u) / (unsigned)p_row);
    if (p_row == 1) d.magic = 0xffffffffu;   // unused: a one-float row is never strided below (rows collapse)
    if ((g_row != p_row || g.grad_col != 0) && (p_row < 2 || p_row > 32768)) { set_error("adam_step: group %d: strided gradient rows need 2 <= param_row <= 32768", i); return B200GS_EINVAL; }
    if ((long long)(g_row / p_row + 2) * UPD_CHUNK >= 0x7fffffffll) { set_error("adam_step: group %d: gradient row too wide", i); return B200GS_EINVAL; }
    chunks += (g.n + UPD_CHUNK - 1) / UPD_CHUNK;
  }
  if (chunks == 0) return B200GS_OK;
  if (chunks > 0x7fffffffll) { set_error("adam_step: too many parameters for one launch"); return B200GS_EINVAL; }
  k.one_minus_beta1 = (float)(1.0 - a->beta1);
  k.beta2 = (float)a->beta2;
  k.one_minus_beta2 = (float)(1.0 - a->beta2);
  k.eps = (float)a->eps;
  cudaStream_t s = (cudaStream_t)stream;
  StageScope _t(ST_ADAM, s);
  adam_kernel<<<(unsigned)chunks, UPD_THREADS, 0, s>>>(k);
  B200GS_LAUNCHED("adam_kernel", false, s);
  return B200GS_OK;
}

int b200gs_densify_stats(int P, const int* radii, const float* dL_dmeans2D, const float* gaussian_weights, float* max_radii2D,
                         float* xyz_gradient_accum, float* denom, float* accum_w, float* denom_w, void* stream) {
  if (P == 0) return B200GS_OK;
  if (P < 0 || !radii || !dL_dmeans2D || !max_radii2D || !xyz_gradient_accum || !denom || (gaussian_weights && (!accum_w || !denom_w))) {
    set_error("densify_stats: bad argument"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  StageScope _t(ST_DENSIFY_STATS, s);
  densify_stats_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, radii, dL_dmeans2D, gaussian_weights, max_radii2D, xyz_gradient_accum,
                                                      denom, accum_w, denom_w);
  B200GS_LAUNCHED("densify_stats_kernel", false, s);
  return B200GS_OK;
}

}  // extern "C"
