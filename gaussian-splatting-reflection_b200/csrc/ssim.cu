// ssim.cu -- the image loss kernel of the training step ("next" row 3):
// SSIM with an 11x11 separable Gaussian window (sigma 1.5), zero "same" padding, forward and the
// gradient with respect to the first image.
//
// What it replaces in the reference: fused_ssim(image, gt_image) (train.py:168-169 ->
// submodules/fused-ssim/fused_ssim/__init__.py:8-47 -> fusedssimCUDA ssim.cu:187 / fusedssim_backwardCUDA
// ssim.cu:288), which computes the same map as the torch ssim() of utils/loss_utils.py:63-101.
//
//   mu1 = G*x1, mu2 = G*x2, s1 = G*x1^2 - mu1^2, s2 = G*x2^2 - mu2^2, s12 = G*x1x2 - mu1 mu2
//   map = (2 mu1 mu2 + C1)(2 s12 + C2) / ((mu1^2 + mu2^2 + C1)(s1 + s2 + C2))
//   backward:  dL/dx1 = G*(dL dm/dmu1) + 2 x1 G*(dL dm/ds1) + x2 G*(dL dm/ds12)      (G symmetric)
//
// Design: one CTA per 32x32 tile of one (batch, channel) plane.  Both image tiles are staged ONCE with
// their 5-pixel halo (zero outside the image); the horizontal pass produces all five filtered
// quantities of a row in one sweep, the vertical pass slides an 11-tap register window down a column
// (4 outputs per thread from 14 loads per quantity).  The reference runs five separate
// shared-memory convolutions with ~25 block barriers per channel; here it is 3 barriers.
// FP32-issue / shared-memory bound on the filter taps, ~24 algorithmic bytes per pixel and channel.
#include "common.cuh"

namespace b200gs {

constexpr int SS_TW = 32, SS_TH = 32;              // output tile
constexpr int SS_R = 5;                            // filter radius
constexpr int SS_SW = SS_TW + 2 * SS_R, SS_SH = SS_TH + 2 * SS_R;
constexpr int SS_THREADS = 256;                    // 32 x 8, four output rows per thread

// gaussian(11, 1.5) of utils/loss_utils.py:47-49, fp32
__constant__ float c_gauss[11] = {0.001028380123898387f, 0.0075987582094967365f, 0.036000773310661316f, 0.10936068743467331f,
                                  0.21300552785396576f,  0.26601171493530273f,   0.21300552785396576f,  0.10936068743467331f,
                                  0.036000773310661316f, 0.0075987582094967365f, 0.001028380123898387f};

template <int Q>
struct SsimSmem {
  float in[3][SS_SH][SS_SW + 1];        // staged planes (forward uses 2, backward 3)
  float h[Q][SS_SH][SS_TW + 1];         // horizontally filtered quantities (row stride 33: conflict-free stores)
};

__device__ __forceinline__ float plane_at(const float* __restrict__ p, int y, int x, int H, int W) {
  return (x >= 0 && x < W && y >= 0 && y < H) ? __ldg(p + (size_t)y * W + x) : 0.0f;
}

// Vertical 11-tap pass for the four rows 4*ty .. 4*ty+3 of column tx: out[k] = sum_i g[i] h[4 ty + k + i][tx].
__device__ __forceinline__ void vertical4(const float (*h)[SS_TW + 1], int tx, int ty, float out[4]) {
  float w[14];
#pragma unroll
  for (int k = 0; k < 14; ++k) w[k] = h[4 * ty + k][tx];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < 11; ++i) s = fmaf(c_gauss[i], w[k + i], s);
    out[k] = s;
  }
}

__global__ void __launch_bounds__(SS_THREADS) ssim_fwd_kernel(int H, int W, float C1, float C2, const float* __restrict__ img1,
                                                              const float* __restrict__ img2, float* __restrict__ ssim_map,
                                                              float* __restrict__ dm_dmu1, float* __restrict__ dm_dsigma1_sq,
                                                              float* __restrict__ dm_dsigma12) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SsimSmem<5>& sm = *reinterpret_cast<SsimSmem<5>*>(smem_raw);
  const int t = threadIdx.x, tx = t & 31, ty = t >> 5;
  const int x0 = blockIdx.x * SS_TW, y0 = blockIdx.y * SS_TH;
  const size_t plane = (size_t)blockIdx.z * H * W;
  const float* p1 = img1 + plane;
  const float* p2 = img2 + plane;

  {   // stage both tiles: all of a thread's loads are issued before the first store waits on one
    constexpr int IT = (SS_SH * SS_SW + SS_THREADS - 1) / SS_THREADS;
    float r1[IT], r2[IT];
#pragma unroll
    for (int k = 0; k < IT; ++k) {
      const int e = t + k * SS_THREADS, ly = e / SS_SW, lx = e - ly * SS_SW;
      r1[k] = plane_at(p1, y0 + ly - SS_R, x0 + lx - SS_R, H, W);
      r2[k] = plane_at(p2, y0 + ly - SS_R, x0 + lx - SS_R, H, W);
    }
#pragma unroll
    for (int k = 0; k < IT; ++k) {
      const int e = t + k * SS_THREADS, ly = e / SS_SW, lx = e - ly * SS_SW;
      if (e < SS_SH * SS_SW) { sm.in[0][ly][lx] = r1[k]; sm.in[1][ly][lx] = r2[k]; }
    }
  }
  __syncthreads();
  // horizontal pass: four consecutive outputs of a staged row per work item, all five quantities from one
  // 14-wide register window of each image
  for (int e = t; e < SS_SH * (SS_TW / 4); e += SS_THREADS) {
    const int ly = e >> 3, lx = (e & 7) * 4;
    float u[14], v[14];
#pragma unroll
    for (int i = 0; i < 14; ++i) { u[i] = sm.in[0][ly][lx + i]; v[i] = sm.in[1][ly][lx + i]; }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float a = 0.f, b = 0.f, aa = 0.f, bb = 0.f, ab = 0.f;
#pragma unroll
      for (int i = 0; i < 11; ++i) {
        const float g = c_gauss[i];
        const float gu = g * u[k + i], gv = g * v[k + i];
        a += gu; b += gv;
        aa = fmaf(gu, u[k + i], aa); bb = fmaf(gv, v[k + i], bb); ab = fmaf(gu, v[k + i], ab);
      }
      sm.h[0][ly][lx + k] = a; sm.h[1][ly][lx + k] = b; sm.h[2][ly][lx + k] = aa; sm.h[3][ly][lx + k] = bb; sm.h[4][ly][lx + k] = ab;
    }
  }
  __syncthreads();
  float mu1[4], mu2[4], e11[4], e22[4], e12[4];
  vertical4(sm.h[0], tx, ty, mu1);
  vertical4(sm.h[1], tx, ty, mu2);
  vertical4(sm.h[2], tx, ty, e11);
  vertical4(sm.h[3], tx, ty, e22);
  vertical4(sm.h[4], tx, ty, e12);
  const int x = x0 + tx;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int y = y0 + 4 * ty + k;
    if (x >= W || y >= H) continue;
    const float m1 = mu1[k], m2 = mu2[k];
    const float mu1_sq = m1 * m1, mu2_sq = m2 * m2, mu1_mu2 = m1 * m2;
    const float sigma1_sq = e11[k] - mu1_sq, sigma2_sq = e22[k] - mu2_sq, sigma12 = e12[k] - mu1_mu2;
    const float Cc = 2.0f * mu1_mu2 + C1, Dd = 2.0f * sigma12 + C2;
    const float Aa = mu1_sq + mu2_sq + C1, Bb = sigma1_sq + sigma2_sq + C2;
    const float inv_ab = 1.0f / (Aa * Bb);
    const size_t o = plane + (size_t)y * W + x;
    ssim_map[o] = Cc * Dd * inv_ab;
    if (dm_dmu1) {   // the three partial derivatives the backward filters (ssim.cu:275-283 of the reference)
      dm_dmu1[o] = (m2 * 2.0f * Dd) * inv_ab - (m2 * 2.0f * Cc) * inv_ab - (m1 * 2.0f * Cc * Dd) * inv_ab / Aa +
                   (m1 * 2.0f * Cc * Dd) * inv_ab / Bb;
      dm_dsigma1_sq[o] = (-Cc * Dd) * inv_ab / Bb;
      dm_dsigma12[o] = (2.0f * Cc) * inv_ab;
    }
  }
}

__global__ void __launch_bounds__(SS_THREADS) ssim_bwd_kernel(int H, int W, const float* __restrict__ img1,
                                                              const float* __restrict__ img2, const float* __restrict__ dL_dmap,
                                                              const float* __restrict__ dm_dmu1,
                                                              const float* __restrict__ dm_dsigma1_sq,
                                                              const float* __restrict__ dm_dsigma12, float* __restrict__ dL_dimg1) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SsimSmem<3>& sm = *reinterpret_cast<SsimSmem<3>*>(smem_raw);
  const int t = threadIdx.x, tx = t & 31, ty = t >> 5;
  const int x0 = blockIdx.x * SS_TW, y0 = blockIdx.y * SS_TH;
  const size_t plane = (size_t)blockIdx.z * H * W;
  const float* pl = dL_dmap + plane;
  const float* pa = dm_dmu1 + plane;
  const float* pb = dm_dsigma1_sq + plane;
  const float* pc = dm_dsigma12 + plane;

  {
    constexpr int IT = (SS_SH * SS_SW + SS_THREADS - 1) / SS_THREADS;
    float rl[IT], ra[IT], rb[IT], rc[IT];
#pragma unroll
    for (int k = 0; k < IT; ++k) {
      const int e = t + k * SS_THREADS, ly = e / SS_SW, lx = e - ly * SS_SW;
      const int y = y0 + ly - SS_R, x = x0 + lx - SS_R;
      rl[k] = plane_at(pl, y, x, H, W); ra[k] = plane_at(pa, y, x, H, W);
      rb[k] = plane_at(pb, y, x, H, W); rc[k] = plane_at(pc, y, x, H, W);
    }
#pragma unroll
    for (int k = 0; k < IT; ++k) {
      const int e = t + k * SS_THREADS, ly = e / SS_SW, lx = e - ly * SS_SW;
      if (e < SS_SH * SS_SW) { sm.in[0][ly][lx] = rl[k] * ra[k]; sm.in[1][ly][lx] = rl[k] * rb[k]; sm.in[2][ly][lx] = rl[k] * rc[k]; }
    }
  }
  __syncthreads();
  for (int e = t; e < SS_SH * (SS_TW / 4); e += SS_THREADS) {
    const int ly = e >> 3, lx = (e & 7) * 4;
    float u[14], v[14], w[14];
#pragma unroll
    for (int i = 0; i < 14; ++i) { u[i] = sm.in[0][ly][lx + i]; v[i] = sm.in[1][ly][lx + i]; w[i] = sm.in[2][ly][lx + i]; }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float a = 0.f, b = 0.f, c = 0.f;
#pragma unroll
      for (int i = 0; i < 11; ++i) {
        const float g = c_gauss[i];
        a = fmaf(g, u[k + i], a); b = fmaf(g, v[k + i], b); c = fmaf(g, w[k + i], c);
      }
      sm.h[0][ly][lx + k] = a; sm.h[1][ly][lx + k] = b; sm.h[2][ly][lx + k] = c;
    }
  }
  __syncthreads();
  float ga[4], gb[4], gc[4];
  vertical4(sm.h[0], tx, ty, ga);
  vertical4(sm.h[1], tx, ty, gb);
  vertical4(sm.h[2], tx, ty, gc);
  const int x = x0 + tx;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int y = y0 + 4 * ty + k;
    if (x >= W || y >= H) continue;
    const size_t o = plane + (size_t)y * W + x;
    dL_dimg1[o] = ga[k] + img1[o] * 2.0f * gb[k] + img2[o] * gc[k];
  }
}

}  // namespace b200gs

using namespace b200gs;

extern "C" {

int b200gs_ssim_forward(float C1, float C2, int B, int CH, int H, int W, const float* img1, const float* img2, float* ssim_map,
                        float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12, void* stream) {
  if (B < 0 || CH < 0 || H < 0 || W < 0) { set_error("ssim_forward: negative size"); return B200GS_EINVAL; }
  if ((long long)B * CH == 0 || H == 0 || W == 0) return B200GS_OK;
  if (!img1 || !img2 || !ssim_map) { set_error("ssim_forward: null image"); return B200GS_EINVAL; }
  const bool train = dm_dmu1 || dm_dsigma1_sq || dm_dsigma12;
  if (train && !(dm_dmu1 && dm_dsigma1_sq && dm_dsigma12)) { set_error("ssim_forward: all three derivative maps or none"); return B200GS_EINVAL; }
  if ((long long)B * CH > 65535) { set_error("ssim_forward: more than 65535 planes"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  static bool attr = false;
  if (!attr) {
    B200GS_CUDA(cudaFuncSetAttribute(ssim_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SsimSmem<5>)));
    B200GS_CUDA(cudaFuncSetAttribute(ssim_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SsimSmem<3>)));
    attr = true;
  }
  StageScope _t(ST_SSIM_FWD, s);
  const dim3 grid((W + SS_TW - 1) / SS_TW, (H + SS_TH - 1) / SS_TH, B * CH);
  ssim_fwd_kernel<<<grid, SS_THREADS, sizeof(SsimSmem<5>), s>>>(H, W, C1, C2, img1, img2, ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12);
  B200GS_LAUNCHED("ssim_fwd_kernel", false, s);
  return B200GS_OK;
}

int b200gs_ssim_backward(int B, int CH, int H, int W, const float* img1, const float* img2, const float* dL_dmap,
                         const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12, float* dL_dimg1,
                         void* stream) {
  if (B < 0 || CH < 0 || H < 0 || W < 0) { set_error("ssim_backward: negative size"); return B200GS_EINVAL; }
  if ((long long)B * CH == 0 || H == 0 || W == 0) return B200GS_OK;
  if (!img1 || !img2 || !dL_dmap || !dm_dmu1 || !dm_dsigma1_sq || !dm_dsigma12 || !dL_dimg1) {
    set_error("ssim_backward: null argument"); return B200GS_EINVAL; }
  if ((long long)B * CH > 65535) { set_error("ssim_backward: more than 65535 planes"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  static bool attr = false;
  if (!attr) {
    B200GS_CUDA(cudaFuncSetAttribute(ssim_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SsimSmem<3>)));
    attr = true;
  }
  StageScope _t(ST_SSIM_BWD, s);
  const dim3 grid((W + SS_TW - 1) / SS_TW, (H + SS_TH - 1) / SS_TH, B * CH);
  ssim_bwd_kernel<<<grid, SS_THREADS, sizeof(SsimSmem<3>), s>>>(H, W, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, dL_dimg1);
  B200GS_LAUNCHED("ssim_bwd_kernel", false, s);
  return B200GS_OK;
}

}  // extern "C"
