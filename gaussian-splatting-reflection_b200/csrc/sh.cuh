// sh.cuh -- spherical-harmonics colour evaluation (degree <= 3) and its backward, shared by the
// surfel and the 3DGS paths (the reference has the same computeColorFromSH in both:
// S|G/cuda_rasterizer/forward.cu:20-71, backward.cu:20-142).
#pragma once
#include "common.cuh"

namespace b200gs {

// A Gaussian's [M][3] coefficient block as registers.  With M == 16 (degree-3 storage, the model's
// layout) a row is 192 B and 16-B aligned, so 12 128-bit accesses replace 48 scalar ones: one thread
// per Gaussian means every access of a warp touches 32 different rows, and the LSU / L1 wavefront count,
// not DRAM, bounds the preprocess kernels.
__device__ __forceinline__ bool sh_row_vec(const float* p, int M) {
  return M == 16 && (reinterpret_cast<uintptr_t>(p) & 15u) == 0;
}
__device__ __forceinline__ void load_sh_row(const float* __restrict__ sh, int M, int deg, float v[48]) {
  if (sh_row_vec(sh, M)) {
#pragma unroll
    for (int g = 0; g < 12; ++g) {
      const float4 q = __ldg(reinterpret_cast<const float4*>(sh) + g);
      v[4 * g] = q.x; v[4 * g + 1] = q.y; v[4 * g + 2] = q.z; v[4 * g + 3] = q.w;
    }
  } else {
    const int n = 3 * min(M, (deg + 1) * (deg + 1));
#pragma unroll
    for (int j = 0; j < 48; ++j) v[j] = j < n ? sh[j] : 0.0f;
  }
}
template <bool ACC>
__device__ __forceinline__ void store_sh_row(float* dsh, int M, int deg, const float o[48]) {
  if (sh_row_vec(dsh, M)) {
#pragma unroll
    for (int g = 0; g < 12; ++g) {
      float4 q = make_float4(o[4 * g], o[4 * g + 1], o[4 * g + 2], o[4 * g + 3]);
      if (ACC) {
        const float4 old = reinterpret_cast<const float4*>(dsh)[g];
        q.x += old.x; q.y += old.y; q.z += old.z; q.w += old.w;
      }
      reinterpret_cast<float4*>(dsh)[g] = q;
    }
  } else {
    const int n = 3 * min(M, (deg + 1) * (deg + 1));   // rows of higher degrees are not written (backward.cu:20-139)
#pragma unroll
    for (int j = 0; j < 48; ++j)
      if (j < n) dsh[j] = ACC ? dsh[j] + o[j] : o[j];
  }
}

// SH degree<=3 colour (forward.cu:20-71).  sh is this Gaussian's [M][3] block (global memory or registers).
__device__ __forceinline__ void sh_to_rgb(int deg, const float* __restrict__ sh, float dx, float dy, float dz,
                                          float out[3], unsigned& clamp_bits) {
  // glm: dir / length(dir) is a component-wise division; keep it a division for the same rounding.
  const float len = __fsqrt_rn(__fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy))));
  const float x = __fdiv_rn(dx, len), y = __fdiv_rn(dy, len), z = __fdiv_rn(dz, len);
  // Operation shapes pinned to the reference build's SASS (computeColorFromSH inlined in
  // preprocessCUDA<3>): every basis term is rounded on its own, then one FFMA per coefficient.
  float r[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) r[c] = __fmul_rn(SH_C0, sh[c]);
  if (deg > 0) {
    const float b1 = __fmul_rn(y, SH_C1), b2 = __fmul_rn(z, SH_C1), b3 = __fmul_rn(x, SH_C1);
#pragma unroll
    for (int c = 0; c < 3; ++c)
      r[c] = __fmaf_rn(-b3, sh[9 + c], __fmaf_rn(b2, sh[6 + c], __fmaf_rn(-b1, sh[3 + c], r[c])));
    if (deg > 1) {
      const float xx = __fmul_rn(x, x), yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
      const float xy = __fmul_rn(x, y), yz = __fmul_rn(y, z), xz = __fmul_rn(x, z);
      const float zz2 = __fadd_rn(zz, zz);
      const float xx_yy = __fsub_rn(xx, yy);
      const float b4 = __fmul_rn(xy, SH_C2_0), b5 = __fmul_rn(yz, SH_C2_1);
      const float b6 = __fmul_rn(__fsub_rn(__fsub_rn(zz2, xx), yy), SH_C2_2);
      const float b7 = __fmul_rn(xz, SH_C2_3), b8 = __fmul_rn(xx_yy, SH_C2_4);
#pragma unroll
      for (int c = 0; c < 3; ++c)
        r[c] = __fmaf_rn(b8, sh[24 + c], __fmaf_rn(b7, sh[21 + c], __fmaf_rn(b6, sh[18 + c],
               __fmaf_rn(b5, sh[15 + c], __fmaf_rn(b4, sh[12 + c], r[c])))));
      if (deg > 2) {
        const float a4 = __fsub_rn(__fmaf_rn(zz, 4.0f, -xx), yy);  // 4zz - xx - yy
        const float b9 = __fmul_rn(__fmul_rn(y, SH_C3_0), __fmaf_rn(xx, 3.0f, -yy));
        const float b10 = __fmul_rn(__fmul_rn(xy, SH_C3_1), z);
        const float b11 = __fmul_rn(__fmul_rn(y, SH_C3_2), a4);
        const float b12 = __fmul_rn(__fmul_rn(z, SH_C3_3), __fmaf_rn(yy, -3.0f, __fmaf_rn(xx, -3.0f, zz2)));
        const float b13 = __fmul_rn(a4, __fmul_rn(x, SH_C3_4));
        const float b14 = __fmul_rn(xx_yy, __fmul_rn(z, SH_C3_5));
        const float b15 = __fmul_rn(__fmul_rn(x, SH_C3_6), __fmaf_rn(yy, -3.0f, xx));
#pragma unroll
        for (int c = 0; c < 3; ++c)
          r[c] = __fmaf_rn(b15, sh[45 + c], __fmaf_rn(b14, sh[42 + c], __fmaf_rn(b13, sh[39 + c],
                 __fmaf_rn(b12, sh[36 + c], __fmaf_rn(b11, sh[33 + c], __fmaf_rn(b10, sh[30 + c],
                 __fmaf_rn(b9, sh[27 + c], r[c])))))));
      }
    }
  }
  clamp_bits = 0;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    r[c] = __fadd_rn(r[c], 0.5f);
    if (r[c] < 0.0f) clamp_bits |= 1u << c;
    out[c] = fmaxf(r[c], 0.0f);
  }
}

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
__device__ __forceinline__ V3 operator*(float f, V3 a) { return {f * a.x, f * a.y, f * a.z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// SH backward (backward.cu:20-139): writes dL_dsh rows (ACC: adds to them, for gradients accumulated
// over the views of a batch), returns dL/d(mean) through the view direction.
template <bool ACC = false>
__device__ __forceinline__ V3 sh_backward(int deg, int M, const float* __restrict__ sh_row, V3 dir_orig, V3 dRGB,
                                          float* dsh_row) {
  float sh[48], dsh[48];
  load_sh_row(sh_row, M, deg, sh);
#pragma unroll
  for (int j = 0; j < 48; ++j) dsh[j] = 0.0f;
  const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
  const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
  V3 dx = {0, 0, 0}, dy = {0, 0, 0}, dz = {0, 0, 0};
  auto S = [&](int k) { return V3{sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]}; };
  auto W = [&](int k, float c) { dsh[3 * k] = c * dRGB.x; dsh[3 * k + 1] = c * dRGB.y; dsh[3 * k + 2] = c * dRGB.z; };
  W(0, SH_C0);
  if (deg > 0) {
    W(1, -SH_C1 * y); W(2, SH_C1 * z); W(3, -SH_C1 * x);
    dx = -SH_C1 * S(3); dy = -SH_C1 * S(1); dz = SH_C1 * S(2);
    if (deg > 1) {
      const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
      W(4, SH_C2_0 * xy); W(5, SH_C2_1 * yz); W(6, SH_C2_2 * (2.f * zz - xx - yy));
      W(7, SH_C2_3 * xz); W(8, SH_C2_4 * (xx - yy));
      dx = dx + (SH_C2_0 * y) * S(4) + (SH_C2_2 * 2.f * -x) * S(6) + (SH_C2_3 * z) * S(7) + (SH_C2_4 * 2.f * x) * S(8);
      dy = dy + (SH_C2_0 * x) * S(4) + (SH_C2_1 * z) * S(5) + (SH_C2_2 * 2.f * -y) * S(6) + (SH_C2_4 * 2.f * -y) * S(8);
      dz = dz + (SH_C2_1 * y) * S(5) + (SH_C2_2 * 2.f * 2.f * z) * S(6) + (SH_C2_3 * x) * S(7);
      if (deg > 2) {
        W(9, SH_C3_0 * y * (3.f * xx - yy)); W(10, SH_C3_1 * xy * z); W(11, SH_C3_2 * y * (4.f * zz - xx - yy));
        W(12, SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy)); W(13, SH_C3_4 * x * (4.f * zz - xx - yy));
        W(14, SH_C3_5 * z * (xx - yy)); W(15, SH_C3_6 * x * (xx - 3.f * yy));
        dx = dx + (SH_C3_0 * 3.f * 2.f * xy) * S(9) + (SH_C3_1 * yz) * S(10) + (SH_C3_2 * -2.f * xy) * S(11) +
             (SH_C3_3 * -3.f * 2.f * xz) * S(12) + (SH_C3_4 * (-3.f * xx + 4.f * zz - yy)) * S(13) +
             (SH_C3_5 * 2.f * xz) * S(14) + (SH_C3_6 * 3.f * (xx - yy)) * S(15);
        dy = dy + (SH_C3_0 * 3.f * (xx - yy)) * S(9) + (SH_C3_1 * xz) * S(10) +
             (SH_C3_2 * (-3.f * yy + 4.f * zz - xx)) * S(11) + (SH_C3_3 * -3.f * 2.f * yz) * S(12) +
             (SH_C3_4 * -2.f * xy) * S(13) + (SH_C3_5 * -2.f * yz) * S(14) + (SH_C3_6 * -3.f * 2.f * xy) * S(15);
        dz = dz + (SH_C3_1 * xy) * S(10) + (SH_C3_2 * 4.f * 2.f * yz) * S(11) +
             (SH_C3_3 * 3.f * (2.f * zz - xx - yy)) * S(12) + (SH_C3_4 * 4.f * 2.f * xz) * S(13) +
             (SH_C3_5 * (xx - yy)) * S(14);
      }
    }
  }
  store_sh_row<ACC>(dsh_row, M, deg, dsh);
  const V3 ddir = {dot(dx, dRGB), dot(dy, dRGB), dot(dz, dRGB)};
  // dnormvdv (auxiliary.h:132-142)
  const V3 v = dir_orig;
  const float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
  const float inv32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
  V3 r;
  r.x = ((+sum2 - v.x * v.x) * ddir.x - v.y * v.x * ddir.y - v.z * v.x * ddir.z) * inv32;
  r.y = (-v.x * v.y * ddir.x + (sum2 - v.y * v.y) * ddir.y - v.z * v.y * ddir.z) * inv32;
  r.z = (-v.x * v.z * ddir.x - v.y * v.z * ddir.y + (sum2 - v.z * v.z) * ddir.z) * inv32;
  return r;
}

}  // namespace b200gs
