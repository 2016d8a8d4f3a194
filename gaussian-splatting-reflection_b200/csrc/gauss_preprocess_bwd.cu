// gauss_preprocess_bwd.cu -- K5g: per-Gaussian chain rule of the 3DGS (EWA) path, one kernel.
//
// Restates computeCov2DCUDA (G/cuda_rasterizer/backward.cu:147-326), BACKWARD::preprocessCUDA
// (:398-449), computeCov3D backward (:330-393) and the SH backward (:23-142) -- the reference runs
// the first as its own kernel and re-reads the intermediates; here one thread carries a Gaussian
// through all of it.  Input is the packed [P,16] accumulator written by the compositing backward.
// Reference behaviours kept: antialias branch rescales dL_dopacity and adds the h-derivative
// (:210-246); denom2inv = 1/(denom^2 + 1e-7) (:250); clamp masks x/y_grad_mul (:182-183); invdepth
// term on dL_dtz (:314-315); dL_dmeans = cov-path value, then += projection term, then += SH term
// (:325,:440,:141); gradient w.r.t. the raw (un-normalised) quaternion (:385-392); Python receives
// the "pixels" variant of dL_dmean2D while the full one feeds dL_dmeans3D (:434-436).
#include "common.cuh"
#include "sh.cuh"
#include "gauss_math.cuh"

namespace b200gs {

struct GPreBwdArgs {
  int P, D, M;
  float mod, tan_fovx, tan_fovy, focal_x, focal_y;
  int antialiasing;
  const float* means3D;
  const int* radii;
  const float* shs;
  const uint8_t* clamped;
  const float* opacities;
  const float* scales;
  const float* rotations;
  const float* cov3D;       // precomputed or the forward's
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  const float* acc;
  int has_invdepth;
  float* dL_dmeans2D;   // [P,3] pixels variant
  float* dL_dconic;     // [P,4] or NULL
  float* dL_dopacity;
  float* dL_dcolors;
  float* dL_dnormals;
  float* dL_drefl;
  float* dL_dinvdepths;
  float* dL_dmeans3D;
  float* dL_dcov3D;
  float* dL_dsh;
  float* dL_dscales;
  float* dL_drot;
};

__device__ __forceinline__ float sqf(float x) { return x * x; }

__global__ void __launch_bounds__(256) gauss_preprocess_bwd_kernel(const GPreBwdArgs a) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.P || !(a.radii[idx] > 0)) return;

  float g[GACC_STRIDE];
  {
    const float4* ap = reinterpret_cast<const float4*>(a.acc + (size_t)idx * GACC_STRIDE);
#pragma unroll
    for (int k = 0; k < GACC_STRIDE / 4; ++k) {
      const float4 v = ap[k];
      g[4 * k] = v.x; g[4 * k + 1] = v.y; g[4 * k + 2] = v.z; g[4 * k + 3] = v.w;
    }
  }
  // A visible Gaussian no pixel blended (occluded, or below the 1/255 cut everywhere) has an all-zero accumulator
  // row: every gradient derived from it below is an exact zero (finite factors times 0), which the caller's
  // zero-filled outputs (G/rasterize_points.cu:186-199 does torch::zeros) already hold -- as in the surfel kernel.
  {
    bool any = false;
#pragma unroll
    for (int k = 0; k < GACC_STRIDE; ++k) any = any || (g[k] != 0.0f);
    if (!any) return;
  }
  a.dL_dmeans2D[3 * idx] = g[2]; a.dL_dmeans2D[3 * idx + 1] = g[3];
  if (a.dL_dconic) reinterpret_cast<float4*>(a.dL_dconic)[idx] = make_float4(g[4], g[5], 0.f, g[6]);
  a.dL_dcolors[3 * idx] = g[8]; a.dL_dcolors[3 * idx + 1] = g[9]; a.dL_dcolors[3 * idx + 2] = g[10];
  a.dL_dnormals[3 * idx] = g[11]; a.dL_dnormals[3 * idx + 1] = g[12]; a.dL_dnormals[3 * idx + 2] = g[13];
  if (a.dL_drefl) a.dL_drefl[idx] = g[14];
  if (a.dL_dinvdepths) a.dL_dinvdepths[idx] = g[15];
  float dL_dopacity = g[7];

  // ---- computeCov2DCUDA (:147-326)
  const float* vm = a.viewmatrix;
  const float* cov3D = a.cov3D + 6 * (size_t)idx;
  const float3 mean = make_float3(a.means3D[3 * idx], a.means3D[3 * idx + 1], a.means3D[3 * idx + 2]);
  const float3 dL_dconic = make_float3(g[4], g[5], g[6]);
  float3 t;
  t.x = vm[0] * mean.x + vm[4] * mean.y + vm[8] * mean.z + vm[12];
  t.y = vm[1] * mean.x + vm[5] * mean.y + vm[9] * mean.z + vm[13];
  t.z = vm[2] * mean.x + vm[6] * mean.y + vm[10] * mean.z + vm[14];
  const float limx = 1.3f * a.tan_fovx, limy = 1.3f * a.tan_fovy;
  const float txtz = t.x / t.z, tytz = t.y / t.z;
  t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
  t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
  const float x_grad_mul = txtz < -limx || txtz > limx ? 0.f : 1.f;
  const float y_grad_mul = tytz < -limy || tytz > limy ? 0.f : 1.f;
  const float h_x = a.focal_x, h_y = a.focal_y;
  M3 J = {{{h_x / t.z, 0.0f, -(h_x * t.x) / (t.z * t.z)}, {0.0f, h_y / t.z, -(h_y * t.y) / (t.z * t.z)}, {0.f, 0.f, 0.f}}};
  M3 W = {{{vm[0], vm[4], vm[8]}, {vm[1], vm[5], vm[9]}, {vm[2], vm[6], vm[10]}}};
  M3 Vrk = {{{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}}};
  const M3 T = mul(W, J);
  const M3 cov2D = mul(mul(transpose(T), transpose(Vrk)), T);
  float c_xx = cov2D.c[0][0], c_xy = cov2D.c[0][1], c_yy = cov2D.c[1][1];
  constexpr float h_var = 0.3f;
  float d_inside_root = 0.f;
  if (a.antialiasing) {
    const float det_cov = c_xx * c_yy - c_xy * c_xy;
    c_xx += h_var; c_yy += h_var;
    const float det_cov_plus_h_cov = c_xx * c_yy - c_xy * c_xy;
    const float h_convolution_scaling = sqrtf(fmaxf(0.000025f, det_cov / det_cov_plus_h_cov));
    const float d_h = dL_dopacity * a.opacities[idx];
    dL_dopacity = dL_dopacity * h_convolution_scaling;
    d_inside_root = (det_cov / det_cov_plus_h_cov) <= 0.000025f ? 0.f : d_h / (2 * h_convolution_scaling);
  } else {
    c_xx += h_var; c_yy += h_var;
  }
  a.dL_dopacity[idx] = dL_dopacity;
  float dL_dc_xx = 0, dL_dc_xy = 0, dL_dc_yy = 0;
  if (a.antialiasing) {
    const float x = c_xx, y = c_yy, z = c_xy, w = h_var;
    const float denom_f = d_inside_root / sqf(w * w + w * (x + y) + x * y - z * z);
    dL_dc_xx = w * (w * y + y * y + z * z) * denom_f;
    dL_dc_yy = w * (w * x + x * x + z * z) * denom_f;
    dL_dc_xy = -2.f * w * z * (w + x + y) * denom_f;
  }
  const float denom = c_xx * c_yy - c_xy * c_xy;
  const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
  float dcov[6];
  if (denom2inv != 0) {
    dL_dc_xx += denom2inv * (-c_yy * c_yy * dL_dconic.x + 2 * c_xy * c_yy * dL_dconic.y + (denom - c_xx * c_yy) * dL_dconic.z);
    dL_dc_yy += denom2inv * (-c_xx * c_xx * dL_dconic.z + 2 * c_xx * c_xy * dL_dconic.y + (denom - c_xx * c_yy) * dL_dconic.x);
    dL_dc_xy += denom2inv * 2 * (c_xy * c_yy * dL_dconic.x - (denom + 2 * c_xy * c_xy) * dL_dconic.y + c_xx * c_xy * dL_dconic.z);
    const float (*Tc)[3] = T.c;
    dcov[0] = (Tc[0][0] * Tc[0][0] * dL_dc_xx + Tc[0][0] * Tc[1][0] * dL_dc_xy + Tc[1][0] * Tc[1][0] * dL_dc_yy);
    dcov[3] = (Tc[0][1] * Tc[0][1] * dL_dc_xx + Tc[0][1] * Tc[1][1] * dL_dc_xy + Tc[1][1] * Tc[1][1] * dL_dc_yy);
    dcov[5] = (Tc[0][2] * Tc[0][2] * dL_dc_xx + Tc[0][2] * Tc[1][2] * dL_dc_xy + Tc[1][2] * Tc[1][2] * dL_dc_yy);
    dcov[1] = 2 * Tc[0][0] * Tc[0][1] * dL_dc_xx + (Tc[0][0] * Tc[1][1] + Tc[0][1] * Tc[1][0]) * dL_dc_xy + 2 * Tc[1][0] * Tc[1][1] * dL_dc_yy;
    dcov[2] = 2 * Tc[0][0] * Tc[0][2] * dL_dc_xx + (Tc[0][0] * Tc[1][2] + Tc[0][2] * Tc[1][0]) * dL_dc_xy + 2 * Tc[1][0] * Tc[1][2] * dL_dc_yy;
    dcov[4] = 2 * Tc[0][2] * Tc[0][1] * dL_dc_xx + (Tc[0][1] * Tc[1][2] + Tc[0][2] * Tc[1][1]) * dL_dc_xy + 2 * Tc[1][1] * Tc[1][2] * dL_dc_yy;
  } else {
#pragma unroll
    for (int i = 0; i < 6; ++i) dcov[i] = 0.f;
  }
#pragma unroll
  for (int i = 0; i < 6; ++i) a.dL_dcov3D[6 * (size_t)idx + i] = dcov[i];

  const float (*Tc)[3] = T.c;
  const float (*V)[3] = Vrk.c;
  const float r0x = Tc[0][0] * V[0][0] + Tc[0][1] * V[0][1] + Tc[0][2] * V[0][2];
  const float r0y = Tc[0][0] * V[1][0] + Tc[0][1] * V[1][1] + Tc[0][2] * V[1][2];
  const float r0z = Tc[0][0] * V[2][0] + Tc[0][1] * V[2][1] + Tc[0][2] * V[2][2];
  const float r1x = Tc[1][0] * V[0][0] + Tc[1][1] * V[0][1] + Tc[1][2] * V[0][2];
  const float r1y = Tc[1][0] * V[1][0] + Tc[1][1] * V[1][1] + Tc[1][2] * V[1][2];
  const float r1z = Tc[1][0] * V[2][0] + Tc[1][1] * V[2][1] + Tc[1][2] * V[2][2];
  const float dL_dT00 = 2 * r0x * dL_dc_xx + r1x * dL_dc_xy;
  const float dL_dT01 = 2 * r0y * dL_dc_xx + r1y * dL_dc_xy;
  const float dL_dT02 = 2 * r0z * dL_dc_xx + r1z * dL_dc_xy;
  const float dL_dT10 = 2 * r1x * dL_dc_yy + r0x * dL_dc_xy;
  const float dL_dT11 = 2 * r1y * dL_dc_yy + r0y * dL_dc_xy;
  const float dL_dT12 = 2 * r1z * dL_dc_yy + r0z * dL_dc_xy;
  const float dL_dJ00 = W.c[0][0] * dL_dT00 + W.c[0][1] * dL_dT01 + W.c[0][2] * dL_dT02;
  const float dL_dJ02 = W.c[2][0] * dL_dT00 + W.c[2][1] * dL_dT01 + W.c[2][2] * dL_dT02;
  const float dL_dJ11 = W.c[1][0] * dL_dT10 + W.c[1][1] * dL_dT11 + W.c[1][2] * dL_dT12;
  const float dL_dJ12 = W.c[2][0] * dL_dT10 + W.c[2][1] * dL_dT11 + W.c[2][2] * dL_dT12;
  const float tz = 1.f / t.z, tz2 = tz * tz, tz3 = tz2 * tz;
  const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
  const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
  float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12;
  if (a.has_invdepth) dL_dtz -= g[15] / (t.z * t.z);
  // transformVec4x3Transpose
  float3 dmean = make_float3(vm[0] * dL_dtx + vm[1] * dL_dty + vm[2] * dL_dtz, vm[4] * dL_dtx + vm[5] * dL_dty + vm[6] * dL_dtz,
                             vm[8] * dL_dtx + vm[9] * dL_dty + vm[10] * dL_dtz);

  // ---- BACKWARD::preprocessCUDA (:398-449): projection path of the (full) 2D-mean gradient
  const float* proj = a.projmatrix;
  const float m_hom_w = proj[3] * mean.x + proj[7] * mean.y + proj[11] * mean.z + proj[15];
  const float m_w = 1.0f / (m_hom_w + 0.0000001f);
  const float mul1 = (proj[0] * mean.x + proj[4] * mean.y + proj[8] * mean.z + proj[12]) * m_w * m_w;
  const float mul2 = (proj[1] * mean.x + proj[5] * mean.y + proj[9] * mean.z + proj[13]) * m_w * m_w;
  const float d2x = g[0], d2y = g[1];
  dmean.x += (proj[0] * m_w - proj[3] * mul1) * d2x + (proj[1] * m_w - proj[3] * mul2) * d2y;
  dmean.y += (proj[4] * m_w - proj[7] * mul1) * d2x + (proj[5] * m_w - proj[7] * mul2) * d2y;
  dmean.z += (proj[8] * m_w - proj[11] * mul1) * d2x + (proj[9] * m_w - proj[11] * mul2) * d2y;

  if (a.shs) {
    const unsigned cb = a.clamped[idx];
    const V3 dRGB = {(cb & 1u) ? 0.f : g[8], (cb & 2u) ? 0.f : g[9], (cb & 4u) ? 0.f : g[10]};
    const V3 dir = {mean.x - a.campos[0], mean.y - a.campos[1], mean.z - a.campos[2]};
    const V3 dm = sh_backward(a.D, a.M, a.shs + (size_t)idx * a.M * 3, dir, dRGB, a.dL_dsh + (size_t)idx * a.M * 3);
    dmean.x += dm.x; dmean.y += dm.y; dmean.z += dm.z;
  }
  a.dL_dmeans3D[3 * idx] = dmean.x; a.dL_dmeans3D[3 * idx + 1] = dmean.y; a.dL_dmeans3D[3 * idx + 2] = dmean.z;

  // ---- computeCov3D backward (:330-393)
  if (a.scales) {
    const float4 q = reinterpret_cast<const float4*>(a.rotations)[idx];
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    M3 R = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
             {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
             {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
    const float sx = a.mod * a.scales[3 * idx], sy = a.mod * a.scales[3 * idx + 1], sz = a.mod * a.scales[3 * idx + 2];
    M3 S = {{{sx, 0.f, 0.f}, {0.f, sy, 0.f}, {0.f, 0.f, sz}}};
    const M3 Mm = mul(S, R);
    M3 dSigma = {{{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]}, {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                  {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}}};
    M3 M2;
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int rr = 0; rr < 3; ++rr) M2.c[c][rr] = 2.0f * Mm.c[c][rr];
    const M3 dL_dM = mul(M2, dSigma);
    const M3 Rt = transpose(R);
    M3 dMt = transpose(dL_dM);
    a.dL_dscales[3 * idx + 0] = Rt.c[0][0] * dMt.c[0][0] + Rt.c[0][1] * dMt.c[0][1] + Rt.c[0][2] * dMt.c[0][2];
    a.dL_dscales[3 * idx + 1] = Rt.c[1][0] * dMt.c[1][0] + Rt.c[1][1] * dMt.c[1][1] + Rt.c[1][2] * dMt.c[1][2];
    a.dL_dscales[3 * idx + 2] = Rt.c[2][0] * dMt.c[2][0] + Rt.c[2][1] * dMt.c[2][1] + Rt.c[2][2] * dMt.c[2][2];
#pragma unroll
    for (int k = 0; k < 3; ++k) { dMt.c[0][k] *= sx; dMt.c[1][k] *= sy; dMt.c[2][k] *= sz; }
    float4 dq;
    dq.x = 2 * z * (dMt.c[0][1] - dMt.c[1][0]) + 2 * y * (dMt.c[2][0] - dMt.c[0][2]) + 2 * x * (dMt.c[1][2] - dMt.c[2][1]);
    dq.y = 2 * y * (dMt.c[1][0] + dMt.c[0][1]) + 2 * z * (dMt.c[2][0] + dMt.c[0][2]) + 2 * r * (dMt.c[1][2] - dMt.c[2][1]) - 4 * x * (dMt.c[2][2] + dMt.c[1][1]);
    dq.z = 2 * x * (dMt.c[1][0] + dMt.c[0][1]) + 2 * r * (dMt.c[2][0] - dMt.c[0][2]) + 2 * z * (dMt.c[1][2] + dMt.c[2][1]) - 4 * y * (dMt.c[2][2] + dMt.c[0][0]);
    dq.w = 2 * r * (dMt.c[0][1] - dMt.c[1][0]) + 2 * x * (dMt.c[2][0] + dMt.c[0][2]) + 2 * y * (dMt.c[1][2] + dMt.c[2][1]) - 4 * z * (dMt.c[1][1] + dMt.c[0][0]);
    reinterpret_cast<float4*>(a.dL_drot)[idx] = dq;
  }
}

int gauss_preprocess_backward(const b200gs_gauss_bwd& f, const GaussGeom& g, const float* acc, cudaStream_t s) {
  GPreBwdArgs a;
  a.P = f.P; a.D = f.sh_degree; a.M = f.sh_coeffs; a.mod = f.scale_modifier;
  a.tan_fovx = f.tan_fovx; a.tan_fovy = f.tan_fovy;
  a.focal_y = f.height / (2.0f * f.tan_fovy);   // G/cuda_rasterizer/rasterizer_impl.cu:402-403
  a.focal_x = f.width / (2.0f * f.tan_fovx);
  a.antialiasing = f.antialiasing;
  a.means3D = f.means3D; a.radii = f.radii; a.shs = f.shs; a.clamped = g.clamped; a.opacities = f.opacities;
  a.scales = f.scales; a.rotations = f.rotations;
  a.cov3D = f.cov3D_precomp ? f.cov3D_precomp : g.cov3D;
  a.viewmatrix = f.viewmatrix; a.projmatrix = f.projmatrix; a.campos = f.campos; a.acc = acc;
  a.has_invdepth = (f.dL_dout_invdepth != nullptr && f.dL_dinvdepths != nullptr) ? 1 : 0;
  a.dL_dmeans2D = f.dL_dmeans2D; a.dL_dconic = f.dL_dconic; a.dL_dopacity = f.dL_dopacity; a.dL_dcolors = f.dL_dcolors;
  a.dL_dnormals = f.dL_dnormals; a.dL_drefl = f.dL_drefl; a.dL_dinvdepths = f.dL_dinvdepths;
  a.dL_dmeans3D = f.dL_dmeans3D; a.dL_dcov3D = f.dL_dcov3D; a.dL_dsh = f.dL_dsh; a.dL_dscales = f.dL_dscales;
  a.dL_drot = f.dL_drotations;
  if (f.shs && !f.dL_dsh) { set_error("dL_dsh must be provided when shs is"); return B200GS_EINVAL; }
  StageScope _t(ST_PREPROCESS_BWD, s);
  gauss_preprocess_bwd_kernel<<<(f.P + 255) / 256, 256, 0, s>>>(a);
  B200GS_LAUNCHED("gauss_preprocess_bwd_kernel", f.debug, s);
  return B200GS_OK;
}

}  // namespace b200gs
