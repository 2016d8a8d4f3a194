// surfel_preprocess_bwd.cu -- K5s: per-Gaussian chain rule of the surfel path.
//
// Restates BACKWARD::preprocessCUDA (S/cuda_rasterizer/backward.cu:605-660) with
// compute_transmat_aabb (:473-603), the SH backward (:20-139), dnormvdv
// (auxiliary.h:132-142) and quat_to_rotmat_vjp (auxiliary.h:242-286), including
// the reference's quirks: W/H re-derived as int(focal*tan*2) (:637-638), the
// scale_modifier ignored when T is rebuilt (:511), the AABB-centre term applied
// only for a non-zero mean2D gradient (:545), and dL_dmean2D overwritten by the
// densification proxy (:657-659).
//
// Input is the packed [P, ACC_STRIDE] accumulator written by the compositing
// backward; this kernel unpacks it into the reference's output tensors.
#include "common.cuh"
#include "sh.cuh"

#ifndef B200GS_PREBWD_CTAS
#define B200GS_PREBWD_CTAS 2   // resident CTAs per SM the register allocation is bounded for (2: 119 registers 0.47 ms; 3: 80 registers, 380 B of spills, 1.40 ms)
#endif

namespace b200gs {

struct PreBwdArgs {
  int P, D, M, W, H;
  const float* means3D;
  const int* radii;
  const float* shs;
  const uint8_t* clamped;
  const float* scales;
  const float* rotations;
  const float* transMat_precomp;
  const SurfelRec* rec;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  const float* acc;
  float* dL_dmeans2D;
  float* dL_dnormal;
  float* dL_dopacity;
  float* dL_dcolors;
  float* dL_drefl;
  float* dL_dmeans3D;
  float* dL_dtransMat;
  float* dL_dsh;
  float* dL_dscales;
  float* dL_drot;
};

// ACC: the six PARAMETER gradients (opacity, refl, means3D, sh, scales, rotations) are added to what
// the output tensors hold -- gradients of a batch of views accumulate in place (b200gs.dp.GradArena,
// SURVEY.md section 8e); rows of Gaussians this view culled are left alone, so no zero-fill per view.
// The view-space outputs (means2D, colors, normal, transMat) are always overwritten.
template <bool ACC>
__global__ void __launch_bounds__(256, B200GS_PREBWD_CTAS) surfel_preprocess_bwd_kernel(const PreBwdArgs a) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.P || !(a.radii[idx] > 0)) return;

  float g[ACC_STRIDE];
  {
    const float4* ap = reinterpret_cast<const float4*>(a.acc + (size_t)idx * ACC_STRIDE);
#pragma unroll
    for (int k = 0; k < ACC_STRIDE / 4; ++k) {
      const float4 v = ap[k];
      g[4 * k] = v.x; g[4 * k + 1] = v.y; g[4 * k + 2] = v.z; g[4 * k + 3] = v.w;
    }
  }
  // A visible Gaussian no pixel blended (occluded, or below the 1/255 cut everywhere) has an all-zero accumulator
  // row: every gradient this kernel would derive from it is an exact zero, which the zero-filled outputs (or, when
  // accumulating, the sums so far) already hold.  On scenes with occlusion that is most visible Gaussians, and it
  // saves their input reads (237 B) and gradient rows (248 B).
  {
    bool any = false;
#pragma unroll
    for (int k = 0; k < ACC_STRIDE; ++k) any = any || (g[k] != 0.0f);
    if (!any) return;
  }
  // plain per-Gaussian sums straight through
  a.dL_dopacity[idx] = ACC ? a.dL_dopacity[idx] + g[11] : g[11];
  a.dL_dcolors[3 * idx + 0] = g[12]; a.dL_dcolors[3 * idx + 1] = g[13]; a.dL_dcolors[3 * idx + 2] = g[14];
  if (a.dL_drefl) a.dL_drefl[idx] = ACC ? a.dL_drefl[idx] + g[15] : g[15];
  a.dL_dnormal[3 * idx + 0] = g[16]; a.dL_dnormal[3 * idx + 1] = g[17]; a.dL_dnormal[3 * idx + 2] = g[18];

  const bool precomp = (a.scales == nullptr);  // backward.cu:639
  const float* vm = a.viewmatrix;
  const float* pm = a.projmatrix;
  const V3 p = {a.means3D[3 * idx], a.means3D[3 * idx + 1], a.means3D[3 * idx + 2]};

  V3 T0, T1, T2;             // T[0], T[1], T[2] = Tu, Tv, Tw
  float Pm[3][4];            // P[c][r] = world2ndc * ndc2pix
  V3 R0 = {0, 0, 0}, R1 = {0, 0, 0}, R2 = {0, 0, 0};
  float4 q = make_float4(0, 0, 0, 0);
  float2 sc = make_float2(0, 0);
  V3 normal = {0, 0, 0};
  if (precomp) {
    const float* tp = a.transMat_precomp + 9 * (size_t)idx;
    T0 = {tp[0], tp[1], tp[2]}; T1 = {tp[3], tp[4], tp[5]}; T2 = {tp[6], tp[7], tp[8]};
  } else {
    q = reinterpret_cast<const float4*>(a.rotations)[idx];
    sc = reinterpret_cast<const float2*>(a.scales)[idx];
    const float s = rsqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
    const float w = q.x * s, x = q.y * s, y = q.z * s, z = q.w * s;
    R0 = {1.f - 2.f * (y * y + z * z), 2.f * (x * y + w * z), 2.f * (x * z - w * y)};
    R1 = {2.f * (x * y - w * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z + w * x)};
    R2 = {2.f * (x * z + w * y), 2.f * (y * z - w * x), 1.f - 2.f * (x * x + y * y)};
    const V3 L0 = sc.x * R0, L1 = sc.y * R1;  // scale_to_mat(scale, 1.0f): modifier ignored (:511)
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      Pm[0][r] = pm[4 * r] * (float(a.W) / 2.0f) + pm[4 * r + 3] * (float(a.W - 1) / 2.0f);
      Pm[1][r] = pm[4 * r + 1] * (float(a.H) / 2.0f) + pm[4 * r + 3] * (float(a.H - 1) / 2.0f);
      Pm[2][r] = pm[4 * r + 3];
    }
    auto rowdot = [&](int c, V3 h, float hw) { return h.x * Pm[c][0] + h.y * Pm[c][1] + h.z * Pm[c][2] + hw * Pm[c][3]; };
    T0 = {rowdot(0, L0, 0.f), rowdot(0, L1, 0.f), rowdot(0, p, 1.f)};
    T1 = {rowdot(1, L0, 0.f), rowdot(1, L1, 0.f), rowdot(1, p, 1.f)};
    T2 = {rowdot(2, L0, 0.f), rowdot(2, L1, 0.f), rowdot(2, p, 1.f)};
    normal = {vm[0] * R2.x + vm[4] * R2.y + vm[8] * R2.z, vm[1] * R2.x + vm[5] * R2.y + vm[9] * R2.z,
              vm[2] * R2.x + vm[6] * R2.y + vm[10] * R2.z};
  }

  V3 dT0 = {g[0], g[1], g[2]}, dT1 = {g[3], g[4], g[5]}, dT2 = {g[6], g[7], g[8]};
  const float dmx = g[9], dmy = g[10];
  if (dmx != 0.0f || dmy != 0.0f) {  // AABB-centre term (:545-560)
    const V3 t = {9.0f, 9.0f, -1.0f};
    const float d = dot(t, T2 * T2);
    const V3 f = (1.0f / d) * t;
    const V3 a0 = dmx * (f * T2), a1 = dmy * (f * T2);
    V3 a3 = dmx * (f * T0) + dmy * (f * T1);
    const V3 dL_df = dmx * (T0 * T2) + dmy * (T1 * T2);
    const float dL_dd = dot(dL_df, f) * (-1.0f / d);
    const V3 dd_dT3 = 2.0f * (t * T2);
    a3 = a3 + dL_dd * dd_dT3;
    dT0 = dT0 + a0; dT1 = dT1 + a1; dT2 = dT2 + a3;
  }
  // dL_dtransMat output: the render sums; in the precomputed-T path the AABB term is folded in (:561-572)
  {
    float* o = a.dL_dtransMat + 9 * (size_t)idx;
    if (precomp) {
      o[0] = dT0.x; o[1] = dT0.y; o[2] = dT0.z; o[3] = dT1.x; o[4] = dT1.y; o[5] = dT1.z;
      o[6] = dT2.x; o[7] = dT2.y; o[8] = dT2.z;
    } else {
#pragma unroll
      for (int k = 0; k < 9; ++k) o[k] = g[k];
    }
  }
  const float dTs2 = precomp ? dT0.z : g[2], dTs5 = precomp ? dT1.z : g[5];

  V3 dmean = {0, 0, 0};
  if (!precomp) {
    // dL_dM = P * transpose(dL_dT): dL_dM[c][r] = sum_k P[k][r] * dL_dT[k][c]   (:578)
    const float dTc[3][3] = {{dT0.x, dT1.x, dT2.x}, {dT0.y, dT1.y, dT2.y}, {dT0.z, dT1.z, dT2.z}};  // [c][k]
    float dM[3][3];
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int r = 0; r < 3; ++r) dM[c][r] = Pm[0][r] * dTc[c][0] + Pm[1][r] * dTc[c][1] + Pm[2][r] * dTc[c][2];
    // dL_dnormal: view -> world, dual-visible sign (:579-585)
    V3 dtn = {vm[0] * g[16] + vm[1] * g[17] + vm[2] * g[18], vm[4] * g[16] + vm[5] * g[17] + vm[6] * g[18],
              vm[8] * g[16] + vm[9] * g[17] + vm[10] * g[18]};
    const V3 pv = {vm[0] * p.x + vm[4] * p.y + vm[8] * p.z + vm[12], vm[1] * p.x + vm[5] * p.y + vm[9] * p.z + vm[13],
                   vm[2] * p.x + vm[6] * p.y + vm[10] * p.z + vm[14]};
    const float cosv = -(pv.x * normal.x + pv.y * normal.y + pv.z * normal.z);
    const float mult = cosv > 0 ? 1.0f : -1.0f;
    dtn = mult * dtn;
    const V3 dRS0 = {dM[0][0], dM[0][1], dM[0][2]}, dRS1 = {dM[1][0], dM[1][1], dM[1][2]};
    // v_R columns (glm: v_R[c][r])
    const V3 c0 = sc.x * dRS0, c1 = sc.y * dRS1, c2 = dtn;
    const float s = rsqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
    const float w = q.x * s, x = q.y * s, y = q.z * s, z = q.w * s;
    float4 dq;  // quat_to_rotmat_vjp (auxiliary.h:242-286); v_R[c][r]: c0=(.x=[0][0],.y=[0][1],.z=[0][2])
    dq.x = 2.f * (x * (c1.z - c2.y) + y * (c2.x - c0.z) + z * (c0.y - c1.x));
    dq.y = 2.f * (-2.f * x * (c1.y + c2.z) + y * (c0.y + c1.x) + z * (c0.z + c2.x) + w * (c1.z - c2.y));
    dq.z = 2.f * (x * (c0.y + c1.x) - 2.f * y * (c0.x + c2.z) + z * (c1.z + c2.y) + w * (c2.x - c0.z));
    dq.w = 2.f * (x * (c0.z + c2.x) + y * (c1.z + c2.y) - 2.f * z * (c0.x + c1.y) + w * (c0.y - c1.x));
    float2 ds = make_float2(dot(dRS0, R0), dot(dRS1, R1));
    if (ACC) {
      const float4 oq = reinterpret_cast<const float4*>(a.dL_drot)[idx];
      const float2 os = reinterpret_cast<const float2*>(a.dL_dscales)[idx];
      dq.x += oq.x; dq.y += oq.y; dq.z += oq.z; dq.w += oq.w;
      ds.x += os.x; ds.y += os.y;
    }
    reinterpret_cast<float4*>(a.dL_drot)[idx] = dq;
    reinterpret_cast<float2*>(a.dL_dscales)[idx] = ds;
    dmean = {dM[2][0], dM[2][1], dM[2][2]};
  }

  if (a.shs) {  // :653-654
    const unsigned cb = a.clamped[idx];
    const V3 dRGB = {(cb & 1u) ? 0.f : g[12], (cb & 2u) ? 0.f : g[13], (cb & 4u) ? 0.f : g[14]};
    const V3 dir = {p.x - a.campos[0], p.y - a.campos[1], p.z - a.campos[2]};
    const V3 dm = sh_backward<ACC>(a.D, a.M, a.shs + (size_t)idx * a.M * 3, dir, dRGB, a.dL_dsh + (size_t)idx * a.M * 3);
    dmean = dmean + dm;
  }
  if (!precomp || a.shs) {
    float* dm3 = a.dL_dmeans3D + 3 * (size_t)idx;
    if (ACC) { dm3[0] += dmean.x; dm3[1] += dmean.y; dm3[2] += dmean.z; }
    else { dm3[0] = dmean.x; dm3[1] = dmean.y; dm3[2] = dmean.z; }
  }

  // densification proxy (:656-659)
  const float depth = precomp ? a.transMat_precomp[9 * (size_t)idx + 8] : a.rec[idx].q[2].x;
  a.dL_dmeans2D[3 * idx + 0] = dTs2 * depth * 0.5f * float(a.W);
  a.dL_dmeans2D[3 * idx + 1] = dTs5 * depth * 0.5f * float(a.H);
}

int surfel_preprocess_backward(const b200gs_surfel_bwd& f, const SurfelGeom& g, const float* acc, cudaStream_t s) {
  PreBwdArgs a;
  a.P = f.P; a.D = f.sh_degree; a.M = f.sh_coeffs;
  // backward.cu:637-638 with focal = size / (2 tan) from rasterizer_impl.cu:402-403, all fp32
  const float focal_y = f.height / (2.0f * f.tan_fovy);
  const float focal_x = f.width / (2.0f * f.tan_fovx);
  a.W = int(focal_x * f.tan_fovx * 2);
  a.H = int(focal_y * f.tan_fovy * 2);
  a.means3D = f.means3D; a.radii = f.radii; a.shs = f.shs; a.clamped = g.clamped;
  a.scales = f.scales; a.rotations = f.rotations; a.transMat_precomp = f.transMat_precomp; a.rec = g.rec;
  a.viewmatrix = f.viewmatrix; a.projmatrix = f.projmatrix; a.campos = f.campos; a.acc = acc;
  a.dL_dmeans2D = f.dL_dmeans2D; a.dL_dnormal = f.dL_dnormal; a.dL_dopacity = f.dL_dopacity;
  a.dL_dcolors = f.dL_dcolors; a.dL_drefl = f.dL_drefl; a.dL_dmeans3D = f.dL_dmeans3D;
  a.dL_dtransMat = f.dL_dtransMat; a.dL_dsh = f.dL_dsh; a.dL_dscales = f.dL_dscales; a.dL_drot = f.dL_drotations;
  if (f.shs && !f.dL_dsh) { set_error("dL_dsh must be provided when shs is"); return B200GS_EINVAL; }
  StageScope _t(ST_PREPROCESS_BWD, s);
  if (f.accumulate) surfel_preprocess_bwd_kernel<true><<<(f.P + 255) / 256, 256, 0, s>>>(a);
  else surfel_preprocess_bwd_kernel<false><<<(f.P + 255) / 256, 256, 0, s>>>(a);
  B200GS_LAUNCHED("surfel_preprocess_bwd_kernel", f.debug, s);
  return B200GS_OK;
}

}  // namespace b200gs
