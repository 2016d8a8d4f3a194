// gauss_preprocess.cu -- K1g: per-Gaussian preprocess of the 3DGS (EWA) path.
//
// Restates FORWARD::preprocessCUDA of the reference's diff-gaussian-rasterization
// (G/cuda_rasterizer/forward.cu:151-269) with computeCov3D (:114-148), computeCov2D (:74-109),
// in_frustum / ndc2Pix / getRect (G/cuda_rasterizer/auxiliary.h) and computeColorFromSH (:20-71).
//
// radii, tiles_touched and the tile rect depend on a long fp32 chain (quaternion -> Sigma -> J W
// projection -> eigenvalues).  To land on the same bits as the reference build, the 3x3 products
// are written as the same scalar expression trees glm's operator* expands to
// (R[c][r] = A[0][r]*B[c][0] + A[1][r]*B[c][1] + A[2][r]*B[c][2], literal zeros kept), so nvcc's
// fp-contraction makes the same fuse/round choices; ndc2Pix is evaluated in double like the
// reference's ((v + 1.0) * S - 1.0) * 0.5.
#include "common.cuh"
#include "sh.cuh"
#include "gauss_math.cuh"

namespace b200gs {

struct GaussPreArgs {
  int P, D, M, W, H, gx, gy;
  float mod, tan_fovx, tan_fovy, focal_x, focal_y;
  int prefiltered, antialiasing;
  const float* means3D;
  const float* scales;
  const float* rotations;
  const float* opacities;
  const float* shs;
  const float* colors_precomp;
  const float* cov3D_precomp;
  const float* normals;
  const float* refl;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  int* radii;
  float* depths;
  uint32_t* tiles_touched;
  uint8_t* clamped;
  uint2* rects;
  GaussRec* rec;
  float* cov3D;
  CullRec* cull;
  int* flags;
};

// computeCov3D (forward.cu:114-148): Sigma = (S R)^T (S R); the quaternion is used as given.
__device__ __forceinline__ void cov3d_from_scale_rot(float sx, float sy, float sz, float mod, float4 q, float* cov3D) {
  M3 S = {{{mod * sx, 0.0f, 0.0f}, {0.0f, mod * sy, 0.0f}, {0.0f, 0.0f, mod * sz}}};
  const float r = q.x, x = q.y, y = q.z, z = q.w;
  M3 R = {{{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
           {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
           {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}}};
  const M3 Mm = mul(S, R);
  const M3 Sigma = mul(transpose(Mm), Mm);
  cov3D[0] = Sigma.c[0][0]; cov3D[1] = Sigma.c[0][1]; cov3D[2] = Sigma.c[0][2];
  cov3D[3] = Sigma.c[1][1]; cov3D[4] = Sigma.c[1][2]; cov3D[5] = Sigma.c[2][2];
}

// computeCov2D (forward.cu:74-109)
__device__ __forceinline__ float3 cov2d(float3 mean, float focal_x, float focal_y, float tan_fovx, float tan_fovy,
                                        const float* cov3D, const float* vm) {
  float3 t;
  t.x = vm[0] * mean.x + vm[4] * mean.y + vm[8] * mean.z + vm[12];
  t.y = vm[1] * mean.x + vm[5] * mean.y + vm[9] * mean.z + vm[13];
  t.z = vm[2] * mean.x + vm[6] * mean.y + vm[10] * mean.z + vm[14];
  const float limx = 1.3f * tan_fovx;
  const float limy = 1.3f * tan_fovy;
  const float txtz = t.x / t.z;
  const float tytz = t.y / t.z;
  t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
  t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;
  M3 J = {{{focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z)},
           {0.0f, focal_y / t.z, -(focal_y * t.y) / (t.z * t.z)},
           {0.0f, 0.0f, 0.0f}}};
  M3 Wm = {{{vm[0], vm[4], vm[8]}, {vm[1], vm[5], vm[9]}, {vm[2], vm[6], vm[10]}}};
  const M3 T = mul(Wm, J);
  M3 Vrk = {{{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}}};
  const M3 cov = mul(mul(transpose(T), transpose(Vrk)), T);
  return make_float3(cov.c[0][0], cov.c[0][1], cov.c[1][1]);
}

__device__ __forceinline__ float ndc2pix(float v, int S) { return (float)((((double)v + 1.0) * (double)S - 1.0) * 0.5); }

// {alpha >= 1/255} = {d^T conic d <= c2 = 2 ln(255 opa)}, d = centre - pixel (forward.cu:355-366), as
// the same fp64 point-conic record the surfel path uses (common.cuh CullRec): always an ellipse
// here, no low-pass disk.  c2 is inflated by 0.2 % + 0.02, far beyond the fp32 error of the kernels'
// own power / expf.  Consumed by count_kept / emit_kept / tile_hits (binning.cu).
__device__ __forceinline__ CullRec gauss_cull_record(float cx, float cy, float3 conic, float opa) {
  CullRec r;
  r.A = r.B = r.C = r.D = r.E = r.F = 0.0;
  r.cx = cx; r.cy = cy; r.r2 = -1.0f; r.flag = CULL_EVERYTHING;
  if (!(opa >= 0.0f)) return r;
  if (opa < 0.0039f) { r.flag = CULL_NOTHING; return r; }
  const double c2 = 2.0 * (double)logf(255.0f * opa) * 1.002 + 0.02;
  double A = conic.x, B = conic.y, C = conic.z;
  const double x = cx, y = cy;
  double D = -(A * x + B * y), E = -(B * x + C * y);
  double F = (A * x + 2.0 * B * y) * x + C * y * y - c2;
  const double m = fmax(fmax(fabs(A), fabs(C)), fabs(B));
  if (!(m > 0.0) || !(m < 1e300)) return r;
  const double im = 1.0 / m;
  A *= im; B *= im; C *= im; D *= im; E *= im; F *= im;
  if (!(fabs(A) <= 1.0) || !(fabs(B) <= 1.0) || !(fabs(C) <= 1.0)) return r;
  if (!(fabs(D) < 1e30) || !(fabs(E) < 1e30) || !(fabs(F) < 1e30)) return r;
  r.A = A; r.B = B; r.C = C; r.D = D; r.E = E; r.F = F;
  r.flag = 0.0f;
  return r;
}

__global__ void __launch_bounds__(256) gauss_preprocess_kernel(const GaussPreArgs a) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.P) return;
  a.radii[idx] = 0;
  a.tiles_touched[idx] = 0;

  const float* vm = a.viewmatrix;
  const float* pm = a.projmatrix;
  const float3 p = make_float3(a.means3D[3 * idx], a.means3D[3 * idx + 1], a.means3D[3 * idx + 2]);
  // in_frustum (G/cuda_rasterizer/auxiliary.h): cull on view-space depth only
  const float vz = vm[2] * p.x + vm[6] * p.y + vm[10] * p.z + vm[14];
  if (vz <= 0.2f) {
    if (a.prefiltered) atomicExch(&a.flags[0], 1);
    return;
  }
  // transformPoint4x4 + perspective division (forward.cu:197-201)
  const float hx = pm[0] * p.x + pm[4] * p.y + pm[8] * p.z + pm[12];
  const float hy = pm[1] * p.x + pm[5] * p.y + pm[9] * p.z + pm[13];
  const float hw = pm[3] * p.x + pm[7] * p.y + pm[11] * p.z + pm[15];
  const float p_w = 1.0f / (hw + 0.0000001f);
  const float projx = hx * p_w, projy = hy * p_w;

  float c3[6];
  if (a.cov3D_precomp != nullptr) {
#pragma unroll
    for (int k = 0; k < 6; ++k) c3[k] = a.cov3D_precomp[6 * (size_t)idx + k];
  } else {
    const float4 q = reinterpret_cast<const float4*>(a.rotations)[idx];
    cov3d_from_scale_rot(a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2], a.mod, q, c3);
#pragma unroll
    for (int k = 0; k < 6; ++k) a.cov3D[6 * (size_t)idx + k] = c3[k];
  }
  float3 cov = cov2d(p, a.focal_x, a.focal_y, a.tan_fovx, a.tan_fovy, c3, vm);

  // forward.cu:215-243
  constexpr float h_var = 0.3f;
  const float det_cov = cov.x * cov.z - cov.y * cov.y;
  cov.x += h_var;
  cov.z += h_var;
  const float det_cov_plus_h_cov = cov.x * cov.z - cov.y * cov.y;
  float h_convolution_scaling = 1.0f;
  if (a.antialiasing) h_convolution_scaling = sqrtf(fmaxf(0.000025f, det_cov / det_cov_plus_h_cov));
  const float det = det_cov_plus_h_cov;
  if (det == 0.0f) return;
  const float det_inv = 1.f / det;
  const float3 conic = make_float3(cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv);
  const float mid = 0.5f * (cov.x + cov.z);
  const float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
  const float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
  const float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
  const float cx = ndc2pix(projx, a.W), cy = ndc2pix(projy, a.H);
  const int ri = (int)my_radius;
  int x0, y0, x1, y1;
  get_rect(cx, cy, ri, a.gx, a.gy, x0, y0, x1, y1);
  if ((x1 - x0) * (y1 - y0) == 0) return;

  float rgb[3];
  unsigned cb = 0;
  if (a.colors_precomp == nullptr) {
    const float* cp = a.campos;
    float shv[48];
    load_sh_row(a.shs + (size_t)idx * a.M * 3, a.M, a.D, shv);
    sh_to_rgb(a.D, shv, p.x - cp[0], p.y - cp[1], p.z - cp[2], rgb, cb);
  } else {
    rgb[0] = a.colors_precomp[3 * idx]; rgb[1] = a.colors_precomp[3 * idx + 1]; rgb[2] = a.colors_precomp[3 * idx + 2];
  }
  a.clamped[idx] = (uint8_t)cb;
  a.depths[idx] = vz;
  a.radii[idx] = ri;
  a.rects[idx] = make_uint2((uint32_t)x0 | ((uint32_t)x1 << 16), (uint32_t)y0 | ((uint32_t)y1 << 16));
  const float opa = a.opacities[idx] * h_convolution_scaling;
  a.cull[idx] = gauss_cull_record(cx, cy, conic, opa);
  GaussRec* r = a.rec + idx;
  r->q[0] = make_float4(cx, cy, conic.x, conic.y);
  r->q[1] = make_float4(conic.z, opa, 1.0f / vz, a.refl ? a.refl[idx] : 0.0f);
  r->q[2] = make_float4(rgb[0], rgb[1], rgb[2], vz);
  r->q[3] = a.normals ? make_float4(a.normals[3 * idx], a.normals[3 * idx + 1], a.normals[3 * idx + 2], 0.0f)
                      : make_float4(0.f, 0.f, 0.f, 0.f);
  a.tiles_touched[idx] = (uint32_t)((y1 - y0) * (x1 - x0));
}

int gauss_preprocess(const b200gs_gauss_fwd& a, const GaussGeom& g, cudaStream_t s) {
  GaussPreArgs k;
  k.P = a.P; k.D = a.sh_degree; k.M = a.sh_coeffs; k.W = a.width; k.H = a.height;
  k.gx = (a.width + TILE_X - 1) / TILE_X; k.gy = (a.height + TILE_Y - 1) / TILE_Y;
  k.mod = a.scale_modifier; k.tan_fovx = a.tan_fovx; k.tan_fovy = a.tan_fovy;
  k.focal_y = a.height / (2.0f * a.tan_fovy);   // G/cuda_rasterizer/rasterizer_impl.cu:227-228
  k.focal_x = a.width / (2.0f * a.tan_fovx);
  k.prefiltered = a.prefiltered; k.antialiasing = a.antialiasing;
  k.means3D = a.means3D; k.scales = a.scales; k.rotations = a.rotations; k.opacities = a.opacities;
  k.shs = a.shs; k.colors_precomp = a.colors_precomp; k.cov3D_precomp = a.cov3D_precomp;
  k.normals = a.normals; k.refl = a.refl_strengths;
  k.viewmatrix = a.viewmatrix; k.projmatrix = a.projmatrix; k.campos = a.campos;
  k.radii = a.radii; k.depths = g.depths; k.tiles_touched = g.tiles_touched; k.clamped = g.clamped;
  k.rects = g.rects; k.rec = g.rec; k.cov3D = g.cov3D; k.cull = g.cull; k.flags = g.flags;
  StageScope _t(ST_PREPROCESS, s);
  B200GS_CUDA(cudaMemsetAsync(g.flags, 0, 4 * sizeof(int), s));
  gauss_preprocess_kernel<<<(a.P + 255) / 256, 256, 0, s>>>(k);
  B200GS_LAUNCHED("gauss_preprocess_kernel", a.debug, s);
  return B200GS_OK;
}

}  // namespace b200gs
