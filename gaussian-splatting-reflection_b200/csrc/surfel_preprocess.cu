// surfel_preprocess.cu -- K1s: per-Gaussian preprocess of the 2DGS/surfel path.
//
// Restates FORWARD::preprocessCUDA (S/cuda_rasterizer/forward.cu:148-253) with
// compute_transmat (:75-115), compute_aabb (:119-145), quat_to_rotmat
// (auxiliary.h:217-239), getRect (auxiliary.h:71-81), in_frustum (:189-214) and
// computeColorFromSH (forward.cu:20-71).
//
// Everything that feeds an integer output (radii, tiles_touched, the depth bits
// of the sort keys, the tile rect) is written with explicit __fmaf_rn/__fmul_rn/
// __fadd_rn in the exact fused/unfused shape nvcc 12.9 produces for the
// reference's glm expression trees (fp-contract on, no fast-math), so those
// outputs are bit-identical to the reference extension built with the same
// toolkit, independent of how this file is optimised.  Terms the reference
// multiplies by a literal 0 (glm's dense mat products with the sparse S and
// ndc2pix matrices) are dropped: for finite inputs fma(x, 0, y) == y.
//
// Output is one packed 80-B record per visible Gaussian (see common.cuh) instead
// of the reference's six SoA arrays, so the compositing kernels stage a Gaussian
// with five 16-B copies.
#include "common.cuh"
#include "sh.cuh"

#ifndef B200GS_PRE_CTAS
#define B200GS_PRE_CTAS 3   // resident CTAs per SM the register allocation is bounded for (2: 86 registers 0.25 ms, 3: 80 registers 0.21 ms, 4: 64 registers 0.23 ms)
#endif

namespace b200gs {

struct SurfelPreArgs {
  int P, D, M, W, H, gx, gy;
  float mod;
  int prefiltered;
  const float* means3D;
  const float* scales;
  const float* rotations;
  const float* opacities;
  const float* shs;
  const float* colors_precomp;
  const float* transMat_precomp;
  const float* refl;
  const uint8_t* mask;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  int* radii;
  float* gaussian_weights;   // zeroed here: the compositing takes an atomicMax on it
  float* depths;
  uint32_t* tiles_touched;
  uint8_t* clamped;
  uint2* rects;
  SurfelRec* rec;
  CullRec* cull;
  int* flags;
};

// Conservative description of the pixel region where this Gaussian can reach
// alpha = min(0.99, opa*exp(-rho/2)) >= 1/255, i.e. rho = min(rho3d, rho2d) <= c2 = 2 ln(255 opa):
// the union of (a) {rho3d <= c2} and (b) the low-pass disk {2 |xy - pix|^2 <= c2}.
// (a): the ray-splat intersection of pixel p = (x, y, 1) is (u, v, 1) ~ cross(x Tw - Tu, y Tw - Tv)
//      = x (Tv x Tw) + y (Tw x Tu) + (Tu x Tv)   (forward.cu:362-376; the x*y term cancels), a
//      LINEAR map of p, so rho3d = u^2 + v^2 <= c2 is the quadratic form
//      f(p) = (a1.p)^2 + (a2.p)^2 - c2 (a3.p)^2 <= 0 with a_i = i-th components of the three cross
//      products.  Unlike the bounding-box closed form of compute_aabb (forward.cu:119-145) this
//      stays valid when the splat crosses the camera plane and the conic is a hyperbola -- the case
//      where the reference's rect degenerates to "the whole screen" and the Gaussian sits in every
//      tile's list without ever contributing.
// Evaluated in fp64 from the fp32 T rows, c2 inflated to 1.002 c2 + 0.02 (>> fp32 error of rho and
// expf), coefficients normalised.  tile_hits_kernel (binning.cu) minimises f over every 8x4 warp
// block of a tile (dilated by half a pixel); the compositing kernels skip (warp, Gaussian) pairs
// that miss, which the reference would have rejected pixel by pixel with `alpha < 1/255` --
// results stay bit-identical.  Doubtful arithmetic -> "everything".
__device__ __forceinline__ CullRec cull_record(const float* Tu, const float* Tv, const float* Tw, float cx, float cy,
                                               float opa) {
  CullRec r;
  r.A = r.B = r.C = r.D = r.E = r.F = 0.0;
  r.cx = cx; r.cy = cy; r.r2 = 0.0f; r.flag = CULL_EVERYTHING;
  if (!(opa >= 0.0f)) return r;                                  // NaN / negative: leave it to the exact test
  if (opa < 0.0039f) { r.flag = CULL_NOTHING; return r; }        // opa*G <= opa < 1/255 always
  const double c2 = 2.0 * (double)logf(255.0f * opa) * 1.002 + 0.02;
  const double u0 = Tu[0], u1 = Tu[1], u2 = Tu[2], v0 = Tv[0], v1 = Tv[1], v2 = Tv[2], w0 = Tw[0], w1 = Tw[1], w2 = Tw[2];
  // N0 = Tv x Tw, N1 = Tw x Tu, N2 = Tu x Tv
  const double n0x = v1 * w2 - v2 * w1, n0y = v2 * w0 - v0 * w2, n0z = v0 * w1 - v1 * w0;
  const double n1x = w1 * u2 - w2 * u1, n1y = w2 * u0 - w0 * u2, n1z = w0 * u1 - w1 * u0;
  const double n2x = u1 * v2 - u2 * v1, n2y = u2 * v0 - u0 * v2, n2z = u0 * v1 - u1 * v0;
  double A = n0x * n0x + n0y * n0y - c2 * n0z * n0z;
  double B = n0x * n1x + n0y * n1y - c2 * n0z * n1z;
  double C = n1x * n1x + n1y * n1y - c2 * n1z * n1z;
  double D = n0x * n2x + n0y * n2y - c2 * n0z * n2z;
  double E = n1x * n2x + n1y * n2y - c2 * n1z * n2z;
  double F = n2x * n2x + n2y * n2y - c2 * n2z * n2z;
  const double m = fmax(fmax(fabs(A), fabs(C)), fabs(B));
  if (!(m > 0.0) || !(m < 1e300)) return r;
  const double im = 1.0 / m;
  A *= im; B *= im; C *= im; D *= im; E *= im; F *= im;
  if (!(fabs(A) <= 1.0) || !(fabs(B) <= 1.0) || !(fabs(C) <= 1.0)) return r;   // also catches NaN
  if (!(fabs(D) < 1e30) || !(fabs(E) < 1e30) || !(fabs(F) < 1e30)) return r;
  if (!(fabsf(cx) < 1e30f) || !(fabsf(cy) < 1e30f)) return r;
  r.A = A; r.B = B; r.C = C; r.D = D; r.E = E; r.F = F;
  r.r2 = (float)(0.5 * c2 * (1.0 + 1e-6));
  r.flag = 0.0f;
  return r;
}

__global__ void __launch_bounds__(256, B200GS_PRE_CTAS) surfel_preprocess_kernel(const SurfelPreArgs a) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.P) return;

  // "If this isn't changed, this Gaussian will not be processed further" (forward.cu:180-183)
  int radius_out = 0;
  uint32_t tiles = 0;

  const float px = a.means3D[3 * idx + 0], py = a.means3D[3 * idx + 1], pz = a.means3D[3 * idx + 2];
  const float* __restrict__ vm = a.viewmatrix;
  const float* __restrict__ pm = a.projmatrix;

  // in_frustum (auxiliary.h:189-214): p_view = transformPoint4x3(p, view); cull z <= 0.2
  const float vx = __fadd_rn(vm[12], dot3_ref(px, vm[0], py, vm[4], pz, vm[8]));
  const float vy = __fadd_rn(vm[13], dot3_ref(px, vm[1], py, vm[5], pz, vm[9]));
  const float vz = __fadd_rn(vm[14], dot3_ref(px, vm[2], py, vm[6], pz, vm[10]));

  bool alive = vz > 0.2f;
  if (!alive && a.prefiltered) atomicExch(&a.flags[0], 1);

  float Tu[3], Tv[3], Tw[3], nrm[3];
  if (alive) {
    if (a.transMat_precomp == nullptr) {
      // quat_to_rotmat (auxiliary.h:217-239); storage (w,x,y,z) = q0..q3
      const float4 q = reinterpret_cast<const float4*>(a.rotations)[idx];
      const float q0 = q.x, q1 = q.y, q2 = q.z, q3 = q.w;
      const float n2 = __fmaf_rn(q2, q2, __fmaf_rn(q1, q1, __fmaf_rn(q0, q0, __fmul_rn(q3, q3))));
      const float s = rsqrtf(n2);
      const float w = __fmul_rn(q0, s), x = __fmul_rn(q1, s), y = __fmul_rn(q2, s), z = __fmul_rn(q3, s);
      // Contraction exactly as in the reference build (SASS of preprocessCUDA<3>, nvcc 12.9):
      // in a*b + c*d the SECOND product is rounded and the first is fused onto it.
      const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
      const float wz = __fmul_rn(w, z), wy = __fmul_rn(w, y), wx = __fmul_rn(w, x);
      // glm column-major R[c][r]
      const float R00 = __fsub_rn(1.0f, __fmul_rn(2.0f, __fadd_rn(yy, zz)));
      const float R01 = __fmul_rn(2.0f, __fmaf_rn(x, y, wz));
      const float R02 = __fmul_rn(2.0f, __fmaf_rn(x, z, -wy));
      const float R10 = __fmul_rn(2.0f, __fmaf_rn(x, y, -wz));
      const float R11 = __fsub_rn(1.0f, __fmul_rn(2.0f, __fmaf_rn(x, x, zz)));
      const float R12 = __fmul_rn(2.0f, __fmaf_rn(y, z, wx));
      const float R20 = __fmul_rn(2.0f, __fmaf_rn(x, z, wy));
      const float R21 = __fmul_rn(2.0f, __fmaf_rn(y, z, -wx));
      const float R22 = __fsub_rn(1.0f, __fmul_rn(2.0f, __fmaf_rn(x, x, yy)));
      // L = R * diag(mod*sx, mod*sy, 1)   (forward.cu:88-90)
      const float2 sc = reinterpret_cast<const float2*>(a.scales)[idx];
      const float S00 = __fmul_rn(a.mod, sc.x), S11 = __fmul_rn(a.mod, sc.y);
      const float L0[3] = {__fmul_rn(S00, R00), __fmul_rn(S00, R01), __fmul_rn(S00, R02)};
      const float L1[3] = {__fmul_rn(S11, R10), __fmul_rn(S11, R11), __fmul_rn(S11, R12)};
      // M = transpose(splat2world) * world2ndc  (forward.cu:93-112); M[c][r], c = clip component
      float M0[4], M1[4], M2[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        M0[c] = __fmaf_rn(L0[2], pm[8 + c], __fmaf_rn(L0[0], pm[c], __fmul_rn(L0[1], pm[4 + c])));
        M1[c] = __fmaf_rn(L1[2], pm[8 + c], __fmaf_rn(L1[0], pm[c], __fmul_rn(L1[1], pm[4 + c])));
        M2[c] = __fadd_rn(pm[12 + c], __fmaf_rn(pz, pm[8 + c], __fmaf_rn(px, pm[c], __fmul_rn(py, pm[4 + c]))));
      }
      // T = M * ndc2pix: T[0][r] = fma((W-1)/2, M[3][r], W/2 * M[0][r]) etc.
      const float hw = __fmul_rn((float)a.W, 0.5f), hw1 = __fmul_rn((float)(a.W - 1), 0.5f);
      const float hh = __fmul_rn((float)a.H, 0.5f), hh1 = __fmul_rn((float)(a.H - 1), 0.5f);
      Tu[0] = __fmaf_rn(hw1, M0[3], __fmul_rn(hw, M0[0]));
      Tu[1] = __fmaf_rn(hw1, M1[3], __fmul_rn(hw, M1[0]));
      Tu[2] = __fmaf_rn(hw1, M2[3], __fmul_rn(hw, M2[0]));
      Tv[0] = __fmaf_rn(hh1, M0[3], __fmul_rn(hh, M0[1]));
      Tv[1] = __fmaf_rn(hh1, M1[3], __fmul_rn(hh, M1[1]));
      Tv[2] = __fmaf_rn(hh1, M2[3], __fmul_rn(hh, M2[1]));
      Tw[0] = M0[3]; Tw[1] = M1[3]; Tw[2] = M2[3];
      // normal = transformVec4x3(L[2] = R[2], viewmatrix)  (forward.cu:113)
      nrm[0] = dot3_ref(R20, vm[0], R21, vm[4], R22, vm[8]);
      nrm[1] = dot3_ref(R20, vm[1], R21, vm[5], R22, vm[9]);
      nrm[2] = dot3_ref(R20, vm[2], R21, vm[6], R22, vm[10]);
    } else {
      const float* tp = a.transMat_precomp + 9 * (size_t)idx;  // forward.cu:200-208
      Tu[0] = tp[0]; Tu[1] = tp[1]; Tu[2] = tp[2];
      Tv[0] = tp[3]; Tv[1] = tp[4]; Tv[2] = tp[5];
      Tw[0] = tp[6]; Tw[1] = tp[7]; Tw[2] = tp[8];
      nrm[0] = 0.0f; nrm[1] = 0.0f; nrm[2] = 1.0f;
    }
    // DUAL_VISIABLE (forward.cu:211-216): cos = -(p_view . normal); cos == 0 -> drop
    const float pn = __fmaf_rn(vz, nrm[2], __fmaf_rn(vx, nrm[0], __fmul_rn(vy, nrm[1])));
    if (pn == 0.0f) alive = false;
    const float mult = pn < 0.0f ? 1.0f : -1.0f;
    nrm[0] = __fmul_rn(nrm[0], mult); nrm[1] = __fmul_rn(nrm[1], mult); nrm[2] = __fmul_rn(nrm[2], mult);
  }

  float cx = 0.f, cy = 0.f;
  int x0 = 0, y0 = 0, x1 = 0, y1 = 0;
  if (alive) {
    // compute_aabb (forward.cu:119-145), cutoff = 3
    const float d = __fmaf_rn(-Tw[2], Tw[2], __fmaf_rn(__fmul_rn(Tw[0], Tw[0]), 9.0f, __fmul_rn(__fmul_rn(Tw[1], Tw[1]), 9.0f)));
    if (d == 0.0f) {
      alive = false;
    } else {
      const float fz = __frcp_rn(d);          // f = (1/d) * (9, 9, -1)
      const float fxy = __fmul_rn(fz, 9.0f);
      cx = __fmaf_rn(__fmul_rn(Tu[2], Tw[2]), -fz,
                     __fmaf_rn(fxy, __fmul_rn(Tu[0], Tw[0]), __fmul_rn(fxy, __fmul_rn(Tu[1], Tw[1]))));
      cy = __fmaf_rn(__fmul_rn(Tv[2], Tw[2]), -fz,
                     __fmaf_rn(fxy, __fmul_rn(Tv[0], Tw[0]), __fmul_rn(fxy, __fmul_rn(Tv[1], Tw[1]))));
      const float du = __fmaf_rn(fxy, __fmul_rn(Tu[0], Tu[0]), __fmul_rn(fxy, __fmul_rn(Tu[1], Tu[1])));
      const float dv = __fmaf_rn(fxy, __fmul_rn(Tv[0], Tv[0]), __fmul_rn(fxy, __fmul_rn(Tv[1], Tv[1])));
      const float h0x = __fmaf_rn(cx, cx, __fmaf_rn(__fmul_rn(Tu[2], Tu[2]), fz, -du));
      const float h0y = __fmaf_rn(cy, cy, __fmaf_rn(__fmul_rn(Tv[2], Tv[2]), fz, -dv));
      const float hx = __fsqrt_rn(fmaxf(h0x, 1e-4f));
      const float hy = __fsqrt_rn(fmaxf(h0y, 1e-4f));
      // radius = ceil(max(max(hx, hy), 3 * 0.707106))  (forward.cu:232)
      const float radius = ceilf(fmaxf(fmaxf(hx, hy), 2.12131786346435546875f));
      const int ri = (int)radius;
      get_rect(cx, cy, ri, a.gx, a.gy, x0, y0, x1, y1);
      tiles = (uint32_t)((x1 - x0) * (y1 - y0));
      if (tiles == 0) alive = false;
      else radius_out = ri;
    }
  }

  a.radii[idx] = radius_out;
  a.gaussian_weights[idx] = 0.0f;
  a.tiles_touched[idx] = alive ? tiles : 0u;
  if (!alive) return;

  // Colour (forward.cu:240-246)
  float rgb[3];
  unsigned cb = 0;
  if (a.colors_precomp == nullptr) {
    const float* cp = a.campos;
    float shv[48];
    load_sh_row(a.shs + (size_t)idx * a.M * 3, a.M, a.D, shv);
    sh_to_rgb(a.D, shv, px - cp[0], py - cp[1], pz - cp[2], rgb, cb);
  } else {
    rgb[0] = a.colors_precomp[3 * idx + 0];
    rgb[1] = a.colors_precomp[3 * idx + 1];
    rgb[2] = a.colors_precomp[3 * idx + 2];
  }
  a.clamped[idx] = (uint8_t)cb;
  a.depths[idx] = vz;
  a.rects[idx] = make_uint2((uint32_t)x0 | ((uint32_t)x1 << 16), (uint32_t)y0 | ((uint32_t)y1 << 16));

  const float opa = a.opacities[idx];
  const float refl = a.refl ? a.refl[idx] : 0.0f;
  const float mk = (a.mask && a.mask[idx]) ? 1.0f : 0.0f;
  a.cull[idx] = cull_record(Tu, Tv, Tw, cx, cy, opa);
  SurfelRec* r = a.rec + idx;
  r->q[0] = make_float4(Tu[0], Tu[1], Tu[2], Tv[0]);
  r->q[1] = make_float4(Tv[1], Tv[2], Tw[0], Tw[1]);
  r->q[2] = make_float4(Tw[2], cx, cy, opa);
  r->q[3] = make_float4(nrm[0], nrm[1], nrm[2], refl);
  r->q[4] = make_float4(rgb[0], rgb[1], rgb[2], mk);
}

int surfel_preprocess(const b200gs_surfel_fwd& a, const SurfelGeom& g, cudaStream_t s) {
  SurfelPreArgs k;
  k.P = a.P; k.D = a.sh_degree; k.M = a.sh_coeffs; k.W = a.width; k.H = a.height;
  k.gx = (a.width + TILE_X - 1) / TILE_X; k.gy = (a.height + TILE_Y - 1) / TILE_Y;
  k.mod = a.scale_modifier; k.prefiltered = a.prefiltered;
  k.means3D = a.means3D; k.scales = a.scales; k.rotations = a.rotations; k.opacities = a.opacities;
  k.shs = a.shs; k.colors_precomp = a.colors_precomp; k.transMat_precomp = a.transMat_precomp;
  k.refl = a.refl_strengths; k.mask = a.env_scope_mask;
  k.viewmatrix = a.viewmatrix; k.projmatrix = a.projmatrix; k.campos = a.campos;
  k.radii = a.radii; k.gaussian_weights = a.gaussian_weights; k.depths = g.depths; k.tiles_touched = g.tiles_touched; k.clamped = g.clamped;
  k.rects = g.rects; k.rec = g.rec; k.cull = g.cull; k.flags = g.flags;
  StageScope _t(ST_PREPROCESS, s);
  B200GS_CUDA(cudaMemsetAsync(g.flags, 0, 4 * sizeof(int), s));
  surfel_preprocess_kernel<<<(a.P + 255) / 256, 256, 0, s>>>(k);
  B200GS_LAUNCHED("surfel_preprocess_kernel", a.debug, s);
  return B200GS_OK;
}

}  // namespace b200gs
