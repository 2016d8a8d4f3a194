// surface.cu -- the surface attributes the renderer derives from the rasterizer's 8 auxiliary planes
// ("next" row 2, second half): one pixel kernel each way for
// gaussian_renderer/__init__.py:143-183 with utils/point_utils.py:9-37,
//   render_normal  = normalize_eps(allmap[2:5] rotated by world_view_transform[:3,:3].T)       :147-148, :182-183
//   depth_expected = nan_to_num(allmap[0] / clamp(alpha, 1e-3)),  depth_median = nan_to_num(allmap[5])   :151-158
//   surf_depth     = depth_expected (1 - depth_ratio) + depth_ratio depth_median                :170
//   surf_normal    = depth_to_normal(view, surf_depth) * alpha.detach()                         :173-176
// where the reference runs ~25 full-image torch kernels (permutes, matmuls, slices, cross, normalize).
// depth_to_normal is a 4-neighbour stencil on the back-projected points; its backward is written as a
// gather (every pixel re-derives the cross products of its four neighbours), so there are no atomics.
#include "common.cuh"

namespace b200gs {

struct SurfCam {
  float m[9];      // world_view_transform[:3,:3], row major as stored: world normal_i = sum_j view normal_j m[i][j]
  float a[9];      // ray matrix: rays_d = (x, y, 1) @ A,  A = intrins.inverse().T @ c2w[:3,:3].T
  float o[3];      // rays_o = c2w[:3,3]
  float ratio;     // pipe.depth_ratio
};

// torch.nan_to_num(x, 0, 0) as the renderer calls it (gaussian_renderer/__init__.py:152,:158: nan=0, posinf=0, neginf
// left at its default): NaN and +inf become 0, -inf becomes the lowest finite float.
__device__ __forceinline__ float nan_to_zero(float v) {
  if (isnan(v) || v == __int_as_float(0x7f800000)) return 0.0f;
  return v == __int_as_float(0xff800000) ? -3.402823466e+38f : v;
}

__device__ __forceinline__ float surf_depth_of(const float* __restrict__ allmap, size_t N, size_t n, float ratio) {
  const float D = allmap[n], alpha = allmap[N + n], med = allmap[5 * N + n];
  const float e = nan_to_zero(D / fmaxf(alpha, 1e-3f));
  return e * (1.0f - ratio) + ratio * nan_to_zero(med);
}

__device__ __forceinline__ void ray_of(const SurfCam& c, int x, int y, float r[3]) {
  const float px = float(x), py = float(y);
#pragma unroll
  for (int k = 0; k < 3; ++k) r[k] = px * c.a[k] + py * c.a[3 + k] + c.a[6 + k];
}

__device__ __forceinline__ void point_of(const SurfCam& c, const float* __restrict__ depth, int W, int x, int y, float p[3]) {
  float r[3];
  ray_of(c, x, y, r);
  const float d = depth[(size_t)y * W + x];
#pragma unroll
  for (int k = 0; k < 3; ++k) p[k] = d * r[k] + c.o[k];
}

// cross(dx, dy) of pixel (x, y), interior pixels only:  dx = P[y+1][x] - P[y-1][x],  dy = P[y][x+1] - P[y][x-1]
__device__ __forceinline__ void stencil(const SurfCam& c, const float* __restrict__ depth, int W, int x, int y, float dx[3],
                                        float dy[3], float cr[3]) {
  float a[3], b[3], l[3], r[3];
  point_of(c, depth, W, x, y + 1, a); point_of(c, depth, W, x, y - 1, b);
  point_of(c, depth, W, x + 1, y, r); point_of(c, depth, W, x - 1, y, l);
#pragma unroll
  for (int k = 0; k < 3; ++k) { dx[k] = a[k] - b[k]; dy[k] = r[k] - l[k]; }
  cr[0] = dx[1] * dy[2] - dx[2] * dy[1];
  cr[1] = dx[2] * dy[0] - dx[0] * dy[2];
  cr[2] = dx[0] * dy[1] - dx[1] * dy[0];
}

__global__ void __launch_bounds__(256) surface_depth_kernel(const SurfCam c, int N, const float* __restrict__ allmap,
                                                            float* __restrict__ surf_depth) {
  const int n = blockIdx.x * 256 + threadIdx.x;
  if (n < N) surf_depth[n] = surf_depth_of(allmap, (size_t)N, (size_t)n, c.ratio);
}

__global__ void __launch_bounds__(256) surface_fwd_kernel(const SurfCam c, int H, int W, const float* __restrict__ allmap,
                                                          const float* __restrict__ surf_depth, float* __restrict__ out_normal,
                                                          float* __restrict__ out_surf_normal) {
  const int n = blockIdx.x * 256 + threadIdx.x;
  const size_t N = (size_t)H * W;
  if (n >= (int)N) return;
  const int y = n / W, x = n - y * W;
  // rendered normal: view -> world, then n / (|n| + 1e-6)
  const float v0 = allmap[2 * N + n], v1 = allmap[3 * N + n], v2 = allmap[4 * N + n];
  float w[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) w[i] = v0 * c.m[3 * i] + v1 * c.m[3 * i + 1] + v2 * c.m[3 * i + 2];
  const float inv = 1.0f / (sqrtf(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]) + 1e-6f);
  out_normal[3 * (size_t)n] = w[0] * inv; out_normal[3 * (size_t)n + 1] = w[1] * inv; out_normal[3 * (size_t)n + 2] = w[2] * inv;
  // pseudo surface normal from the depth map, times the (detached) accumulated alpha
  float sn[3] = {0.f, 0.f, 0.f};
  if (x > 0 && x < W - 1 && y > 0 && y < H - 1) {
    float dx[3], dy[3], cr[3];
    stencil(c, surf_depth, W, x, y, dx, dy, cr);
    const float s = allmap[N + n] / fmaxf(sqrtf(cr[0] * cr[0] + cr[1] * cr[1] + cr[2] * cr[2]), 1e-12f);   // F.normalize eps
    sn[0] = cr[0] * s; sn[1] = cr[1] * s; sn[2] = cr[2] * s;
  }
  out_surf_normal[n] = sn[0]; out_surf_normal[N + n] = sn[1]; out_surf_normal[2 * N + n] = sn[2];
}

// dL/d(dx), dL/d(dy) of interior pixel (x, y) for the upstream gradient of its surf_normal
__device__ __forceinline__ void stencil_grads(const SurfCam& c, const float* __restrict__ depth, const float* __restrict__ allmap,
                                              const float* __restrict__ g_sn, int H, int W, int x, int y, float gdx[3], float gdy[3]) {
  const size_t N = (size_t)H * W, q = (size_t)y * W + x;
  float dx[3], dy[3], cr[3];
  stencil(c, depth, W, x, y, dx, dy, cr);
  const float len = sqrtf(cr[0] * cr[0] + cr[1] * cr[1] + cr[2] * cr[2]);
  const float alpha = allmap[N + q];
  const float g[3] = {g_sn[q] * alpha, g_sn[N + q] * alpha, g_sn[2 * N + q] * alpha};
  float gc[3];
  if (len > 1e-12f) {               // n = c / |c|:  gc = (g - n (n.g)) / |c|
    const float il = 1.0f / len;
    const float nn[3] = {cr[0] * il, cr[1] * il, cr[2] * il};
    const float ng = nn[0] * g[0] + nn[1] * g[1] + nn[2] * g[2];
#pragma unroll
    for (int k = 0; k < 3; ++k) gc[k] = (g[k] - nn[k] * ng) * il;
  } else {                          // clamped denominator: n = c / eps
#pragma unroll
    for (int k = 0; k < 3; ++k) gc[k] = g[k] * 1e12f;
  }
  gdx[0] = dy[1] * gc[2] - dy[2] * gc[1]; gdx[1] = dy[2] * gc[0] - dy[0] * gc[2]; gdx[2] = dy[0] * gc[1] - dy[1] * gc[0];   // dy x gc
  gdy[0] = gc[1] * dx[2] - gc[2] * dx[1]; gdy[1] = gc[2] * dx[0] - gc[0] * dx[2]; gdy[2] = gc[0] * dx[1] - gc[1] * dx[0];   // gc x dx
}

__global__ void __launch_bounds__(256) surface_bwd_kernel(const SurfCam c, int H, int W, const float* __restrict__ allmap,
                                                          const float* __restrict__ surf_depth, const float* __restrict__ g_normal,
                                                          const float* __restrict__ g_depth, const float* __restrict__ g_sn,
                                                          float* __restrict__ g_allmap) {
  const int n = blockIdx.x * 256 + threadIdx.x;
  const size_t N = (size_t)H * W;
  if (n >= (int)N) return;
  const int y = n / W, x = n - y * W;
  // ---- rendered normal
  float gv[3] = {0.f, 0.f, 0.f};
  if (g_normal) {
    const float v0 = allmap[2 * N + n], v1 = allmap[3 * N + n], v2 = allmap[4 * N + n];
    float w[3], g[3], gw[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) { w[i] = v0 * c.m[3 * i] + v1 * c.m[3 * i + 1] + v2 * c.m[3 * i + 2]; g[i] = g_normal[3 * (size_t)n + i]; }
    const float L = sqrtf(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]), inv = 1.0f / (L + 1e-6f);
    const float wg = w[0] * g[0] + w[1] * g[1] + w[2] * g[2];
    const float k2 = L > 0.0f ? wg * inv * inv / L : 0.0f;       // d(1 / (L + eps)) = -dL / (L + eps)^2, dL = w.dw / L
#pragma unroll
    for (int i = 0; i < 3; ++i) gw[i] = g[i] * inv - w[i] * k2;
#pragma unroll
    for (int j = 0; j < 3; ++j) gv[j] = gw[0] * c.m[j] + gw[1] * c.m[3 + j] + gw[2] * c.m[6 + j];
  }
  g_allmap[2 * N + n] = gv[0]; g_allmap[3 * N + n] = gv[1]; g_allmap[4 * N + n] = gv[2];
  // ---- depth: direct gradient + the four neighbours' normals this pixel's point enters
  float gd = g_depth ? g_depth[n] : 0.0f;
  if (g_sn) {
    float gp[3] = {0.f, 0.f, 0.f}, a[3], b[3];
    auto interior = [&](int xx, int yy) { return xx > 0 && xx < W - 1 && yy > 0 && yy < H - 1; };
    if (interior(x, y - 1)) { stencil_grads(c, surf_depth, allmap, g_sn, H, W, x, y - 1, a, b); gp[0] += a[0]; gp[1] += a[1]; gp[2] += a[2]; }   // P[y'+1]
    if (interior(x, y + 1)) { stencil_grads(c, surf_depth, allmap, g_sn, H, W, x, y + 1, a, b); gp[0] -= a[0]; gp[1] -= a[1]; gp[2] -= a[2]; }   // P[y'-1]
    if (interior(x - 1, y)) { stencil_grads(c, surf_depth, allmap, g_sn, H, W, x - 1, y, a, b); gp[0] += b[0]; gp[1] += b[1]; gp[2] += b[2]; }   // P[x'+1]
    if (interior(x + 1, y)) { stencil_grads(c, surf_depth, allmap, g_sn, H, W, x + 1, y, a, b); gp[0] -= b[0]; gp[1] -= b[1]; gp[2] -= b[2]; }   // P[x'-1]
    float r[3];
    ray_of(c, x, y, r);
    gd += gp[0] * r[0] + gp[1] * r[1] + gp[2] * r[2];
  }
  const float D = allmap[n], alpha = allmap[N + n], med = allmap[5 * N + n];
  const float ca = fmaxf(alpha, 1e-3f);
  const float e = D / ca;
  const float ge = (isnan(e) || isinf(e)) ? 0.0f : gd * (1.0f - c.ratio);           // nan_to_num passes finite values only
  g_allmap[n] = ge / ca;
  g_allmap[N + n] = alpha >= 1e-3f ? -ge * D / (ca * ca) : 0.0f;                     // clamp(min): gradient where alpha >= min
  g_allmap[5 * N + n] = (isnan(med) || isinf(med)) ? 0.0f : gd * c.ratio;
  g_allmap[6 * N + n] = 0.0f;
  g_allmap[7 * N + n] = 0.0f;
}

static int surf_cam(const float* view3x3, const float* raymat, const float* rays_o, float ratio, SurfCam& c) {
  if (!view3x3 || !raymat || !rays_o) return 1;
  for (int i = 0; i < 9; ++i) { c.m[i] = view3x3[i]; c.a[i] = raymat[i]; }
  for (int i = 0; i < 3; ++i) c.o[i] = rays_o[i];
  c.ratio = ratio;
  return 0;
}

}  // namespace b200gs

using namespace b200gs;

extern "C" {

int b200gs_surface_forward(int H, int W, const float* allmap, const float* view3x3, const float* raymat, const float* rays_o,
                           float depth_ratio, float* out_normal, float* out_surf_depth, float* out_surf_normal, void* stream) {
  if (H < 0 || W < 0) { set_error("surface_forward: negative size"); return B200GS_EINVAL; }
  if (H == 0 || W == 0) return B200GS_OK;
  SurfCam c;
  if (surf_cam(view3x3, raymat, rays_o, depth_ratio, c) || !allmap || !out_normal || !out_surf_depth || !out_surf_normal) {
    set_error("surface_forward: bad argument"); return B200GS_EINVAL; }
  if ((long long)H * W >= 0x7fffffffll / 8) { set_error("surface_forward: image too large"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  StageScope _t(ST_SURFACE_FWD, s);
  const int N = H * W;
  surface_depth_kernel<<<(N + 255) / 256, 256, 0, s>>>(c, N, allmap, out_surf_depth);
  B200GS_LAUNCHED("surface_depth_kernel", false, s);
  surface_fwd_kernel<<<(N + 255) / 256, 256, 0, s>>>(c, H, W, allmap, out_surf_depth, out_normal, out_surf_normal);
  B200GS_LAUNCHED("surface_fwd_kernel", false, s);
  return B200GS_OK;
}

int b200gs_surface_backward(int H, int W, const float* allmap, const float* view3x3, const float* raymat, const float* rays_o,
                            float depth_ratio, const float* surf_depth, const float* dL_dnormal, const float* dL_dsurf_depth,
                            const float* dL_dsurf_normal, float* dL_dallmap, void* stream) {
  if (H < 0 || W < 0) { set_error("surface_backward: negative size"); return B200GS_EINVAL; }
  if (H == 0 || W == 0) return B200GS_OK;
  SurfCam c;
  if (surf_cam(view3x3, raymat, rays_o, depth_ratio, c) || !allmap || !surf_depth || !dL_dallmap) {
    set_error("surface_backward: bad argument"); return B200GS_EINVAL; }
  if ((long long)H * W >= 0x7fffffffll / 8) { set_error("surface_backward: image too large"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  StageScope _t(ST_SURFACE_BWD, s);
  const int N = H * W;
  surface_bwd_kernel<<<(N + 255) / 256, 256, 0, s>>>(c, H, W, allmap, surf_depth, dL_dnormal, dL_dsurf_depth, dL_dsurf_normal, dL_dallmap);
  B200GS_LAUNCHED("surface_bwd_kernel", false, s);
  return B200GS_OK;
}

}  // extern "C"
