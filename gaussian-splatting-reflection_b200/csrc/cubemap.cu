// cubemap.cu -- K6: filtered cubemap environment lookup and its gradient.
//
// Restates C/src/cubemapencoder.cu (C = submodules/cubemapencoder of the reference):
//   Compute_Cubemap_UV :147-187, Compute_Seamless_Index :189-263, EdgeTable :66-106
//   (the LEFT_TOP_AS_ORIGIN branch, :20), the three forward kernels :297-427, the three
//   backward kernels :509-710 and Compute_Cubemap_UV_Backward :265-292.
//
// This file is compiled with --use_fast_math, like the reference extension
// (C/setup.py:6-11), so divisions / floor land on the same approximate
// instructions and texel selection agrees with the reference wherever possible.
//
// B200 design: one thread per ray; the 1.2 MB texture lives in L2.  A single
// `Taps` descriptor (4 texel offsets + 4 weights) covers nearest, in-face bilinear,
// seamless edge and seamless vertex cases, so forward and backward share one
// code path.  The cross-face neighbour rules are a 24-entry table in constant
// memory instead of a branch ladder.  Backward: rays of a warp (neighbouring
// pixels) mostly hit the same few texels, so texel gradients are first combined
// across the warp (match.any + log-step peer reduction, all channels at once)
// and only one RED.ADD per distinct texel and channel goes to L2.
#include "common.cuh"

namespace b200gs {

// Neighbour-face rule for leaving `face` through an edge.  edge: 0 = u<0 (flag 1),
// 1 = u>=L (flag 2), 2 = v<0 (flag 4), 3 = v>=L (flag 8).  x/y source codes:
// 0 -> 0, 1 -> L-1, 2 -> x, 3 -> L-1-x, 4 -> y, 5 -> L-1-y.
struct EdgeRule { unsigned char face, xs, ys; };
__constant__ EdgeRule c_edge[6][4] = {
    {{4, 1, 4}, {5, 0, 4}, {3, 1, 2}, {2, 1, 2}},
    {{5, 1, 4}, {4, 0, 4}, {3, 0, 3}, {2, 0, 3}},
    {{1, 5, 1}, {0, 4, 1}, {4, 2, 1}, {5, 3, 1}},
    {{1, 5, 0}, {0, 4, 0}, {4, 2, 0}, {5, 3, 0}},
    {{1, 1, 4}, {0, 0, 4}, {3, 2, 0}, {2, 2, 0}},
    {{0, 1, 4}, {1, 0, 4}, {3, 3, 1}, {2, 3, 1}},
};

__device__ __forceinline__ int edge_src(int code, int x, int y, int L) {
  switch (code) {
    case 0: return 0;
    case 1: return L - 1;
    case 2: return x;
    case 3: return L - 1 - x;
    case 4: return y;
    default: return L - 1 - y;
  }
}

struct Taps {
  int off[4];      // texel offset of channel 0: (face*C*L + y)*L + x ; channel stride L*L
  float kx, ky;
  int face, flag;
  bool vertex;     // 3 taps, the 4th value is their mean
  float u, v;      // face coordinates before the texel mapping (for the direction gradient)
};

// Face selection and (u,v) (cubemapencoder.cu:147-187).
__device__ __forceinline__ void face_uv(float x, float y, float z, int& face, float& u, float& v) {
  int dim = 0;
  float mv = fabsf(x);
  if (fabsf(y) > mv) { mv = fabsf(y); dim = 1; }
  if (fabsf(z) > mv) { mv = fabsf(z); dim = 2; }
  if (dim == 0) {
    u = z / x; v = y / x;
    if (x >= 0.f) { face = 0; u = -u; v = -v; } else { face = 1; u = -u; }
  } else if (dim == 1) {
    u = x / y; v = z / y;
    if (y >= 0.f) { face = 2; } else { face = 3; u = -u; v = -v; }
  } else {
    u = x / z; v = y / z;
    if (z >= 0.f) { face = 4; v = -v; } else { face = 5; }
  }
}

__device__ __forceinline__ int tex_off(int face, int x, int y, int C, int L) { return (face * C * L + y) * L + x; }

__device__ __forceinline__ int neighbour(int face, int edge, int x, int y, int C, int L) {
  const EdgeRule r = c_edge[face][edge];
  return tex_off(r.face, edge_src(r.xs, x, y, L), edge_src(r.ys, x, y, L), C, L);
}

// mode 0 nearest, 1 bilinear in-face, 2 bilinear seamless.
template <int MODE>
__device__ __forceinline__ Taps make_taps(float dx, float dy, float dz, int C, int L) {
  Taps t;
  face_uv(dx, dy, dz, t.face, t.u, t.v);
  t.flag = 0; t.vertex = false;
  const float fu = (t.u * 0.5f + 0.5f) * float(L);
  const float fv = (-t.v * 0.5f + 0.5f) * float(L);       // LEFT_TOP_AS_ORIGIN: v -> -v (:203)
  if (MODE == 0) {                                         // :411-421
    int ux = int(fu), uy = int(fv);
    ux = imin(imax(ux, 0), L - 1); uy = imin(imax(uy, 0), L - 1);
    t.off[0] = t.off[1] = t.off[2] = t.off[3] = tex_off(t.face, ux, uy, C, L);
    t.kx = 0.f; t.ky = 0.f;
    return t;
  }
  int x0 = int(floorf(fu - 0.5f)), y0 = int(floorf(fv - 0.5f));
  int x1 = x0 + 1, y1 = y0 + 1;
  t.kx = fu - float(x0) - 0.5f;
  t.ky = fv - float(y0) - 0.5f;
  x0 = imin(imax(x0, 0), L - 1); x1 = imin(imax(x1, 0), L - 1);
  y0 = imin(imax(y0, 0), L - 1); y1 = imin(imax(y1, 0), L - 1);
  if (MODE == 2) {                                         // :221-262
    int eu = -1, ev = -1;
    if (fu < 0.5f) { t.flag |= 1; t.kx = 0.5f - fu; eu = 0; }
    else if (fu >= float(L) - 0.5f) { t.flag |= 2; eu = 1; }
    if (fv < 0.5f) { t.flag |= 4; t.ky = 0.5f - fv; ev = 2; }
    else if (fv >= float(L) - 0.5f) { t.flag |= 8; ev = 3; }
    if (eu >= 0 && ev >= 0) {
      t.vertex = true;
      t.off[0] = tex_off(t.face, x0, y0, C, L);
      t.off[1] = neighbour(t.face, eu, x0, y0, C, L);
      t.off[2] = neighbour(t.face, ev, x0, y0, C, L);
      t.off[3] = t.off[0];
      return t;
    }
    if (eu >= 0) {
      t.off[0] = tex_off(t.face, x0, y0, C, L);
      t.off[1] = neighbour(t.face, eu, x0, y0, C, L);
      t.off[2] = tex_off(t.face, x0, y1, C, L);
      t.off[3] = neighbour(t.face, eu, x0, y1, C, L);
      return t;
    }
    if (ev >= 0) {
      t.off[0] = tex_off(t.face, x0, y0, C, L);
      t.off[1] = tex_off(t.face, x1, y0, C, L);
      t.off[2] = neighbour(t.face, ev, x0, y0, C, L);
      t.off[3] = neighbour(t.face, ev, x1, y0, C, L);
      return t;
    }
  }
  t.off[0] = tex_off(t.face, x0, y0, C, L);
  t.off[1] = tex_off(t.face, x1, y0, C, L);
  t.off[2] = tex_off(t.face, x0, y1, C, L);
  t.off[3] = tex_off(t.face, x1, y1, C, L);
  return t;
}

template <int MODE>
__global__ void __launch_bounds__(256) cubemap_fwd_kernel(const float* __restrict__ inputs, const float* __restrict__ cubemap,
                                                          const float* __restrict__ fail, float* __restrict__ outputs,
                                                          unsigned B, int C, int L) {
  const unsigned n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= B) return;
  const float dx = inputs[3 * (size_t)n], dy = inputs[3 * (size_t)n + 1], dz = inputs[3 * (size_t)n + 2];
  if (dx == 0.f && dy == 0.f && dz == 0.f) {              // :311-316
    for (int c = 0; c < C; ++c) outputs[(size_t)c * B + n] = fail[c];
    return;
  }
  const Taps t = make_taps<MODE>(dx, dy, dz, C, L);
  const int cs = L * L;
  for (int c = 0; c < C; ++c) {
    const float* tex = cubemap + (size_t)c * cs;
    if (MODE == 0) { outputs[(size_t)c * B + n] = tex[t.off[0]]; continue; }
    const float v00 = tex[t.off[0]], v01 = tex[t.off[1]], v10 = tex[t.off[2]];
    const float v11 = t.vertex ? (v00 + v01 + v10) / 3.f : tex[t.off[3]];
    outputs[(size_t)c * B + n] = (1 - t.ky) * ((1 - t.kx) * v00 + t.kx * v01) + t.ky * ((1 - t.kx) * v10 + t.kx * v11);
  }
}

// ------------------------------------------------------------------ backward
// Sum `v[0..NC)` over the lanes of `peers` (all holding the same texel); the result is valid in
// the lowest lane of the group.  Log-step: every round each still-active lane pulls the values
// of the next active peer above it, then every second active lane drops out.
template <int NC>
__device__ __forceinline__ void peer_reduce(unsigned active, unsigned peers, float* v) {
  const int lane = threadIdx.x & 31;
  int rel = __popc(peers & ((1u << lane) - 1u));        // my rank inside the group
  unsigned above = peers & (0xfffffffeu << lane);       // group members above me
  while (__any_sync(active, above != 0u)) {
    const int next = __ffs(above) - 1;                  // -1: nothing to pull
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const float o = __shfl_sync(active, v[c], next < 0 ? lane : next);
      if (next >= 0) v[c] += o;
    }
    const unsigned keep = __ballot_sync(active, (rel & 1) == 0);   // even ranks stay
    above &= keep;
    rel >>= 1;
  }
}

// d(u,v) -> d(direction) (cubemapencoder.cu:265-292); gu, gv are dL/du, dL/dv of the face coordinates.
__device__ __forceinline__ void uv_backward(int face, float x, float y, float z, float gu, float gv, float g[3]) {
  if (face < 2) {
    if (face == 0) { gu = -gu; gv = -gv; } else { gu = -gu; }
    g[0] = -(z * gu + y * gv) / (x * x); g[1] = 1.f / x * gv; g[2] = 1.f / x * gu;
  } else if (face < 4) {
    if (face == 3) { gu = -gu; gv = -gv; }
    g[0] = 1.f / y * gu; g[1] = -(x * gu + z * gv) / (y * y); g[2] = 1.f / y * gv;
  } else {
    if (face == 4) { gv = -gv; }
    g[0] = 1.f / z * gu; g[1] = 1.f / z * gv; g[2] = -(x * gu + y * gv) / (z * z);
  }
}

// NC > 0: channel count known at compile time (warp-aggregated scatter); NC == 0: generic C, plain REDs.
template <int MODE, int NC>
__global__ void __launch_bounds__(256) cubemap_bwd_kernel(const float* __restrict__ grad_out, const float* __restrict__ cubemap,
                                                          const float* __restrict__ inputs, float* __restrict__ grad_cubemap,
                                                          float* __restrict__ grad_inputs, float* __restrict__ grad_fail,
                                                          unsigned B, int Crt, int L) {
  const int C = NC > 0 ? NC : Crt;
  const unsigned n = blockIdx.x * blockDim.x + threadIdx.x;
  const bool in_range = n < B;
  float dx = 0.f, dy = 0.f, dz = 0.f;
  if (in_range) { dx = inputs[3 * (size_t)n]; dy = inputs[3 * (size_t)n + 1]; dz = inputs[3 * (size_t)n + 2]; }
  const bool zero = dx == 0.f && dy == 0.f && dz == 0.f;
  if (in_range && zero) {                                  // :525-533
    for (int c = 0; c < C; ++c) atomicAdd(grad_fail + c, grad_out[(size_t)c * B + n]);
    if (MODE != 0) { grad_inputs[3 * (size_t)n] = 0.f; grad_inputs[3 * (size_t)n + 1] = 0.f; grad_inputs[3 * (size_t)n + 2] = 0.f; }
  }
  const bool live = in_range && !zero;
  const int cs = L * L;
  Taps t;
  t.off[0] = t.off[1] = t.off[2] = t.off[3] = -1;
  t.kx = t.ky = 0.f; t.vertex = false; t.flag = 0; t.face = 0;
  if (live) t = make_taps<MODE>(dx, dy, dz, C, L);

  // texel weights (:546-558; vertex: kx*ky/3 extra on each of the three taps)
  float w[4];
  if (MODE == 0) { w[0] = 1.f; w[1] = w[2] = w[3] = 0.f; }
  else {
    const float extra = t.vertex ? t.ky * t.kx / 3.f : 0.f;
    w[0] = (1 - t.ky) * (1 - t.kx) + extra;
    w[1] = (1 - t.ky) * t.kx + extra;
    w[2] = t.ky * (1 - t.kx) + extra;
    w[3] = t.vertex ? 0.f : t.ky * t.kx;
  }
  const int ntaps = MODE == 0 ? 1 : 4;

  if (NC > 0) {
    float go[NC > 0 ? NC : 1];
#pragma unroll
    for (int c = 0; c < NC; ++c) go[c] = live ? grad_out[(size_t)c * B + n] : 0.f;
    const unsigned active = __activemask();                 // whole warp: no early returns above
    for (int k = 0; k < ntaps; ++k) {
      const bool has = live && !(k == 3 && t.vertex);
      const int key = has ? t.off[k] : -1 - (int)(threadIdx.x & 31);   // unique key for idle lanes
      const unsigned peers = __match_any_sync(active, key);
      float v[NC > 0 ? NC : 1];
#pragma unroll
      for (int c = 0; c < NC; ++c) v[c] = w[k] * go[c];
      if (__any_sync(active, peers != (1u << (threadIdx.x & 31)))) peer_reduce<(NC > 0 ? NC : 1)>(active, peers, v);
      if (has && (peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0u) {     // group leader
#pragma unroll
        for (int c = 0; c < NC; ++c) atomicAdd(grad_cubemap + (size_t)c * cs + key, v[c]);
      }
    }
  } else if (live) {
    for (int c = 0; c < C; ++c) {
      const float g = grad_out[(size_t)c * B + n];
      for (int k = 0; k < ntaps; ++k)
        if (!(k == 3 && t.vertex)) atomicAdd(grad_cubemap + (size_t)c * cs + t.off[k], w[k] * g);
    }
  }

  if (MODE == 0 || !live) return;                           // nearest has no direction gradient (:749-751)
  float gsum[3] = {0.f, 0.f, 0.f};
  for (int c = 0; c < C; ++c) {                             // :560-578
    const float* tex = cubemap + (size_t)c * cs;
    const float g = grad_out[(size_t)c * B + n];
    const float v00 = tex[t.off[0]], v01 = tex[t.off[1]], v10 = tex[t.off[2]];
    const float v11 = t.vertex ? (v00 + v01 + v10) / 3.f : tex[t.off[3]];
    float gu = (1 - t.ky) * (v01 - v00) + t.ky * (v11 - v10);
    float gv = (1 - t.kx) * (v10 - v00) + t.kx * (v11 - v01);
    gu *= 0.5f * float(L) * g;
    gv *= 0.5f * float(L) * g;
    if (t.flag & 1) gu = -gu;
    if (t.flag & 4) gv = -gv;
    gv = -gv;                                               // LEFT_TOP_AS_ORIGIN
    float gd[3];
    uv_backward(t.face, dx, dy, dz, gu, gv, gd);
    gsum[0] += gd[0]; gsum[1] += gd[1]; gsum[2] += gd[2];
  }
  grad_inputs[3 * (size_t)n] = gsum[0]; grad_inputs[3 * (size_t)n + 1] = gsum[1]; grad_inputs[3 * (size_t)n + 2] = gsum[2];
}

// ------------------------------------------------------------------ fused reflection shading
// SURVEY.md section 8f, "next" row 2 (first half): the reflection pass the fork adds around the
// rasterizer, gaussian_renderer/__init__.py:22-35 and :199-201, as ONE pixel kernel each way:
//   rays_d = sample_camera_rays(HWK, R, T)            utils/general_utils.py:177-197
//   refl   = rays_d - 2 n (rays_d . n)                 reflection(), gaussian_renderer/__init__.py:22-24
//   env    = sigmoid(env_map(refl))                    sample_cubemap_color(), :26-30 (cubemap lookup above)
//   final  = (1 - r) base + r env                      :201
// where the reference runs ~12 full-image torch kernels and the [C,B] cubemap kernel.  Three channels.
struct ShadeCam {
  float kinv[9];   // inverse intrinsics, row major: pixel_camera = Kinv (x, y, 1)
  float rot[9];    // the camera's stored R (3DGS keeps it transposed), row major: used as R p
  float t[3];
};

__device__ __forceinline__ void shade_ray(const ShadeCam& c, int x, int y, float d[3]) {
  const float px = float(x), py = float(y);
  // pixel_camera = xy1 @ inv(K).T ; pixel_world = (pixel_camera - T) @ R.T' ; rays_o = -R'.T @ T  (R' = R.T)
  const float pc[3] = {c.kinv[0] * px + c.kinv[1] * py + c.kinv[2], c.kinv[3] * px + c.kinv[4] * py + c.kinv[5],
                       c.kinv[6] * px + c.kinv[7] * py + c.kinv[8]};
  const float q[3] = {pc[0] - c.t[0], pc[1] - c.t[1], pc[2] - c.t[2]};
  float w[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const float pw = c.rot[3 * i] * q[0] + c.rot[3 * i + 1] * q[1] + c.rot[3 * i + 2] * q[2];
    const float ro = -(c.rot[3 * i] * c.t[0] + c.rot[3 * i + 1] * c.t[1] + c.rot[3 * i + 2] * c.t[2]);
    w[i] = pw - ro;
  }
  // rays_d / torch.norm(rays_d): an IEEE square root and three IEEE divisions, as the torch lines do it (this file is built
  // with --use_fast_math for the lookup's sake; an approximate rsqrt here moves ten times as many rays across a texel
  // boundary of the lookup, where the direction gradient jumps)
  const float nn = __fsqrt_rn(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  d[0] = __fdiv_rn(w[0], nn); d[1] = __fdiv_rn(w[1], nn); d[2] = __fdiv_rn(w[2], nn);
}

template <int MODE>
__global__ void __launch_bounds__(256) reflect_shade_fwd_kernel(const ShadeCam cam, int H, int W, const float* __restrict__ normal,
                                                                const float* __restrict__ base, const float* __restrict__ refl,
                                                                const float* __restrict__ cubemap, const float* __restrict__ fail,
                                                                int L, float* __restrict__ out_final, float* __restrict__ out_env) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  const int N = H * W;
  if (n >= N) return;
  const int y = n / W, x = n - y * W;
  float d[3];
  shade_ray(cam, x, y, d);
  const float nx = normal[3 * (size_t)n], ny = normal[3 * (size_t)n + 1], nz = normal[3 * (size_t)n + 2];
  const float dn = d[0] * nx + d[1] * ny + d[2] * nz;
  const float rx = d[0] - 2.f * nx * dn, ry = d[1] - 2.f * ny * dn, rz = d[2] - 2.f * nz * dn;
  float e[3];
  if (rx == 0.f && ry == 0.f && rz == 0.f) {
    e[0] = fail[0]; e[1] = fail[1]; e[2] = fail[2];
  } else {
    const Taps t = make_taps<MODE>(rx, ry, rz, 3, L);
    const int cs = L * L;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float* tex = cubemap + (size_t)c * cs;
      if (MODE == 0) { e[c] = tex[t.off[0]]; continue; }
      const float v00 = tex[t.off[0]], v01 = tex[t.off[1]], v10 = tex[t.off[2]];
      const float v11 = t.vertex ? (v00 + v01 + v10) / 3.f : tex[t.off[3]];
      e[c] = (1 - t.ky) * ((1 - t.kx) * v00 + t.kx * v01) + t.ky * ((1 - t.kx) * v10 + t.kx * v11);
    }
  }
  const float r = refl[n];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float env = 1.0f / (1.0f + expf(-e[c]));
    out_env[(size_t)c * N + n] = env;
    out_final[(size_t)c * N + n] = (1.0f - r) * base[(size_t)c * N + n] + r * env;
  }
}

template <int MODE>
__global__ void __launch_bounds__(256) reflect_shade_bwd_kernel(const ShadeCam cam, int H, int W, const float* __restrict__ normal,
                                                                const float* __restrict__ base, const float* __restrict__ refl,
                                                                const float* __restrict__ cubemap, int L,
                                                                const float* __restrict__ env, const float* __restrict__ g_final,
                                                                const float* __restrict__ g_env, float* __restrict__ g_base,
                                                                float* __restrict__ g_refl, float* __restrict__ g_normal,
                                                                float* __restrict__ grad_cubemap, float* __restrict__ grad_fail) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  const int N = H * W;
  const bool in_range = n < N;
  float d[3] = {0.f, 0.f, 1.f}, nrm[3] = {0.f, 0.f, 0.f}, go[3] = {0.f, 0.f, 0.f};
  float dn = 0.f, rx = 0.f, ry = 0.f, rz = 0.f;
  if (in_range) {
    const int y = n / W, x = n - y * W;
    shade_ray(cam, x, y, d);
    nrm[0] = normal[3 * (size_t)n]; nrm[1] = normal[3 * (size_t)n + 1]; nrm[2] = normal[3 * (size_t)n + 2];
    dn = d[0] * nrm[0] + d[1] * nrm[1] + d[2] * nrm[2];
    rx = d[0] - 2.f * nrm[0] * dn; ry = d[1] - 2.f * nrm[1] * dn; rz = d[2] - 2.f * nrm[2] * dn;
    const float r = refl[n];
    float gr = 0.f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float gf = g_final[(size_t)c * N + n], ev = env[(size_t)c * N + n], b = base[(size_t)c * N + n];
      g_base[(size_t)c * N + n] = (1.0f - r) * gf;
      gr += gf * (ev - b);
      const float ge = r * gf + (g_env ? g_env[(size_t)c * N + n] : 0.f);
      go[c] = ge * ev * (1.0f - ev);                         // through the sigmoid, down to the lookup's output
    }
    g_refl[n] = gr;
  }
  const bool zero = rx == 0.f && ry == 0.f && rz == 0.f;
  if (in_range && zero) {
#pragma unroll
    for (int c = 0; c < 3; ++c) atomicAdd(grad_fail + c, go[c]);
    g_normal[3 * (size_t)n] = 0.f; g_normal[3 * (size_t)n + 1] = 0.f; g_normal[3 * (size_t)n + 2] = 0.f;
  }
  const bool live = in_range && !zero;
  const int cs = L * L;
  Taps t;
  t.off[0] = t.off[1] = t.off[2] = t.off[3] = -1;
  t.kx = t.ky = 0.f; t.vertex = false; t.flag = 0; t.face = 0;
  if (live) t = make_taps<MODE>(rx, ry, rz, 3, L);
  float w[4];
  if (MODE == 0) { w[0] = 1.f; w[1] = w[2] = w[3] = 0.f; }
  else {
    const float extra = t.vertex ? t.ky * t.kx / 3.f : 0.f;
    w[0] = (1 - t.ky) * (1 - t.kx) + extra;
    w[1] = (1 - t.ky) * t.kx + extra;
    w[2] = t.ky * (1 - t.kx) + extra;
    w[3] = t.vertex ? 0.f : t.ky * t.kx;
  }
  const int ntaps = MODE == 0 ? 1 : 4;
  const unsigned active = __activemask();                   // whole warp: no early returns above
  for (int k = 0; k < ntaps; ++k) {                         // warp-aggregated texture scatter, as in cubemap_bwd_kernel
    const bool has = live && !(k == 3 && t.vertex);
    const int key = has ? t.off[k] : -1 - (int)(threadIdx.x & 31);
    const unsigned peers = __match_any_sync(active, key);
    float v[3] = {w[k] * go[0], w[k] * go[1], w[k] * go[2]};
    if (__any_sync(active, peers != (1u << (threadIdx.x & 31)))) peer_reduce<3>(active, peers, v);
    if (has && (peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0u) {
#pragma unroll
      for (int c = 0; c < 3; ++c) atomicAdd(grad_cubemap + (size_t)c * cs + key, v[c]);
    }
  }
  if (!live) return;
  float gd[3] = {0.f, 0.f, 0.f};                            // dL/d(reflected direction)
  if (MODE != 0) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float* tex = cubemap + (size_t)c * cs;
      const float v00 = tex[t.off[0]], v01 = tex[t.off[1]], v10 = tex[t.off[2]];
      const float v11 = t.vertex ? (v00 + v01 + v10) / 3.f : tex[t.off[3]];
      float gu = (1 - t.ky) * (v01 - v00) + t.ky * (v11 - v10);
      float gv = (1 - t.kx) * (v10 - v00) + t.kx * (v11 - v01);
      gu *= 0.5f * float(L) * go[c];
      gv *= 0.5f * float(L) * go[c];
      if (t.flag & 1) gu = -gu;
      if (t.flag & 4) gv = -gv;
      gv = -gv;
      float g3[3];
      uv_backward(t.face, rx, ry, rz, gu, gv, g3);
      gd[0] += g3[0]; gd[1] += g3[1]; gd[2] += g3[2];
    }
  }
  // refl = d - 2 n (d.n):  dL/dn = -2 [ (d.n) gd + (gd.n) d ]
  const float gn = gd[0] * nrm[0] + gd[1] * nrm[1] + gd[2] * nrm[2];
#pragma unroll
  for (int i = 0; i < 3; ++i) g_normal[3 * (size_t)n + i] = -2.f * (dn * gd[i] + gn * d[i]);
}

}  // namespace b200gs

using namespace b200gs;

template <int MODE>
static int launch_bwd(const float* go, const float* cm, const float* in, float* gcm, float* gin, float* gf, unsigned B, int C,
                      int L, cudaStream_t s) {
  const unsigned blocks = (B + 255u) / 256u;
  switch (C) {
    case 1: cubemap_bwd_kernel<MODE, 1><<<blocks, 256, 0, s>>>(go, cm, in, gcm, gin, gf, B, C, L); break;
    case 2: cubemap_bwd_kernel<MODE, 2><<<blocks, 256, 0, s>>>(go, cm, in, gcm, gin, gf, B, C, L); break;
    case 3: cubemap_bwd_kernel<MODE, 3><<<blocks, 256, 0, s>>>(go, cm, in, gcm, gin, gf, B, C, L); break;
    case 4: cubemap_bwd_kernel<MODE, 4><<<blocks, 256, 0, s>>>(go, cm, in, gcm, gin, gf, B, C, L); break;
    case 6: cubemap_bwd_kernel<MODE, 6><<<blocks, 256, 0, s>>>(go, cm, in, gcm, gin, gf, B, C, L); break;
    default: cubemap_bwd_kernel<MODE, 0><<<blocks, 256, 0, s>>>(go, cm, in, gcm, gin, gf, B, C, L); break;
  }
  B200GS_LAUNCHED("cubemap_bwd_kernel", false, s);
  return B200GS_OK;
}

static int cubemap_mode(uint32_t interp, uint32_t seamless) { return interp == 0 ? 0 : (seamless == 0 ? 1 : 2); }

extern "C" {

int b200gs_cubemap_forward(const float* inputs, const float* cubemap, const float* fail_value, float* outputs,
                           uint32_t interp, uint32_t seamless, uint32_t B, uint32_t C, uint32_t L, void* stream) {
  if (B == 0) return B200GS_OK;
  if (!inputs || !cubemap || !fail_value || !outputs || C == 0 || L == 0) { set_error("cubemap_forward: bad argument"); return B200GS_EINVAL; }
  if ((uint64_t)6 * C * L * L >= 0x7fffffffull) { set_error("cubemap_forward: texture too large for 32-bit texel offsets"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  StageScope _t(ST_CUBEMAP_FWD, s);
  const unsigned blocks = (B + 255u) / 256u;
  switch (cubemap_mode(interp, seamless)) {
    case 0: cubemap_fwd_kernel<0><<<blocks, 256, 0, s>>>(inputs, cubemap, fail_value, outputs, B, (int)C, (int)L); break;
    case 1: cubemap_fwd_kernel<1><<<blocks, 256, 0, s>>>(inputs, cubemap, fail_value, outputs, B, (int)C, (int)L); break;
    default: cubemap_fwd_kernel<2><<<blocks, 256, 0, s>>>(inputs, cubemap, fail_value, outputs, B, (int)C, (int)L); break;
  }
  B200GS_LAUNCHED("cubemap_fwd_kernel", false, s);
  return B200GS_OK;
}

int b200gs_cubemap_backward(const float* grad_outputs, const float* inputs, const float* cubemap, float* grad_cubemap,
                            float* grad_inputs, float* grad_fail, uint32_t interp, uint32_t seamless, uint32_t B,
                            uint32_t C, uint32_t L, void* stream) {
  if (B == 0) return B200GS_OK;
  if (!grad_outputs || !inputs || !cubemap || !grad_cubemap || !grad_inputs || !grad_fail || C == 0 || L == 0) {
    set_error("cubemap_backward: bad argument"); return B200GS_EINVAL; }
  if ((uint64_t)6 * C * L * L >= 0x7fffffffull) { set_error("cubemap_backward: texture too large for 32-bit texel offsets"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  StageScope _t(ST_CUBEMAP_BWD, s);
  const int mode = cubemap_mode(interp, seamless);
  if (mode == 0) {  // "NEAREST has no grad" (cubemapencoder.cu:749-751)
    B200GS_CUDA(cudaMemsetAsync(grad_inputs, 0, (size_t)B * 3 * sizeof(float), s));
    return launch_bwd<0>(grad_outputs, cubemap, inputs, grad_cubemap, grad_inputs, grad_fail, B, (int)C, (int)L, s);
  }
  if (mode == 1) return launch_bwd<1>(grad_outputs, cubemap, inputs, grad_cubemap, grad_inputs, grad_fail, B, (int)C, (int)L, s);
  return launch_bwd<2>(grad_outputs, cubemap, inputs, grad_cubemap, grad_inputs, grad_fail, B, (int)C, (int)L, s);
}


static int shade_cam(const float* Kinv, const float* R, const float* T, ShadeCam& c) {
  if (!Kinv || !R || !T) return 1;
  for (int i = 0; i < 9; ++i) { c.kinv[i] = Kinv[i]; c.rot[i] = R[i]; }
  for (int i = 0; i < 3; ++i) c.t[i] = T[i];
  return 0;
}

int b200gs_reflect_shade_forward(int H, int W, const float* Kinv, const float* R, const float* T, const float* normal,
                                 const float* base_color, const float* refl_strength, const float* cubemap,
                                 const float* fail_value, uint32_t L, uint32_t interp, uint32_t seamless, float* out_final,
                                 float* out_env, void* stream) {
  if (H < 0 || W < 0) { set_error("reflect_shade_forward: negative size"); return B200GS_EINVAL; }
  if (H == 0 || W == 0) return B200GS_OK;
  ShadeCam cam;
  if (shade_cam(Kinv, R, T, cam) || !normal || !base_color || !refl_strength || !cubemap || !fail_value || !out_final || !out_env || L == 0) {
    set_error("reflect_shade_forward: bad argument"); return B200GS_EINVAL; }
  if ((uint64_t)6 * 3 * L * L >= 0x7fffffffull || (long long)H * W >= 0x7fffffffll) { set_error("reflect_shade_forward: too large"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  StageScope _t(ST_SHADE_FWD, s);
  const unsigned blocks = (unsigned)(((long long)H * W + 255) / 256);
  switch (cubemap_mode(interp, seamless)) {
    case 0: reflect_shade_fwd_kernel<0><<<blocks, 256, 0, s>>>(cam, H, W, normal, base_color, refl_strength, cubemap, fail_value, (int)L, out_final, out_env); break;
    case 1: reflect_shade_fwd_kernel<1><<<blocks, 256, 0, s>>>(cam, H, W, normal, base_color, refl_strength, cubemap, fail_value, (int)L, out_final, out_env); break;
    default: reflect_shade_fwd_kernel<2><<<blocks, 256, 0, s>>>(cam, H, W, normal, base_color, refl_strength, cubemap, fail_value, (int)L, out_final, out_env); break;
  }
  B200GS_LAUNCHED("reflect_shade_fwd_kernel", false, s);
  return B200GS_OK;
}

int b200gs_reflect_shade_backward(int H, int W, const float* Kinv, const float* R, const float* T, const float* normal,
                                  const float* base_color, const float* refl_strength, const float* cubemap, uint32_t L,
                                  uint32_t interp, uint32_t seamless, const float* env_color, const float* dL_dfinal,
                                  const float* dL_denv, float* dL_dbase, float* dL_drefl, float* dL_dnormal,
                                  float* grad_cubemap, float* grad_fail, void* stream) {
  if (H < 0 || W < 0) { set_error("reflect_shade_backward: negative size"); return B200GS_EINVAL; }
  if (H == 0 || W == 0) return B200GS_OK;
  ShadeCam cam;
  if (shade_cam(Kinv, R, T, cam) || !normal || !base_color || !refl_strength || !cubemap || !env_color || !dL_dfinal || !dL_dbase ||
      !dL_drefl || !dL_dnormal || !grad_cubemap || !grad_fail || L == 0) {
    set_error("reflect_shade_backward: bad argument"); return B200GS_EINVAL; }
  if ((uint64_t)6 * 3 * L * L >= 0x7fffffffull || (long long)H * W >= 0x7fffffffll) { set_error("reflect_shade_backward: too large"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  StageScope _t(ST_SHADE_BWD, s);
  const unsigned blocks = (unsigned)(((long long)H * W + 255) / 256);
  switch (cubemap_mode(interp, seamless)) {
    case 0: reflect_shade_bwd_kernel<0><<<blocks, 256, 0, s>>>(cam, H, W, normal, base_color, refl_strength, cubemap, (int)L, env_color, dL_dfinal, dL_denv, dL_dbase, dL_drefl, dL_dnormal, grad_cubemap, grad_fail); break;
    case 1: reflect_shade_bwd_kernel<1><<<blocks, 256, 0, s>>>(cam, H, W, normal, base_color, refl_strength, cubemap, (int)L, env_color, dL_dfinal, dL_denv, dL_dbase, dL_drefl, dL_dnormal, grad_cubemap, grad_fail); break;
    default: reflect_shade_bwd_kernel<2><<<blocks, 256, 0, s>>>(cam, H, W, normal, base_color, refl_strength, cubemap, (int)L, env_color, dL_dfinal, dL_denv, dL_dbase, dL_drefl, dL_dnormal, grad_cubemap, grad_fail); break;
  }
  B200GS_LAUNCHED("reflect_shade_bwd_kernel", false, s);
  return B200GS_OK;
}

}  // extern "C"
