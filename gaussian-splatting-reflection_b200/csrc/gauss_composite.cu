// gauss_composite.cu -- K3g / K4g: per-tile compositing of the 3DGS (EWA) path, forward and
// reverse-order backward.
//
// Restates FORWARD::renderCUDA (G/cuda_rasterizer/forward.cu:274-411) and BACKWARD::renderCUDA
// (G/cuda_rasterizer/backward.cu:452-690).  Same machinery as the surfel kernels: one CTA per
// 16x16 tile, 8x4 pixel block per warp, 256-entry batches of packed 64-B records gathered with
// cp.async into a double-buffered ring, the compacted per-tile instance lists with 8-bit warp-block
// hit masks of tile_hits_kernel (binning.cu), and in the backward a warp reduce-scatter so each (tile, Gaussian) partial leaves as one 64-B vector of
// RED.ADD.F32 instead of the reference's 15-16 atomicAdds per contributing pixel.
//
// The forward arithmetic is pinned to the reference build's contraction (SASS of renderCUDA<3>):
//   power = fma(fma(dx, con.x*dx, (con.z*dy)*dy), -0.5, -((con.y*dx)*dy));  acc = fma(T, alpha*f, acc).
#include "surfel_composite.cuh"

#ifndef B200GS_GBWD_CTAS
#define B200GS_GBWD_CTAS 6   // resident CTAs (of 128 threads, two pixels each) per SM the register allocation is bounded for
#endif

namespace b200gs {

constexpr int GBATCH = 256;

struct GaussEval {
  float dx, dy, G, alpha;
};

// false when the reference `continue`s (power > 0 or alpha < 1/255).
__device__ __forceinline__ bool gauss_eval(const float4& q0, const float4& q1, float fx, float fy, GaussEval& e) {
  e.dx = __fsub_rn(q0.x, fx);
  e.dy = __fsub_rn(q0.y, fy);
  const float a = __fmul_rn(__fmul_rn(e.dy, q1.x), e.dy);            // (con.z*dy)*dy
  const float b = __fmul_rn(__fmul_rn(e.dx, q0.w), e.dy);            // (con.y*dx)*dy
  const float s = __fmaf_rn(e.dx, __fmul_rn(e.dx, q0.z), a);         // con.x*dx*dx + ...
  const float power = __fmaf_rn(s, -0.5f, -b);
  if (power > 0.0f) return false;
  e.G = expf(power);
  e.alpha = fminf(0.99f, __fmul_rn(q1.y, e.G));
  return !(e.alpha < 1.0f / 255.0f);
}

struct GCompFwdArgs {
  int W, H, gx;
  const uint2* ranges;        // ranges into the COMPACTED instance list
  const uint32_t* point_list; // compacted ids
  const uint32_t* list_pos;   // position of every compacted entry in the tile's full list
  const GaussRec* rec;
  const uint8_t* hit;         // per compacted entry: warp blocks that can reach alpha >= 1/255
  const float* bg;
  float* accum;
  uint32_t* n_contrib;
  float* out_color;
  float* out_normal;
  float* out_refl;
  float* out_invdepth;
  // depth slices (common.cuh, as in surfel_composite_fwd.cu): this launch composites slice k of S
  int k, S;
  const uint2* uranges;       // slice k's ranges in the UNcompacted list (their lengths number the contributors)
  const uint32_t* base_k;     // [tiles] list position at which slice k starts in every tile
  uint32_t* base_next;        // [tiles] the same for slice k+1 (written here)
  uint8_t* tile_done;         // [tiles] set when every pixel of the tile has saturated
  uint32_t* live;             // [0] tiles not done
  float* state;               // [GSTATE_PLANES,N] compositing state of unfinished pixels between slices
};

__global__ void __launch_bounds__(GBATCH) gauss_composite_fwd_kernel(const GCompFwdArgs a) {
  __shared__ GaussRec s_rec[2][GBATCH];
  __shared__ uint32_t s_pos[2][GBATCH];
  __shared__ uint8_t s_hit[2][GBATCH];

  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int tile = blockIdx.x;
  const int tile_x = tile % a.gx, tile_y = tile / a.gx;
  int lx, ly;
  tile_pixel(t, lx, ly);
  const int pxi = tile_x * TILE_X + lx, pyi = tile_y * TILE_Y + ly;
  const bool inside = pxi < a.W && pyi < a.H;
  const float fx = (float)pxi, fy = (float)pyi;
  const size_t N = (size_t)a.W * a.H;
  const size_t pix_id = (size_t)a.W * pyi + pxi;
  const bool sliced = a.S > 1;
  if (sliced && a.k > 0 && a.tile_done[tile]) return;      // finished (and its outputs written) by an earlier slice
  const uint32_t base = a.k > 0 ? a.base_k[tile] : 0u;     // contributors are numbered through all slices of the tile
  const uint2 range = a.ranges[tile];
  const int n = (int)(range.y - range.x);
  const int rounds = (n + GBATCH - 1) / GBATCH;

  bool done = !inside;
  float T = 1.0f;
  uint32_t last_contributor = 0;
  float C0 = 0.f, C1 = 0.f, C2 = 0.f, N0 = 0.f, N1 = 0.f, N2 = 0.f, refl = 0.f, invd = 0.f;
  if (sliced && a.k > 0 && inside) {                        // pick up where the previous slice left this pixel
    const float* st = a.state + pix_id;
    const float Ts = st[0];
    done = Ts < 0.0f; T = fabsf(Ts);                        // T >= 1e-4 > 0: the sign carries `done`
    C0 = st[1 * N]; C1 = st[2 * N]; C2 = st[3 * N];
    N0 = st[4 * N]; N1 = st[5 * N]; N2 = st[6 * N];
    refl = st[7 * N]; invd = st[8 * N];
    last_contributor = __float_as_uint(st[9 * N]);
  }

  const uint32_t* pl = a.point_list + range.x;
  const uint8_t* ph = a.hit + range.x;
  const uint32_t* pp = a.list_pos + range.x;
  unsigned hit_cur = 0u, hit_next = 0u;
  uint32_t pos_next = 0u;
  auto stage = [&](int buf, int id) {
    hit_cur = hit_next;                  // 0 for padding entries: never visited
    if (id >= 0) {
      s_pos[buf][t] = pos_next;
      const GaussRec* g = a.rec + id;
      GaussRec* d = &s_rec[buf][t];
#pragma unroll
      for (int k = 0; k < 4; ++k) cp_async16(&d->q[k], &g->q[k]);
    }
    cp_async_commit();
  };
  int id_next = -1;
  if (t < n) { id_next = (int)pl[t]; hit_next = ph[t]; pos_next = pp[t]; }
  stage(0, id_next);
  id_next = -1; hit_next = 0u;
  if (GBATCH + t < n) { id_next = (int)pl[GBATCH + t]; hit_next = ph[GBATCH + t]; pos_next = pp[GBATCH + t]; }

  for (int i = 0; i < rounds; ++i) {
    cp_async_wait_all();
    s_hit[i & 1][t] = (uint8_t)hit_cur;
    if (__syncthreads_count(done) == GBATCH) break;
    if (i + 1 < rounds) {
      stage((i + 1) & 1, id_next);
      const int nx = (i + 2) * GBATCH + t;
      id_next = -1; hit_next = 0u;
      if (nx < n) { id_next = (int)pl[nx]; hit_next = ph[nx]; pos_next = pp[nx]; }
    }
    const int buf = i & 1;
    const GaussRec* rb = s_rec[buf];
    bool warp_done = false;
#pragma unroll 1
    for (int r = 0; r < GBATCH / 32 && !warp_done; ++r) {
      unsigned m = __ballot_sync(0xffffffffu, (s_hit[buf][r * 32 + lane] >> warp) & 1u);
      while (m) {
        if (__all_sync(0xffffffffu, done)) { warp_done = true; break; }
        const int j = r * 32 + __ffs(m) - 1;
        m &= m - 1;
        const float4 q0 = rb[j].q[0], q1 = rb[j].q[1];
        GaussEval e;
        if (!gauss_eval(q0, q1, fx, fy, e) || done) continue;
        const float test_T = __fmul_rn(T, __fsub_rn(1.0f, e.alpha));
        if (test_T < 0.0001f) { done = true; continue; }
        const float4 q2 = rb[j].q[2], q3 = rb[j].q[3];
        C0 = __fmaf_rn(T, __fmul_rn(e.alpha, q2.x), C0);
        C1 = __fmaf_rn(T, __fmul_rn(e.alpha, q2.y), C1);
        C2 = __fmaf_rn(T, __fmul_rn(e.alpha, q2.z), C2);
        N0 = __fmaf_rn(T, __fmul_rn(e.alpha, q3.x), N0);
        N1 = __fmaf_rn(T, __fmul_rn(e.alpha, q3.y), N1);
        N2 = __fmaf_rn(T, __fmul_rn(e.alpha, q3.z), N2);
        refl = __fmaf_rn(T, __fmul_rn(e.alpha, q1.w), refl);
        invd = __fmaf_rn(T, __fmul_rn(e.alpha, q1.z), invd);
        T = test_T;
        last_contributor = base + s_pos[buf][j] + 1u;   // 1-based position in the tile's (full) list
      }
    }
  }
  cp_async_wait_all();

  bool finalize = true;
  if (sliced) {
    const bool last = a.k == a.S - 1;
    const bool all_done = __syncthreads_count(done) == GBATCH;
    finalize = last || all_done;
    if (t == 0 && !last) {
      const uint2 ur = a.uranges[tile];
      a.base_next[tile] = base + (ur.y - ur.x);
      if (all_done) { a.tile_done[tile] = 1; atomicSub(a.live, 1u); }   // later slices emit nothing into this tile
    }
    if (!finalize && inside) {
      float* st = a.state + pix_id;
      st[0] = done ? -T : T;
      st[1 * N] = C0; st[2 * N] = C1; st[3 * N] = C2;
      st[4 * N] = N0; st[5 * N] = N1; st[6 * N] = N2;
      st[7 * N] = refl; st[8 * N] = invd;
      st[9 * N] = __uint_as_float(last_contributor);
    }
  }

  if (finalize && inside) {  // forward.cu:395-410
    a.accum[pix_id] = T;
    a.n_contrib[pix_id] = last_contributor;
    a.out_color[pix_id] = __fmaf_rn(T, a.bg[0], C0);
    a.out_color[pix_id + N] = __fmaf_rn(T, a.bg[1], C1);
    a.out_color[pix_id + 2 * N] = __fmaf_rn(T, a.bg[2], C2);
    a.out_normal[pix_id] = N0;
    a.out_normal[pix_id + N] = N1;
    a.out_normal[pix_id + 2 * N] = N2;
    a.out_refl[pix_id] = refl;
    if (a.out_invdepth) a.out_invdepth[pix_id] = invd;
  }
}

int gauss_composite_forward(const b200gs_gauss_fwd& f, const GaussGeom& g, const GaussImage& im, const Binning& b,
                            int k, int S, cudaStream_t s) {
  GCompFwdArgs a;
  a.W = f.width; a.H = f.height;
  a.gx = (f.width + TILE_X - 1) / TILE_X;
  const int gy = (f.height + TILE_Y - 1) / TILE_Y;
  const size_t tiles = (size_t)a.gx * gy;
  a.k = k; a.S = S;
  a.uranges = im.ranges + (size_t)k * (tiles + 1);
  a.base_k = im.tile_base + (size_t)k * tiles; a.base_next = im.tile_base + (size_t)(k + 1) * tiles;
  a.tile_done = im.tile_done; a.live = im.live; a.state = im.state;
  a.ranges = im.cranges + (size_t)k * (tiles + 1); a.point_list = b.comp_id; a.list_pos = b.comp_pos; a.rec = g.rec; a.hit = b.comp_hit; a.bg = f.background;
  a.accum = im.accum; a.n_contrib = im.n_contrib;
  a.out_color = f.out_color; a.out_normal = f.out_normal; a.out_refl = f.out_refl; a.out_invdepth = f.out_invdepth;
  StageScope _t(ST_COMPOSITE_FWD, s);
  gauss_composite_fwd_kernel<<<a.gx * gy, GBATCH, 0, s>>>(a);
  B200GS_LAUNCHED("gauss_composite_fwd_kernel", f.debug, s);
  return B200GS_OK;
}

// ------------------------------------------------------------------ backward
struct GCompBwdArgs {
  int W, H, gx;
  const uint2* ranges;        // compacted lists, as in the forward
  const uint32_t* point_list;
  const uint32_t* list_pos;
  const GaussRec* rec;
  const uint8_t* hit;
  const float* bg;
  const float* accum;
  const uint32_t* n_contrib;
  const float* dL_dpix;      // [3,N]
  const float* dL_dnormal;   // [3,N]
  const float* dL_drefl;     // [N] or NULL
  const float* dL_dinvd;     // [N] or NULL
  float* acc;                // [P, GACC_STRIDE]
  int tiles;
  const int* plan;            // device copy of the forward's slice plan ([1] = S)
  const uint32_t* tile_base;  // [MAX_SLICES+1][tiles]
};

constexpr int GPAIR = PAIR_THREADS;   // the backward: two pixels per thread, 128 threads per tile (surfel_composite.cuh)

// Everything the backward keeps per pixel.
struct GaussPixelBwd {
  float fx, fy;
  uint32_t last_contributor;          // 0 for pixels outside the image
  float T_final, T;
  float dpix0, dpix1, dpix2, dn0, dn1, dn2, drefl, dinv, bg_dot, mpix0, mpix1;
  float acc_q, last_q, acc_m, last_m, last_alpha;

  __device__ __forceinline__ void load(const GCompBwdArgs& a, int pxi, int pyi) {
    const bool inside = pxi < a.W && pyi < a.H;
    fx = (float)pxi; fy = (float)pyi;
    const size_t N = (size_t)a.W * a.H;
    const size_t pix_id = (size_t)a.W * pyi + pxi;
    T_final = inside ? a.accum[pix_id] : 0.0f;
    T = T_final;
    last_contributor = inside ? a.n_contrib[pix_id] : 0u;
    dpix0 = dpix1 = dpix2 = dn0 = dn1 = dn2 = drefl = dinv = 0.f;
    if (inside) {
      dpix0 = a.dL_dpix[pix_id]; dpix1 = a.dL_dpix[pix_id + N]; dpix2 = a.dL_dpix[pix_id + 2 * N];
      dn0 = a.dL_dnormal[pix_id]; dn1 = a.dL_dnormal[pix_id + N]; dn2 = a.dL_dnormal[pix_id + 2 * N];
      drefl = a.dL_drefl ? a.dL_drefl[pix_id] : 0.0f;
      dinv = a.dL_dinvd ? a.dL_dinvd[pix_id] : 0.0f;
    }
    bg_dot = 0.0f;
    bg_dot += a.bg[0] * dpix0;
    bg_dot += a.bg[1] * dpix1;
    bg_dot += a.bg[2] * dpix2;
    // the running-sum quirk (backward.cu:614): the mean2D gradient uses 3 t0 + 2 t1 + t2 of the colour terms
    mpix0 = 3.0f * dpix0; mpix1 = 2.0f * dpix1;
    acc_q = last_q = acc_m = last_m = last_alpha = 0.f;
  }

  // One contributing (pixel, Gaussian) pair: advances the recurrences and ADDS the pair's 16 partials to g.
  __device__ __forceinline__ void contribute(const GaussEval& e, const float4& q0, const float4& q1, const float4& q2,
                                             const float4& q3, float ddelx_dx, float ddely_dy, float* g) {
    const float alpha = e.alpha, G = e.G;
    const float inv_1ma = rcp_newton(__fsub_rn(1.0f, alpha));   // as in the surfel backward
    T = T * inv_1ma;
    const float w = alpha * T;                      // dchannel_dcolor
    const float om = 1.f - last_alpha;
    // backward.cu:596-650: every channel enters dL_dalpha as (value - value accumulated behind) *
    // upstream gradient.  The recurrence is linear, so the eight per-channel recurrences collapse
    // into one on q = sum_ch value_ch * upstream_ch, plus one on the colour-only weighted sum the
    // mean2D running-sum quirk (:614) needs.
    const float cm = q2.x * mpix0 + q2.y * mpix1 + q2.z * dpix2;
    const float qv = q2.x * dpix0 + q2.y * dpix1 + q2.z * dpix2 + q3.x * dn0 + q3.y * dn1 + q3.z * dn2 +
                     q1.w * drefl + q1.z * dinv;
    acc_q = last_alpha * last_q + om * acc_q; last_q = qv;
    acc_m = last_alpha * last_m + om * acc_m; last_m = cm;
    last_alpha = alpha;
    g[8] += w * dpix0; g[9] += w * dpix1; g[10] += w * dpix2;
    g[11] += w * dn0; g[12] += w * dn1; g[13] += w * dn2;
    g[14] += w * drefl; g[15] += w * dinv;
    const float bgterm = (-T_final * inv_1ma) * bg_dot;
    const float dL_dalpha = (qv - acc_q) * T + bgterm;
    const float dL_dalpha_m = (cm - acc_m) * T + bgterm;
    const float opa = q1.y;
    const float dL_dG = opa * dL_dalpha, dL_dG_m = opa * dL_dalpha_m;
    const float gdx = G * e.dx, gdy = G * e.dy;
    const float dG_ddelx = -gdx * q0.z - gdy * q0.w;
    const float dG_ddely = -gdy * q1.x - gdx * q0.w;
    g[0] += dL_dG * dG_ddelx * ddelx_dx;
    g[1] += dL_dG * dG_ddely * ddely_dy;
    g[2] += dL_dG_m * dG_ddelx * ddelx_dx;
    g[3] += dL_dG_m * dG_ddely * ddely_dy;
    g[4] += -0.5f * gdx * e.dx * dL_dG;
    g[5] += -0.5f * gdx * e.dy * dL_dG;
    g[6] += -0.5f * gdy * e.dy * dL_dG;
    g[7] += G * dL_dalpha;
  }
};

__global__ void __launch_bounds__(GPAIR, B200GS_GBWD_CTAS) gauss_composite_bwd_kernel(const GCompBwdArgs a) {
  __shared__ GaussRec s_rec[2][GPAIR];
  __shared__ int s_id[2][GPAIR];
  __shared__ uint32_t s_pos[2][GPAIR];
  __shared__ uint8_t s_hit[2][GPAIR];
  __shared__ uint32_t s_last[GPAIR / 32];
  __shared__ uint32_t s_tile_last;
  __shared__ int s_n;                   // live prefix of the slice's compacted list

  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int tile = blockIdx.x;
  const int tile_x = tile % a.gx, tile_y = tile / a.gx;
  int lx, ly;
  pair_pixel(t, lx, ly);
  GaussPixelBwd px[2];
  px[0].load(a, tile_x * TILE_X + lx, tile_y * TILE_Y + ly);
  px[1].load(a, tile_x * TILE_X + lx, tile_y * TILE_Y + ly + 4);
  const unsigned warp_bits = pair_hit_bits(warp);       // the two 8x4 blocks of this warp's 8x8 pixels

  // lane -> accumulator component owned after the reduce-scatter (16 values, 32 lanes)
  const int b16 = (lane >> 4) & 1, b8 = (lane >> 3) & 1, b4 = (lane >> 2) & 1, b2 = (lane >> 1) & 1;
  const int comp = b16 * 8 + b8 * 4 + b4 * 2 + b2;
  const bool writer = (lane & 1) == 0;
  const float ddelx_dx = 0.5f * a.W, ddely_dy = 0.5f * a.H;

  // Entries at or behind the last contributor of every pixel of the tile are never used
  // (backward.cu:560: `contributor >= last_contributor` -> continue).
  const uint32_t warp_last = __reduce_max_sync(0xffffffffu, max(px[0].last_contributor, px[1].last_contributor));
  if (lane == 0) s_last[warp] = warp_last;
  __syncthreads();
  if (t == 0) {
    uint32_t tl = 0;
#pragma unroll
    for (int w = 0; w < GPAIR / 32; ++w) tl = max(tl, s_last[w]);
    s_tile_last = tl;
  }
  __syncthreads();
  const uint32_t tile_last = s_tile_last;
  if (tile_last == 0u) return;

  // The tile's list is the concatenation of its depth slices' lists (one slice unless the forward was sliced): walk
  // the slices back to front, each slice's entries back to front.  The per-pixel recurrences simply carry on.
  const int S = a.plan[1];
  for (int sl = S - 1; sl >= 0; --sl) {
    const uint32_t base = sl == 0 ? 0u : a.tile_base[(size_t)sl * a.tiles + tile];   // list position of the slice's first entry
    if (base >= tile_last) continue;
    const uint2 range = a.ranges[(size_t)sl * (a.tiles + 1) + tile];
    if (range.y == range.x) continue;
    const uint32_t* pl = a.point_list + range.x;
    const uint8_t* ph = a.hit + range.x;
    const uint32_t* pp = a.list_pos + range.x;
    // list positions ascend, so the live part is a prefix of the slice's compacted list; find its length by bisection
    __syncthreads();                                  // everyone has left the previous slice's shared buffers
    if (t == 0) {
      const uint32_t lim = tile_last - base;
      int lo = 0, hi = (int)(range.y - range.x);     // first index with pos >= lim
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (pp[mid] < lim) lo = mid + 1; else hi = mid;
      }
      s_n = lo;
    }
    __syncthreads();
    const int n = s_n;
    const int rounds = (n + GPAIR - 1) / GPAIR;
    if (n == 0) continue;

    unsigned hit_cur = 0u, hit_next = 0u;
    uint32_t pos_next = 0u;
    auto stage = [&](int buf, int id) {
      hit_cur = hit_next;
      if (id >= 0) {
        s_pos[buf][t] = pos_next;
        const GaussRec* g = a.rec + id;
        GaussRec* d = &s_rec[buf][t];
#pragma unroll
        for (int k = 0; k < 4; ++k) cp_async16(&d->q[k], &g->q[k]);
        s_id[buf][t] = id;
      }
      cp_async_commit();
    };
    int id_next = -1;
    if (t < n) { id_next = (int)pl[n - 1 - t]; hit_next = ph[n - 1 - t]; pos_next = pp[n - 1 - t]; }
    stage(0, id_next);
    id_next = -1; hit_next = 0u;
    if (GPAIR + t < n) { const int q = n - 1 - (GPAIR + t); id_next = (int)pl[q]; hit_next = ph[q]; pos_next = pp[q]; }

    for (int i = 0; i < rounds; ++i) {
      cp_async_wait_all();
      s_hit[i & 1][t] = (uint8_t)hit_cur;
      __syncthreads();
      if (i + 1 < rounds) {
        stage((i + 1) & 1, id_next);
        const int nx = (i + 2) * GPAIR + t;
        id_next = -1; hit_next = 0u;
        if (nx < n) { const int q = n - 1 - nx; id_next = (int)pl[q]; hit_next = ph[q]; pos_next = pp[q]; }
      }
      const int buf = i & 1;
      const int count = min(GPAIR, n - i * GPAIR);
      const GaussRec* rb = s_rec[buf];
#pragma unroll 1
      for (int r = 0; r < (count + 31) >> 5; ++r) {
        // entries of other warp blocks, and entries behind every pixel's last contributor, are skipped
        const int en = r * 32 + lane;
        unsigned m = __ballot_sync(0xffffffffu, en < count && (s_hit[buf][en] & warp_bits) && base + s_pos[buf][en] < warp_last);
        while (m) {
          const int j = r * 32 + __ffs(m) - 1;
          m &= m - 1;
          const uint32_t contributor = base + s_pos[buf][j];   // 0-based position in the tile's full list
          const float4 q0 = rb[j].q[0], q1 = rb[j].q[1], q2 = rb[j].q[2], q3 = rb[j].q[3];

          float g[16];
#pragma unroll
          for (int k = 0; k < 16; ++k) g[k] = 0.0f;
#pragma unroll
          for (int p = 0; p < 2; ++p) {
            GaussEval e;
            bool c = contributor < px[p].last_contributor;
            c = gauss_eval(q0, q1, px[p].fx, px[p].fy, e) && c;
            if (c) px[p].contribute(e, q0, q1, q2, q3, ddelx_dx, ddely_dy, g);
          }
          // warp reduce-scatter 16 -> 8 -> 4 -> 2 -> 1, then the last butterfly (a visit without any contributing
          // pixel adds zeros: cheaper than a vote and a branch on every visit)
          rs_step<8>(g, 16, b16);
          rs_step<4>(g, 8, b8);
          rs_step<2>(g, 4, b4);
          rs_step<1>(g, 2, b2);
          g[0] += __shfl_xor_sync(0xffffffffu, g[0], 1);
          if (writer) atomicAdd(a.acc + (size_t)s_id[buf][j] * GACC_STRIDE + comp, g[0]);
        }
      }
    }
    cp_async_wait_all();
  }   // slices
}

int gauss_composite_backward(const b200gs_gauss_bwd& f, const GaussGeom& g, const GaussImage& im, const Binning& b,
                             float* acc, cudaStream_t s) {
  GCompBwdArgs a;
  a.W = f.width; a.H = f.height;
  a.gx = (f.width + TILE_X - 1) / TILE_X;
  const int gy = (f.height + TILE_Y - 1) / TILE_Y;
  a.ranges = im.cranges; a.point_list = b.comp_id; a.list_pos = b.comp_pos; a.rec = g.rec; a.hit = b.comp_hit; a.bg = f.background;
  a.accum = im.accum; a.n_contrib = im.n_contrib;
  a.dL_dpix = f.dL_dout_color; a.dL_dnormal = f.dL_dout_normal; a.dL_drefl = f.dL_dout_refl; a.dL_dinvd = f.dL_dout_invdepth;
  a.acc = acc;
  a.tiles = a.gx * gy; a.plan = g.plan; a.tile_base = im.tile_base;
  StageScope _t(ST_COMPOSITE_BWD, s);
  B200GS_CUDA(cudaMemsetAsync(acc, 0, (size_t)f.P * GACC_STRIDE * sizeof(float), s));
  gauss_composite_bwd_kernel<<<a.gx * gy, GPAIR, 0, s>>>(a);
  B200GS_LAUNCHED("gauss_composite_bwd_kernel", f.debug, s);
  return B200GS_OK;
}

}  // namespace b200gs
