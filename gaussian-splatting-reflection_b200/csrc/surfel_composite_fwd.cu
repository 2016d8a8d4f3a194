// surfel_composite_fwd.cu -- K3s: per-tile front-to-back compositing of colour,
// reflection strength and the 8 auxiliary planes of the surfel path.
//
// Restates FORWARD::renderCUDA (S/cuda_rasterizer/forward.cu:258-489).  One CTA
// per 16x16 tile, one pixel per thread.  Batches of 256 packed 80-B Gaussian
// records are gathered with cp.async into a double-buffered shared-memory ring
// (the gather of batch i+1 and the id load of batch i+2 overlap the blending of
// batch i); the inner loop is warp-synchronous so whole warps leave as soon as
// all their pixels are saturated, and the per-Gaussian max blending weight is
// reduced with redux.sync + one shared atomicMax per (warp, Gaussian) instead of
// the reference's racy global compare/exchange per pixel.
#include "surfel_composite.cuh"

namespace b200gs {

struct CompFwdArgs {
  int W, H, gx;
  const uint2* ranges;        // ranges into the COMPACTED instance list
  const uint32_t* point_list; // compacted ids
  const uint32_t* list_pos;   // position of every compacted entry in the tile's full list
  const SurfelRec* rec;
  const uint8_t* hit;         // per compacted entry: warp blocks that can reach alpha >= 1/255 (tile_hits_kernel)
  const float* bg;
  float* accum;         // [3,N]
  uint32_t* n_contrib;  // [2,N]
  float* out_color;
  float* out_others;
  float* out_refl;
  float* gaussian_weights;
  // depth slices (common.cuh): this launch composites slice k of S; S == 1 is the whole list in one go
  int k, S;
  const uint2* uranges;       // slice k's ranges in the UNcompacted list (their lengths number the contributors)
  const uint32_t* base_k;     // [tiles] list position at which slice k starts in every tile
  uint32_t* base_next;        // [tiles] the same for slice k+1 (written here)
  uint8_t* tile_done;         // [tiles] set when every pixel of the tile has saturated
  uint32_t* live;             // [0] tiles not done
  float* state;               // [STATE_PLANES,N] compositing state of unfinished pixels between slices
};

constexpr int BATCH = 256;

// Cull audit (b200gs_surfel_fwd.cull_audit): the AUDIT instantiation ignores the hit masks, evaluates
// every (warp, entry) pair like the reference does and counts the pairs that pass the
// reference's alpha cut although their mask bit is clear -- [0] with the forward's
// `unstable` threshold, [1] with the backward's.  Both must stay 0.
__device__ unsigned long long g_cull_violations[2];
// Workload statistics gathered by the AUDIT instantiation only: [0] (warp, entry) pairs visited with
// the hit bit set, [1] those with >= 1 contributing lane, [2] lane contributions, [3] sum over
// (warp, batch) of the busiest lane's contributions, [4] (warp, batch) count.
__device__ unsigned long long g_cull_stats[8];

template <bool AUDIT>
__global__ void __launch_bounds__(BATCH) surfel_composite_fwd_kernel(const CompFwdArgs a) {
  __shared__ SurfelRec s_rec[2][BATCH];
  __shared__ int s_id[2][BATCH];
  __shared__ uint32_t s_pos[2][BATCH];
  __shared__ int s_wmax[2][BATCH];
  __shared__ uint8_t s_hit[2][BATCH];   // bit w: the entry's cull box overlaps warp w's 8x4 pixel block

  const int t = threadIdx.x, lane = t & 31;
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
  const int rounds = (n + BATCH - 1) / BATCH;

  bool done = !inside;
  float T = 1.0f;
  uint32_t last_contributor = 0;
  float C0 = 0.f, C1 = 0.f, C2 = 0.f, refl = 0.f, mask = 0.f;
  float N0 = 0.f, N1 = 0.f, N2 = 0.f, D = 0.f, M1 = 0.f, M2 = 0.f, distortion = 0.f, median_depth = 0.f;
  float median_contributor = -1.0f;
  if (sliced && a.k > 0 && inside) {                        // pick up where the previous slice left this pixel
    const float* st = a.state + pix_id;
    const float Ts = st[0];
    done = Ts < 0.0f; T = fabsf(Ts);                        // T >= 1e-4 > 0: the sign carries `done`
    C0 = st[1 * N]; C1 = st[2 * N]; C2 = st[3 * N];
    N0 = st[4 * N]; N1 = st[5 * N]; N2 = st[6 * N];
    D = st[7 * N]; M1 = st[8 * N]; M2 = st[9 * N]; distortion = st[10 * N];
    median_depth = st[11 * N]; median_contributor = st[12 * N]; refl = st[13 * N]; mask = st[14 * N];
    last_contributor = __float_as_uint(st[15 * N]);
  }

  s_wmax[0][t] = 0;
  s_wmax[1][t] = 0;

  const uint32_t* pl = a.point_list + range.x;
  const uint8_t* ph = a.hit + range.x;
  const uint32_t* pp = a.list_pos + range.x;
  unsigned hit_cur = 0u;                 // hit mask of the entry this thread staged for the batch about to run
  unsigned hit_next = 0u;
  uint32_t pos_next = 0u;
  auto stage = [&](int buf, int id) {
    hit_cur = hit_next;                  // 0 for padding entries: never visited
    if (id >= 0) {
      s_pos[buf][t] = base + pos_next + 1u;   // the entry's 1-based position in the tile's (full) list
      const SurfelRec* g = a.rec + id;
      SurfelRec* d = &s_rec[buf][t];
#pragma unroll
      for (int k = 0; k < 5; ++k) cp_async16(&d->q[k], &g->q[k]);
      s_id[buf][t] = id;
    }
    cp_async_commit();
  };
  int id_next = -1;
  if (t < n) { id_next = (int)pl[t]; hit_next = ph[t]; pos_next = pp[t]; }
  stage(0, id_next);
  id_next = -1; hit_next = 0u;
  if (BATCH + t < n) { id_next = (int)pl[BATCH + t]; hit_next = ph[BATCH + t]; pos_next = pp[BATCH + t]; }

  const int warp = t >> 5;
  int last_batch = -1;
  for (int i = 0; i < rounds; ++i) {
    cp_async_wait_all();
    // s_hit[i&1] was last read while blending batch i-2; everyone left that before the barrier of i-1
    s_hit[i & 1][t] = (uint8_t)hit_cur;
    // batch i has landed for everyone, and everyone is past batch i-1
    if (__syncthreads_count(done) == BATCH) break;
    if (i > 0) {  // publish and clear the per-Gaussian weight maxima of batch i-1
      const int b = (i - 1) & 1;
      const int v = s_wmax[b][t];
      if (v > 0) { atomicMax(reinterpret_cast<int*>(a.gaussian_weights) + s_id[b][t], v); s_wmax[b][t] = 0; }
    }
    if (i + 1 < rounds) {
      stage((i + 1) & 1, id_next);
      const int nx = (i + 2) * BATCH + t;
      id_next = -1; hit_next = 0u;
      if (nx < n) { id_next = (int)pl[nx]; hit_next = ph[nx]; pos_next = pp[nx]; }
    }
    last_batch = i;
    const int buf = i & 1;
    const SurfelRec* rb = s_rec[buf];
    bool warp_done = false;
    int lane_cnt = 0;
#pragma unroll 1
    for (int r = 0; r < BATCH / 32 && !warp_done; ++r) {
      unsigned m = __ballot_sync(0xffffffffu, AUDIT ? (i * BATCH + r * 32 + lane < n)
                                                          : (bool)((s_hit[buf][r * 32 + lane] >> warp) & 1u));
      while (m) {
        if (__all_sync(0xffffffffu, done)) { warp_done = true; break; }
        const int j = r * 32 + __ffs(m) - 1;
        m &= m - 1;
        const uint32_t contributor = s_pos[buf][j];              // 1-based position in the tile's (full) list
        const float4 q0 = rb[j].q[0], q1 = rb[j].q[1], q2 = rb[j].q[2];
        SurfelHit h;
        bool c = surfel_eval<false>(q0, q1, q2, fx, fy, h);
        if (AUDIT && inside && !((s_hit[buf][j] >> warp) & 1u)) {
          SurfelHit hb;
          if (c) atomicAdd(&g_cull_violations[0], 1ull);
          if (surfel_eval<true>(q0, q1, q2, fx, fy, hb)) atomicAdd(&g_cull_violations[1], 1ull);
        }
        c = c && !done;
        float w = 0.0f;
        if (c) {
          const float test_T = __fmul_rn(T, __fsub_rn(1.0f, h.alpha));
          if (test_T < 0.0001f) {
            done = true;
            c = false;
          } else {
            const float4 q3 = rb[j].q[3], q4 = rb[j].q[4];
            w = __fmul_rn(T, h.alpha);
            const float A = __fsub_rn(1.0f, T);
            const float m_ = ndc_depth(h.depth);
            const float mm = __fmul_rn(m_, m_);
            // distortion += (m*m*A + M2 - 2*m*M1) * w   (forward.cu:424)
            const float e = __fmaf_rn(-M1, __fadd_rn(m_, m_), __fmaf_rn(A, mm, M2));
            distortion = __fmaf_rn(e, w, distortion);
            D = __fmaf_rn(h.depth, w, D);
            M1 = __fmaf_rn(m_, w, M1);
            M2 = __fmaf_rn(mm, w, M2);
            if (T > 0.5f) { median_depth = h.depth; median_contributor = (float)contributor; }
            N0 = __fmaf_rn(w, q3.x, N0); N1 = __fmaf_rn(w, q3.y, N1); N2 = __fmaf_rn(w, q3.z, N2);
            C0 = __fmaf_rn(w, q4.x, C0); C1 = __fmaf_rn(w, q4.y, C1); C2 = __fmaf_rn(w, q4.z, C2);
            refl = __fmaf_rn(w, q3.w, refl);
            if (q4.w != 0.0f) mask = 1.0f;
            T = test_T;
            last_contributor = contributor;
          }
        }
        if (AUDIT) {
          const unsigned cm = __ballot_sync(0xffffffffu, c);
          lane_cnt += c ? 1 : 0;
          if (lane == 0 && ((s_hit[buf][j] >> warp) & 1u)) {
            atomicAdd(&g_cull_stats[0], 1ull);
            if (cm) atomicAdd(&g_cull_stats[1], 1ull);
            atomicAdd(&g_cull_stats[2], (unsigned long long)__popc(cm));
          }
        }
        // gaussian_weights[id] = max over pixels of w  (forward.cu:457-459, made exact).  w is 0 for a pixel that did not
        // blend the entry and > 0 otherwise; 97 % of the visits have a contributing pixel, so no vote in front of it.
        {
          const int wm = __reduce_max_sync(0xffffffffu, __float_as_int(w));
          if (lane == 0 && wm > 0) atomicMax(&s_wmax[buf][j], wm);
        }
      }
    }
    if (AUDIT) {
      const int mx = __reduce_max_sync(0xffffffffu, lane_cnt);
      if (lane == 0) { atomicAdd(&g_cull_stats[3], (unsigned long long)mx); atomicAdd(&g_cull_stats[4], 1ull); }
    }
  }
  __syncthreads();
  if (last_batch >= 0) {
    const int b = last_batch & 1;
    const int v = s_wmax[b][t];
    if (v > 0) atomicMax(reinterpret_cast<int*>(a.gaussian_weights) + s_id[b][t], v);
  }
  cp_async_wait_all();

  bool finalize = true;
  if (sliced) {
    const bool last = a.k == a.S - 1;
    const bool all_done = __syncthreads_count(done) == BATCH;
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
      st[7 * N] = D; st[8 * N] = M1; st[9 * N] = M2; st[10 * N] = distortion;
      st[11 * N] = median_depth; st[12 * N] = median_contributor; st[13 * N] = refl; st[14 * N] = mask;
      st[15 * N] = __uint_as_float(last_contributor);
    }
  }

  if (finalize && inside) {  // forward.cu:465-488
    a.accum[pix_id] = T;
    a.accum[pix_id + N] = M1;
    a.accum[pix_id + 2 * N] = M2;
    a.n_contrib[pix_id] = last_contributor;
    a.n_contrib[pix_id + N] = (uint32_t)median_contributor;  // -1 saturates to 0 (cvt.rzi.u32.f32)
    a.out_color[pix_id] = __fmaf_rn(T, a.bg[0], C0);
    a.out_color[pix_id + N] = __fmaf_rn(T, a.bg[1], C1);
    a.out_color[pix_id + 2 * N] = __fmaf_rn(T, a.bg[2], C2);
    a.out_refl[pix_id] = refl;
    a.out_others[pix_id + 0 * N] = D;
    a.out_others[pix_id + 1 * N] = __fsub_rn(1.0f, T);
    a.out_others[pix_id + 2 * N] = N0;
    a.out_others[pix_id + 3 * N] = N1;
    a.out_others[pix_id + 4 * N] = N2;
    a.out_others[pix_id + 5 * N] = median_depth;
    a.out_others[pix_id + 6 * N] = distortion;
    a.out_others[pix_id + 7 * N] = mask;
  }
}

int surfel_composite_forward(const b200gs_surfel_fwd& f, const SurfelGeom& g, const SurfelImage& im, const Binning& b,
                             int k, int S, cudaStream_t s) {
  CompFwdArgs a;
  a.W = f.width; a.H = f.height;
  a.gx = (f.width + TILE_X - 1) / TILE_X;
  const int gy = (f.height + TILE_Y - 1) / TILE_Y;
  const size_t tiles = (size_t)a.gx * gy;
  a.k = k; a.S = S;
  a.ranges = im.cranges + (size_t)k * (tiles + 1); a.uranges = im.ranges + (size_t)k * (tiles + 1);
  a.base_k = im.tile_base + (size_t)k * tiles; a.base_next = im.tile_base + (size_t)(k + 1) * tiles;
  a.tile_done = im.tile_done; a.live = im.live; a.state = im.state;
  a.point_list = b.comp_id; a.list_pos = b.comp_pos; a.rec = g.rec; a.hit = b.comp_hit; a.bg = f.background;
  a.accum = im.accum; a.n_contrib = im.n_contrib;
  a.out_color = f.out_color; a.out_others = f.out_others; a.out_refl = f.out_refl;
  a.gaussian_weights = f.gaussian_weights;
  StageScope _t(ST_COMPOSITE_FWD, s);
  if (f.cull_audit)
    surfel_composite_fwd_kernel<true><<<a.gx * gy, BATCH, 0, s>>>(a);
  else
    surfel_composite_fwd_kernel<false><<<a.gx * gy, BATCH, 0, s>>>(a);
  B200GS_LAUNCHED("surfel_composite_fwd_kernel", f.debug, s);
  return B200GS_OK;
}

}  // namespace b200gs

extern "C" {
int b200gs_debug_cull_stats(unsigned long long* out8, int reset) {
  using namespace b200gs;
  B200GS_CUDA(cudaDeviceSynchronize());
  if (out8) B200GS_CUDA(cudaMemcpyFromSymbol(out8, g_cull_stats, 8 * sizeof(unsigned long long)));
  if (reset) {
    const unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    B200GS_CUDA(cudaMemcpyToSymbol(g_cull_stats, z, sizeof(z)));
  }
  return B200GS_OK;
}
int b200gs_debug_cull_violations(unsigned long long* out2, int reset) {
  using namespace b200gs;
  B200GS_CUDA(cudaDeviceSynchronize());
  if (out2) B200GS_CUDA(cudaMemcpyFromSymbol(out2, g_cull_violations, 2 * sizeof(unsigned long long)));
  if (reset) {
    const unsigned long long z[2] = {0ull, 0ull};
    B200GS_CUDA(cudaMemcpyToSymbol(g_cull_violations, z, sizeof(z)));
  }
  return B200GS_OK;
}
}
