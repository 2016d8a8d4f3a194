// surfel_composite_bwd.cu -- K4s: reverse-order (back-to-front) backward of the
// surfel compositing.
//
// Restates BACKWARD::renderCUDA (S/cuda_rasterizer/backward.cu:143-470).  Same
// tiling and staging as the forward kernel.  Where the reference issues up to 17
// global atomicAdds per contributing (pixel, Gaussian) pair, this kernel keeps
// the per-pixel recurrences in registers, forms the 19 per-Gaussian partials of
// every lane, sums them over the warp (through shared memory: one 80-B row per lane,
// then lane c adds column c), and 19 lanes flush the warp's contribution with ONE
// vector RED.ADD.F32 into a contiguous 80-B accumulator row (ACC_STRIDE floats per Gaussian).
//
// The kernel is bound by instruction issue (ncu: 74-83 % of the issue slots, DRAM 1 %), and a quarter of its
// instructions per (warp, Gaussian) visit were that reduction plus the per-entry loop overhead
// (profiles/r2_ncu_summary.md: 84 of 342).  Every thread therefore owns TWO pixels -- a warp covers an 8x8 pixel
// block, a CTA of four warps the 16x16 tile (surfel_composite.cuh: pair_pixel) -- so one reduction, one RED and one
// trip through the entry loop serve 64 pixels.  The reduction was a 21-shuffle recursive-halving reduce-scatter in
// registers until the end of round 2 (102 instructions with its selects); the shared-memory form is ~75 and took
// K4 from 1.52 to 1.49 ms on the bench view.
#include "surfel_composite.cuh"

namespace b200gs {

struct CompBwdArgs {
  int W, H, gx;
  int tiles;                  // depth slices: ranges / tile_base of slice k start at k * (tiles + 1) / k * tiles
  const int* plan;            // device copy of the forward's slice plan: [1] = number of slices
  const uint32_t* tile_base;  // [slices+1][tiles] list position at which slice k starts in every tile
  const uint2* ranges;        // [slices][tiles+1] ranges into the COMPACTED instance list
  const uint32_t* point_list; // compacted ids
  const uint32_t* list_pos;   // position of every compacted entry in the tile's full list (ascending)
  const SurfelRec* rec;
  const uint8_t* hit;         // per compacted entry
  const float* bg;
  const float* accum;         // [3,N] final T, M1, M2
  const uint32_t* n_contrib;  // [2,N]
  const float* dL_dpix;       // [3,N]
  const float* dL_dothers;    // [8,N]
  const float* dL_drefl;      // [N] or NULL
  float* acc;                 // [P, ACC_STRIDE]
};

constexpr int BATCH = PAIR_THREADS;   // threads per CTA = entries staged per batch

// Everything the backward keeps per pixel: the forward's final values, the upstream gradients and the recurrences.
struct PixelBwd {
  float fx, fy;
  uint32_t last_contributor, median_contributor;     // last_contributor == 0 for pixels outside the image
  float T_final, T, final_D, final_D2, final_A;
  float dpix0, dpix1, dpix2, drefl, ddepth, daccum, dreg, dn0, dn1, dn2, dmedian, bg_dot;
  float acc_q, last_q, last_dL_dT, last_alpha;

  __device__ __forceinline__ void load(const CompBwdArgs& a, int pxi, int pyi) {
    const bool inside = pxi < a.W && pyi < a.H;
    fx = (float)pxi; fy = (float)pyi;
    const size_t N = (size_t)a.W * a.H;
    const size_t pix_id = (size_t)a.W * pyi + pxi;
    T_final = inside ? a.accum[pix_id] : 0.0f;
    T = T_final;
    last_contributor = inside ? a.n_contrib[pix_id] : 0u;
    median_contributor = inside ? a.n_contrib[pix_id + N] : 0u;
    final_D = inside ? a.accum[pix_id + N] : 0.0f;
    final_D2 = inside ? a.accum[pix_id + 2 * N] : 0.0f;
    final_A = 1.0f - T_final;
    dpix0 = dpix1 = dpix2 = drefl = ddepth = daccum = dreg = dn0 = dn1 = dn2 = dmedian = 0.0f;
    if (inside) {
      dpix0 = a.dL_dpix[pix_id]; dpix1 = a.dL_dpix[pix_id + N]; dpix2 = a.dL_dpix[pix_id + 2 * N];
      drefl = a.dL_drefl ? a.dL_drefl[pix_id] : 0.0f;
      ddepth = a.dL_dothers[pix_id + 0 * N];
      daccum = a.dL_dothers[pix_id + 1 * N];
      dn0 = a.dL_dothers[pix_id + 2 * N]; dn1 = a.dL_dothers[pix_id + 3 * N]; dn2 = a.dL_dothers[pix_id + 4 * N];
      dmedian = a.dL_dothers[pix_id + 5 * N];
      dreg = a.dL_dothers[pix_id + 6 * N];
    }
    bg_dot = 0.0f;
    bg_dot += a.bg[0] * dpix0;
    bg_dot += a.bg[1] * dpix1;
    bg_dot += a.bg[2] * dpix2;
    acc_q = last_q = last_dL_dT = last_alpha = 0.0f;
  }

  // One contributing (pixel, Gaussian) pair: advances the recurrences and ADDS the pair's 19 partials to g.
  __device__ __forceinline__ void contribute(const SurfelHit& h, const float4& q3, const float4& q4, uint32_t contributor,
                                             float* g) {
    const float alpha = h.alpha, G = h.G;
    // One reciprocal for (1 - alpha) and one for p.z instead of the reference's divisions
    // (backward.cu:341,:412,:437-438): nothing cancels behind them.
    // T is a running product over the tile's list: MUFU.RCP plus one Newton step (within half a unit in the last place)
    const float inv_1ma = rcp_newton(__fsub_rn(1.0f, alpha));
    T = T * inv_1ma;
    const float w = alpha * T;
    const float om = 1.0f - last_alpha;
    const float c_d = h.depth;
    // Colour, reflection strength, depth, alpha and normal (backward.cu:346-364,:392-402) all
    // enter dL_dalpha as (value - value accumulated behind) * upstream gradient.  The recurrence
    // is linear, so the nine per-channel recurrences of the reference collapse into ONE on
    // q = sum_ch value_ch * upstream_ch.
    const float qv = q4.x * dpix0 + q4.y * dpix1 + q4.z * dpix2 + q3.w * drefl + c_d * ddepth + daccum +
                     q3.x * dn0 + q3.y * dn1 + q3.z * dn2;
    acc_q = last_alpha * last_q + om * acc_q;
    last_q = qv;
    float dL_dalpha = qv - acc_q;
    g[12] += w * dpix0; g[13] += w * dpix1; g[14] += w * dpix2;
    g[15] += w * drefl;
    g[16] += w * dn0; g[17] += w * dn1; g[18] += w * dn2;
    // distortion (:366-390).  final_D2 + m^2 final_A - 2 m final_D and m final_A - final_D cancel in fp32, so
    // m_d and both expressions are written in exactly the rounding shape of the reference build (SASS of
    // renderCUDA<3>, nvcc 12.9: IEEE -0.2/c_d, +1, *far/(far-near); fma(2m, -D, fma(A, m*m, D2)); fma(A, m, -D)):
    // with the forward's final_* bit-identical, dL_dweight and the last_dL_dT recurrence are then the reference's
    // bit for bit, and only T (reciprocal instead of division, nothing cancels there) differs in the last place.
    const float m_d = ndc_depth(c_d);
    const float dmd_dd = (FAR_N * NEAR_N) * rcp_approx(__fmul_rn(c_d, __fmul_rn(c_d, FAR_N - NEAR_N)));   // a plain factor: no cancellation behind it
    float dL_dz = 0.0f;
    if (contributor == median_contributor - 1u) dL_dz = dmedian;
    const float dL_dweight =
        __fmul_rn(dreg, __fmaf_rn(__fadd_rn(m_d, m_d), -final_D, __fmaf_rn(final_A, __fmul_rn(m_d, m_d), final_D2)));
    dL_dalpha += __fsub_rn(dL_dweight, last_dL_dT);
    last_dL_dT = __fmaf_rn(alpha, dL_dweight, __fmul_rn(last_dL_dT, __fsub_rn(1.0f, alpha)));
    const float dL_dmd = __fmul_rn(dreg, __fmul_rn(__fadd_rn(w, w), __fmaf_rn(final_A, m_d, -final_D)));
    dL_dz = __fmaf_rn(dL_dmd, dmd_dd, dL_dz);

    dL_dalpha *= T;
    last_alpha = alpha;
    dL_dalpha += (-T_final * inv_1ma) * bg_dot;  // :411-414
    const float dL_dG = h.opa * dL_dalpha;
    dL_dz += w * ddepth;
    if (h.rho3d <= h.rho2d) {  // :423-453
      const float dsx = dL_dG * -G * h.sx + dL_dz * h.Twx;
      const float dsy = dL_dG * -G * h.sy + dL_dz * h.Twy;
      const float dsx_pz = dsx * h.rpz, dsy_pz = dsy * h.rpz;
      const float dpx = dsx_pz, dpy = dsy_pz, dpz = -(dsx_pz * h.sx + dsy_pz * h.sy);
      // dL_dk = cross(l, dL_dp), dL_dl = cross(dL_dp, k)
      const float dkx = h.ly * dpz - h.lz * dpy, dky = h.lz * dpx - h.lx * dpz, dkz = h.lx * dpy - h.ly * dpx;
      const float dlx = dpy * h.kz - dpz * h.ky, dly = dpz * h.kx - dpx * h.kz, dlz = dpx * h.ky - dpy * h.kx;
      g[0] -= dkx; g[1] -= dky; g[2] -= dkz;
      g[3] -= dlx; g[4] -= dly; g[5] -= dlz;
      g[6] += fx * dkx + fy * dlx + dL_dz * h.sx;
      g[7] += fx * dky + fy * dly + dL_dz * h.sy;
      g[8] += fx * dkz + fy * dlz + dL_dz * 1.0f;
    } else {  // :454-464
      g[9] += dL_dG * (-G * FILTER_INV_SQUARE * h.dx);
      g[10] += dL_dG * (-G * FILTER_INV_SQUARE * h.dy);
      g[6] += h.sx * dL_dz;
      g[7] += h.sy * dL_dz;
      g[8] += dL_dz;
    }
    g[11] += G * dL_dalpha;  // :467
  }
};

// Register allocation bounded for 4 resident CTAs of 128 threads per SM.
__global__ void __launch_bounds__(BATCH, 4) surfel_composite_bwd_kernel(const CompBwdArgs a) {
  __shared__ SurfelRec s_rec[2][BATCH];
  __shared__ int s_id[2][BATCH];
  __shared__ uint32_t s_pos[2][BATCH];
  __shared__ uint8_t s_hit[2][BATCH];   // bit b: the entry can reach alpha >= 1/255 in 8x4 pixel block b of the tile
  __shared__ int s_n;                   // live prefix of the slice's compacted list
  __shared__ uint32_t s_last[BATCH / 32];
  __shared__ uint32_t s_tile_last;
  __shared__ __align__(16) float s_red[BATCH / 32][32][ACC_STRIDE];   // per warp: the lanes' partials of one visit

  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int tile = blockIdx.x;
  const int tile_x = tile % a.gx, tile_y = tile / a.gx;
  int lx, ly;
  pair_pixel(t, lx, ly);
  PixelBwd px[2];
  px[0].load(a, tile_x * TILE_X + lx, tile_y * TILE_Y + ly);
  px[1].load(a, tile_x * TILE_X + lx, tile_y * TILE_Y + ly + 4);
  const unsigned warp_bits = pair_hit_bits(warp);       // the two 8x4 blocks of this warp's 8x8 pixels

  // Entries at or behind the last contributor of every pixel of the tile are never used
  // (backward.cu:290: `contributor >= last_contributor` -> continue).
  const uint32_t warp_last = __reduce_max_sync(0xffffffffu, max(px[0].last_contributor, px[1].last_contributor));
  if (lane == 0) s_last[warp] = warp_last;
  __syncthreads();
  if (t == 0) {
    uint32_t tl = 0;
#pragma unroll
    for (int w = 0; w < BATCH / 32; ++w) tl = max(tl, s_last[w]);
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
    const int rounds = (n + BATCH - 1) / BATCH;
    if (n == 0) continue;

    unsigned hit_cur = 0u, hit_next = 0u;
    uint32_t pos_next = 0u;
    auto stage = [&](int buf, int id) {
      hit_cur = hit_next;
      if (id >= 0) {
        s_pos[buf][t] = base + pos_next;       // 0-based position in the tile's full list
        const SurfelRec* g = a.rec + id;
        SurfelRec* d = &s_rec[buf][t];
#pragma unroll
        for (int k = 0; k < 5; ++k) cp_async16(&d->q[k], &g->q[k]);
        s_id[buf][t] = id;
      }
      cp_async_commit();
    };
    int id_next = -1;
    if (t < n) { id_next = (int)pl[n - 1 - t]; hit_next = ph[n - 1 - t]; pos_next = pp[n - 1 - t]; }
    stage(0, id_next);
    id_next = -1; hit_next = 0u;
    if (BATCH + t < n) { const int q = n - 1 - (BATCH + t); id_next = (int)pl[q]; hit_next = ph[q]; pos_next = pp[q]; }

    for (int i = 0; i < rounds; ++i) {
      cp_async_wait_all();
      s_hit[i & 1][t] = (uint8_t)hit_cur;
      __syncthreads();
      if (i + 1 < rounds) {
        stage((i + 1) & 1, id_next);
        const int nx = (i + 2) * BATCH + t;
        id_next = -1; hit_next = 0u;
        if (nx < n) { const int q = n - 1 - nx; id_next = (int)pl[q]; hit_next = ph[q]; pos_next = pp[q]; }
      }
      const int buf = i & 1;
      const int count = min(BATCH, n - i * BATCH);
      const SurfelRec* rb = s_rec[buf];
#pragma unroll 1
      for (int r = 0; r < (count + 31) >> 5; ++r) {
        // entries of other warp blocks, and entries behind every pixel's last contributor, are skipped
        const int e = r * 32 + lane;
        unsigned m = __ballot_sync(0xffffffffu, e < count && (s_hit[buf][e] & warp_bits) && s_pos[buf][e] < warp_last);
        while (m) {
          const int j = r * 32 + __ffs(m) - 1;
          m &= m - 1;
          const uint32_t contributor = s_pos[buf][j];          // 0-based position in the tile's full list
          const float4 q0 = rb[j].q[0], q1 = rb[j].q[1], q2 = rb[j].q[2], q3 = rb[j].q[3], q4 = rb[j].q[4];

          float g[20];
#pragma unroll
          for (int k = 0; k < 20; ++k) g[k] = 0.0f;
          SurfelHit h[2];
          bool c[2];
#pragma unroll
          for (int p = 0; p < 2; ++p)      // both pixels' tests first, straight-line: two chains in flight
            c[p] = surfel_eval_straight<true>(q0, q1, q2, px[p].fx, px[p].fy, h[p]) && contributor < px[p].last_contributor;
#pragma unroll
          for (int p = 0; p < 2; ++p)
            if (c[p]) px[p].contribute(h[p], q3, q4, contributor, g);

          // Sum over the warp through shared memory: every lane stores its 20 partials as one 80-B row (five 16-B
          // stores, conflict-free: the rows of eight lanes cover all 32 banks), then lane c adds column c of the 32
          // rows (conflict-free: the lanes of one load read consecutive words).  5 STS + 32 LDS + 33 FADD against the
          // 21 shuffles + 20 adds + 42 selects of a recursive-halving reduce-scatter in registers.
          {
            float4* row = reinterpret_cast<float4*>(s_red[warp][lane]);
#pragma unroll
            for (int k = 0; k < 5; ++k) row[k] = make_float4(g[4 * k], g[4 * k + 1], g[4 * k + 2], g[4 * k + 3]);
            __syncwarp();
            float s0 = 0.0f, s1 = 0.0f;
            if (lane < ACC_STRIDE) {
#pragma unroll
              for (int r = 0; r < 32; r += 2) { s0 += s_red[warp][r][lane]; s1 += s_red[warp][r + 1][lane]; }
            }
            __syncwarp();
            // (16-B column loads by lanes 4 q + g over the rows 4 k + g, then two butterflies: 1.63 ms against 1.49 --
            // bank conflicts of the strided 16-B loads)
            if (lane < 19) atomicAdd(a.acc + (size_t)s_id[buf][j] * ACC_STRIDE + lane, s0 + s1);
          }
        }
      }
    }
    cp_async_wait_all();
  }   // slices
}

int surfel_composite_backward(const b200gs_surfel_bwd& f, const SurfelGeom& g, const SurfelImage& im, const Binning& b,
                              float* acc, cudaStream_t s) {
  CompBwdArgs a;
  a.W = f.width; a.H = f.height;
  a.gx = (f.width + TILE_X - 1) / TILE_X;
  const int gy = (f.height + TILE_Y - 1) / TILE_Y;
  a.tiles = a.gx * gy; a.plan = g.plan; a.tile_base = im.tile_base;
  a.ranges = im.cranges; a.point_list = b.comp_id; a.list_pos = b.comp_pos; a.rec = g.rec; a.hit = b.comp_hit; a.bg = f.background;
  a.accum = im.accum; a.n_contrib = im.n_contrib;
  a.dL_dpix = f.dL_dout_color; a.dL_dothers = f.dL_dout_others; a.dL_drefl = f.dL_dout_refl;
  a.acc = acc;
  StageScope _t(ST_COMPOSITE_BWD, s);
  B200GS_CUDA(cudaMemsetAsync(acc, 0, (size_t)f.P * ACC_STRIDE * sizeof(float), s));
  surfel_composite_bwd_kernel<<<a.gx * gy, BATCH, 0, s>>>(a);
  B200GS_LAUNCHED("surfel_composite_bwd_kernel", f.debug, s);
  return B200GS_OK;
}

}  // namespace b200gs
