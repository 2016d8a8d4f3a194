// binning.cu -- K2: prefix sum of tiles_touched, (tile|depth, id) instance
// emission, 64-bit stable sort and per-tile ranges.
//
// Restates rasterizer_impl.cu:70-111 (duplicateWithKeys), :116-138
// (identifyTileRanges), :35-50 (getHigherMsb) and the two CUB calls at :282 and
// :308-313 of S/cuda_rasterizer/ (the same lines +-2 in G/).  Results are
// integer and must be bit-identical to the reference: keys are
// (tile_id << 32) | float_bits(depth), sorted stably on bits [0, 32+bit).
#include "common.cuh"
#include "radix_sort.cuh"

namespace b200gs {

// ------------------------------------------------------------------ scan
// Single-pass inclusive prefix sum with decoupled look-back.  Tile order is
// taken from a ticket counter so a tile only ever waits on tiles that already
// started.  state[t] = (flag << 32) | value, flag 1 = tile aggregate, 2 = inclusive prefix.
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

__global__ void __launch_bounds__(SCAN_THREADS) scan_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                            int n, unsigned long long* state, uint32_t* ticket) {
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  __shared__ uint32_t s_prefix;
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  const int base = tile * SCAN_TILE + threadIdx.x * SCAN_ITEMS;

  uint32_t v[SCAN_ITEMS];
  if (base + SCAN_ITEMS <= n) {
    const uint4* p = reinterpret_cast<const uint4*>(in + base);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS / 4; ++k) {
      const uint4 q = p[k];
      v[4 * k] = q.x; v[4 * k + 1] = q.y; v[4 * k + 2] = q.z; v[4 * k + 3] = q.w;
    }
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = (base + k < n) ? in[base + k] : 0u;
  }
  uint32_t sum = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) { sum += v[k]; v[k] = sum; }

  // block-wide exclusive scan of the per-thread sums
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += t;
  }
  if (lane == 31) s_warp[warp] = inc;
  __syncthreads();
  uint32_t warp_off = 0, block_sum = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; ++w) {
    const uint32_t t = s_warp[w];
    if (w < warp) warp_off += t;
    block_sum += t;
  }
  const uint32_t thread_excl = warp_off + inc - sum;

  if (warp == 0) {
    volatile unsigned long long* st = state;
    if (lane == 0) st[tile] = ((tile == 0 ? 2ull : 1ull) << 32) | block_sum;
    uint32_t excl = 0;
    if (tile > 0) {
      int look = (int)tile - 1;
      while (true) {
        const int t = look - lane;
        unsigned long long w = 2ull << 32;  // tiles before 0: inclusive prefix 0
        if (t >= 0) { do { w = st[t]; } while ((w >> 32) == 0ull); }
        const bool full = (w >> 32) == 2ull;
        const unsigned ballot = __ballot_sync(0xffffffffu, full);
        const int first = __ffs(ballot) - 1;  // -1 if none
        uint32_t val = (first < 0 || lane <= first) ? (uint32_t)w : 0u;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
        excl += val;
        if (first >= 0) break;
        look -= 32;
      }
      if (lane == 0) st[tile] = (2ull << 32) | (unsigned long long)(excl + block_sum);
    }
    if (lane == 0) s_prefix = excl;
  }
  __syncthreads();
  const uint32_t off = s_prefix + thread_excl;
  if (base + SCAN_ITEMS <= n) {
    uint4* p = reinterpret_cast<uint4*>(out + base);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS / 4; ++k)
      p[k] = make_uint4(v[4 * k] + off, v[4 * k + 1] + off, v[4 * k + 2] + off, v[4 * k + 3] + off);
  } else {
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k)
      if (base + k < n) out[base + k] = v[k] + off;
  }
}

size_t scan_state_words(int n) {
  const size_t tiles = ((size_t)n + SCAN_TILE - 1) / SCAN_TILE;
  return 2 * (tiles + 2);
}

int inclusive_scan_u32(const uint32_t* in, uint32_t* out, int n, uint32_t* state, cudaStream_t s, bool debug) {
  if (n <= 0) return B200GS_OK;
  const int tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  StageScope _t(ST_SCAN, s);
  B200GS_CUDA(cudaMemsetAsync(state, 0, scan_state_words(n) * sizeof(uint32_t), s));
  unsigned long long* st = reinterpret_cast<unsigned long long*>(state);
  scan_kernel<<<tiles, SCAN_THREADS, 0, s>>>(in, out, n, st + 1, state);  // word 0 = ticket
  B200GS_LAUNCHED("scan_kernel", debug, s);
  return B200GS_OK;
}

// ------------------------------------------------------------------ emission
// duplicateWithKeys (rasterizer_impl.cu:70-111): rect rows outer, columns inner,
// key = tile << 32 | depth bits, value = Gaussian id.  One thread per Gaussian
// for small rects; rects of >= 32 tiles are handed to the whole warp.
__global__ void __launch_bounds__(256) duplicate_kernel(int P, const float* __restrict__ depths,
                                                        const uint2* __restrict__ rects,
                                                        const uint32_t* __restrict__ offsets, uint64_t* __restrict__ keys,
                                                        uint32_t* __restrict__ vals, int gx) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  uint32_t off = 0, cnt = 0;
  if (idx < P) {
    off = idx == 0 ? 0u : offsets[idx - 1];
    cnt = offsets[idx] - off;
  }
  uint2 rc = make_uint2(0, 0);
  uint32_t dbits = 0;
  if (cnt > 0) { rc = rects[idx]; dbits = __float_as_uint(depths[idx]); }
  const bool big = cnt >= 32;
  if (cnt > 0 && !big) {
    const int x0 = rc.x & 0xffff, x1 = rc.x >> 16, y0 = rc.y & 0xffff, y1 = rc.y >> 16;
    for (int y = y0; y < y1; ++y)
      for (int x = x0; x < x1; ++x) {
        keys[off] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | dbits;
        vals[off] = (uint32_t)idx;
        ++off;
      }
  }
  unsigned todo = __ballot_sync(0xffffffffu, big);
  while (todo) {
    const int src = __ffs(todo) - 1;
    todo &= todo - 1;
    const uint32_t rx = __shfl_sync(0xffffffffu, rc.x, src), ry = __shfl_sync(0xffffffffu, rc.y, src);
    const uint32_t o = __shfl_sync(0xffffffffu, off, src), c = __shfl_sync(0xffffffffu, cnt, src);
    const uint32_t db = __shfl_sync(0xffffffffu, dbits, src);
    const int x0 = rx & 0xffff, x1 = rx >> 16, y0 = ry & 0xffff;
    const int w = x1 - x0;
    const uint32_t id = (uint32_t)(idx - lane + src);
    for (uint32_t i = lane; i < c; i += 32) {
      const int y = y0 + (int)(i / w), x = x0 + (int)(i % w);
      keys[o + i] = ((uint64_t)(uint32_t)(y * gx + x) << 32) | db;
      vals[o + i] = id;
    }
  }
}

// identifyTileRanges (rasterizer_impl.cu:116-138)
__global__ void __launch_bounds__(256) tile_ranges_kernel(int L, const uint64_t* __restrict__ keys, uint2* ranges) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= L) return;
  const uint32_t cur = (uint32_t)(keys[idx] >> 32);
  if (idx == 0) {
    ranges[cur].x = 0;
  } else {
    const uint32_t prev = (uint32_t)(keys[idx - 1] >> 32);
    if (cur != prev) { ranges[prev].y = idx; ranges[cur].x = idx; }
  }
  if (idx == L - 1) ranges[cur].y = L;
}

// ------------------------------------------------------------------ tile hits
// For every sorted (tile, Gaussian) instance: which of the tile's eight 8x4 warp blocks (bit w:
// column w&1, row w>>1, the layout of tile_pixel()) can the Gaussian reach alpha >= 1/255 in?
// Block = pixel centres [8c, 8c+7] x [4r, 4r+3] relative to the tile origin, tested as the
// rectangle dilated by half a pixel against the CullRec ellipse (exact convex test: the x-extent
// of ellipse /\ row band is attained at the ellipse's extreme point if that lies in the band, else
// on one of the band's two boundary lines) and against the low-pass disk.  One byte per instance;
// the list itself is untouched (the reference's list order and positions are what n_contrib counts).
__global__ void __launch_bounds__(256) tile_hits_kernel(int R, const uint64_t* __restrict__ keys,
                                                        const uint32_t* __restrict__ point_list,
                                                        const CullRec* __restrict__ cull, int gx,
                                                        uint8_t* __restrict__ hit) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= R) return;
  const uint32_t tile = (uint32_t)(keys[i] >> 32);
  const uint32_t id = point_list[i];
  const float4 ca = __ldg(&cull[id].a), cb = __ldg(&cull[id].b);
  unsigned mask;
  if (ca.w < 0.0f) {
    mask = (ca.w == CULL_NOTHING) ? 0u : 0xFFu;
  } else {
    const float ox = (float)((tile % (uint32_t)gx) * TILE_X) - 0.5f - ca.x;  // left edge of column 0, relative to e
    const float oy = (float)((tile / (uint32_t)gx) * TILE_Y) - 0.5f - ca.y;  // top edge of row band 0
    const float hx = ca.z, hy2 = ca.w, sl = cb.x, k = cb.y, dyR = cb.z, r2 = cb.w;
    const float INF = __int_as_float(0x7f800000);
    float lo[5], hi[5];
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const float dy = oy + 4.0f * j;
      const float rem = hy2 - dy * dy;
      const float h = k * sqrtf(fmaxf(rem, 0.0f));
      lo[j] = rem >= 0.0f ? sl * dy - h : INF;
      hi[j] = rem >= 0.0f ? sl * dy + h : -INF;
    }
    // disk: squared distance from e to the column / row intervals
    float ddx2[2], ddy2[4];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const float d = fmaxf(fmaxf(ox + 8.0f * c, -(ox + 8.0f * (c + 1))), 0.0f);
      ddx2[c] = d * d;
    }
    mask = 0u;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const float ya = oy + 4.0f * r, yb = oy + 4.0f * (r + 1);
      const float d = fmaxf(fmaxf(ya, -yb), 0.0f);
      ddy2[r] = d * d;
      const float xmax = (dyR >= ya && dyR <= yb) ? hx : fmaxf(hi[r], hi[r + 1]);
      const float xmin = (-dyR >= ya && -dyR <= yb) ? -hx : fminf(lo[r], lo[r + 1]);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const float xa = ox + 8.0f * c, xb = ox + 8.0f * (c + 1);
        const bool ell = xmin <= xb && xmax >= xa;
        const bool disk = ddx2[c] + ddy2[r] <= r2;
        if (ell || disk) mask |= 1u << (r * 2 + c);
      }
    }
  }
  hit[i] = (uint8_t)mask;
}

int surfel_tile_hits(const CullRec* cull, const Binning& b, int R, int W, cudaStream_t s, bool debug) {
  const int gx = (W + TILE_X - 1) / TILE_X;
  StageScope _t(ST_TILE_HITS, s);
  tile_hits_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, b.keys, b.point_list, cull, gx, b.hit);
  B200GS_LAUNCHED("tile_hits_kernel", debug, s);
  return B200GS_OK;
}

// getHigherMsb (rasterizer_impl.cu:35-50)
static uint32_t higher_msb(uint32_t n) {
  uint32_t msb = sizeof(n) * 4, step = msb;
  while (step > 1) {
    step /= 2;
    if (n >> msb) msb += step; else msb -= step;
  }
  if (n >> msb) msb++;
  return msb;
}

Binning Binning::carve(void* base, size_t R, size_t* bytes) {
  Carver c(base);
  Binning b;
  b.point_list = c.take<uint32_t>(R);
  b.point_list_unsorted = c.take<uint32_t>(R);
  b.keys = c.take<uint64_t>(R);
  b.keys_unsorted = c.take<uint64_t>(R);
  b.hit = c.take<uint8_t>(R);
  b.sort_temp_bytes = radix_sort_temp_bytes(R);
  b.sort_temp = c.take<char>(b.sort_temp_bytes);
  if (bytes) *bytes = c.used + 128;
  return b;
}

int bin_and_sort(const float* depths, const uint2* rects, const uint32_t* offsets, int P, int R, int W, int H,
                 const Binning& b, uint2* ranges, cudaStream_t s, bool debug) {
  const int gx = (W + TILE_X - 1) / TILE_X, gy = (H + TILE_Y - 1) / TILE_Y;
  B200GS_CUDA(cudaMemsetAsync(ranges, 0, (size_t)gx * gy * sizeof(uint2), s));
  if (R <= 0) return B200GS_OK;
  {
  StageScope _t(ST_DUPLICATE, s);
  duplicate_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, depths, rects, offsets, b.keys_unsorted, b.point_list_unsorted, gx);
  B200GS_LAUNCHED("duplicate_kernel", debug, s);
  }
  const int end_bit = 32 + (int)higher_msb((uint32_t)(gx * gy));
  int rc;
  {
  StageScope _t(ST_SORT, s);
  rc = radix_sort_pairs(b.keys_unsorted, b.keys, b.point_list_unsorted, b.point_list, R, end_bit, b.sort_temp,
                            b.sort_temp_bytes, s, debug);
  }
  if (rc) return rc;
  StageScope _t(ST_RANGES, s);
  tile_ranges_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, b.keys, ranges);
  B200GS_LAUNCHED("tile_ranges_kernel", debug, s);
  return B200GS_OK;
}

}  // namespace b200gs
