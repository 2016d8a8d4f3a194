// binning.cu -- K2: depth order of the Gaussians, prefix sum of tiles_touched, (tile, id)
// instance emission, stable sort by tile, per-tile ranges, sub-tile hit masks + compaction.
//
// Restates rasterizer_impl.cu:70-111 (duplicateWithKeys), :116-138
// (identifyTileRanges), :35-50 (getHigherMsb) and the two CUB calls at :282 and
// :308-313 of S/cuda_rasterizer/ (the same lines +-2 in G/).  Results are
// integer and must be bit-identical to the reference, whose keys are
// (tile_id << 32) | float_bits(depth), sorted stably on bits [0, 32+bit) from an emission in
// (Gaussian id, rect row, rect column) order.  Here the same order is produced as: stable sort
// of the Gaussians by depth bits (ties keep id order) -> emission in that order -> stable sort
// of the instances by tile id (radix_sort.cuh).  The 64-bit keys are only materialised by
// b200gs_inspect_binning for the parity tests.
#include "common.cuh"
#include "radix_sort.cuh"
#include "cull_rows.cuh"

namespace b200gs {

// ------------------------------------------------------------------ scan
// Single-pass inclusive prefix sum with decoupled look-back.  Tile order is
// taken from a ticket counter so a tile only ever waits on tiles that already
// started.  state[t] = (flag << 32) | value, flag 1 = tile aggregate, 2 = inclusive prefix.
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 16;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

// `gather` (optional): the scan runs over in[gather[i]].
__global__ void __launch_bounds__(SCAN_THREADS) scan_kernel(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                            int n, unsigned long long* state, uint32_t* ticket,
                                                            const uint32_t* __restrict__ gather) {
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_warp[SCAN_THREADS / 32];
  __shared__ uint32_t s_prefix;
  if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
  __syncthreads();
  const uint32_t tile = s_tile;
  const int base = tile * SCAN_TILE + threadIdx.x * SCAN_ITEMS;

  uint32_t v[SCAN_ITEMS];
  if (gather != nullptr) {
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) v[k] = (base + k < n) ? in[gather[base + k]] : 0u;
  } else if (base + SCAN_ITEMS <= n) {
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

int inclusive_scan_u32(const uint32_t* in, uint32_t* out, int n, uint32_t* state, cudaStream_t s, bool debug,
                       const uint32_t* gather) {
  if (n <= 0) return B200GS_OK;
  const int tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  StageScope _t(ST_SCAN, s);
  B200GS_CUDA(cudaMemsetAsync(state, 0, scan_state_words(n) * sizeof(uint32_t), s));
  unsigned long long* st = reinterpret_cast<unsigned long long*>(state);
  scan_kernel<<<tiles, SCAN_THREADS, 0, s>>>(in, out, n, st + 1, state, gather);  // word 0 = ticket
  B200GS_LAUNCHED("scan_kernel", debug, s);
  return B200GS_OK;
}

// ------------------------------------------------------------------ depth order
// Sort key of a Gaussian: the bits of its view depth (what the reference puts in the low key
// word); culled Gaussians get the largest key -- they emit nothing, so where they land is moot.
__global__ void __launch_bounds__(256) depth_keys_kernel(int P, const float* __restrict__ depths,
                                                         const uint32_t* __restrict__ tiles_touched,
                                                         uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  keys[idx] = tiles_touched[idx] > 0u ? __float_as_uint(depths[idx]) : 0xffffffffu;
  vals[idx] = (uint32_t)idx;
}

DepthOrder DepthOrder::carve(Carver& c, size_t P) {
  DepthOrder o;
  o.keys_a = c.take<uint32_t>(P);
  o.vals_a = c.take<uint32_t>(P);
  o.keys_b = c.take<uint32_t>(P);
  o.vals_b = c.take<uint32_t>(P);
  o.ws_bytes = radix_workspace_bytes(P, 4);
  o.ws = c.take<char>(o.ws_bytes);
  return o;
}

// -> o.vals_a = Gaussian ids in (depth bits, id) order; point_offsets = inclusive prefix sum of
// tiles_touched in that order (the last entry is the instance count R).
int depth_order_and_scan(const float* depths, const uint32_t* tiles_touched, int P, const DepthOrder& o,
                         uint32_t* point_offsets, uint32_t* scan_state, cudaStream_t s, bool debug) {
  if (P <= 0) return B200GS_OK;
  {
    StageScope _t(ST_DEPTH_SORT, s);
    int rc = radix_workspace_clear(o.ws, (size_t)P, 4, s);
    if (rc) return rc;
    depth_keys_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, depths, tiles_touched, o.keys_a, o.vals_a);
    B200GS_LAUNCHED("depth_keys_kernel", debug, s);
    const RadixPlan plan = radix_plan(32);   // 4 passes: result back in (keys_a, vals_a)
    rc = radix_histogram(o.keys_a, P, plan, o.ws, s, debug);
    if (rc) return rc;
    rc = radix_sort_pairs_u32(o.keys_a, o.vals_a, o.keys_b, o.vals_b, P, plan, o.ws, s, debug);
    if (rc) return rc;
  }
  return inclusive_scan_u32(tiles_touched, point_offsets, P, scan_state, s, debug, o.vals_a);
}

// ------------------------------------------------------------------ emission
// duplicateWithKeys (rasterizer_impl.cu:70-111): rect rows outer, columns inner, value =
// Gaussian id; the key is the tile id only (the depth order is already in the emission order).
// Thread i handles the i-th Gaussian in depth order; one thread per Gaussian for small rects,
// rects of >= 32 tiles are handed to the whole warp.
__global__ void __launch_bounds__(256) duplicate_kernel(int P, const uint32_t* __restrict__ order,
                                                        const uint2* __restrict__ rects,
                                                        const uint32_t* __restrict__ offsets, uint32_t* __restrict__ keys,
                                                        uint32_t* __restrict__ vals, int gx) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  uint32_t off = 0, cnt = 0, id = 0;
  if (i < P) {
    off = i == 0 ? 0u : offsets[i - 1];
    cnt = offsets[i] - off;
  }
  uint2 rc = make_uint2(0, 0);
  if (cnt > 0) { id = order[i]; rc = rects[id]; }
  const bool big = cnt >= 32;
  if (cnt > 0 && !big) {
    const int x0 = rc.x & 0xffff, x1 = rc.x >> 16, y0 = rc.y & 0xffff, y1 = rc.y >> 16;
    for (int y = y0; y < y1; ++y)
      for (int x = x0; x < x1; ++x) {
        keys[off] = (uint32_t)(y * gx + x);
        vals[off] = id;
        ++off;
      }
  }
  unsigned todo = __ballot_sync(0xffffffffu, big);
  while (todo) {
    const int src = __ffs(todo) - 1;
    todo &= todo - 1;
    const uint32_t rx = __shfl_sync(0xffffffffu, rc.x, src), ry = __shfl_sync(0xffffffffu, rc.y, src);
    const uint32_t o = __shfl_sync(0xffffffffu, off, src), c = __shfl_sync(0xffffffffu, cnt, src);
    const uint32_t gid = __shfl_sync(0xffffffffu, id, src);
    const int x0 = rx & 0xffff, x1 = rx >> 16, y0 = ry & 0xffff;
    const int w = x1 - x0;
    for (uint32_t k = lane; k < c; k += 32) {
      const int y = y0 + (int)(k / w), x = x0 + (int)(k % w);
      keys[o + k] = (uint32_t)(y * gx + x);
      vals[o + k] = gid;
    }
  }
}

// ------------------------------------------------------------------ culled emission
// Default mode of the surfel path (exact_lists = 0): instances whose tile the Gaussian
// cannot reach with alpha >= 1/255 (cull_rows.cuh) are never emitted, so everything downstream --
// sort, ranges, block masks, compositing -- works on the few per cent of the reference's list that
// can contribute.  Images, radii, weights and gradients are unchanged (the reference rejects those
// pairs pixel by pixel); the lists are a subsequence of the reference's.  The exact-lists mode
// keeps duplicate_kernel above and reproduces the reference's lists bit for bit.
// Work distribution: a Gaussian's cost is its number of tile rows, which varies from 1 to the whole
// image height, so one thread per Gaussian leaves most lanes idle (ncu: 8 of 32 active).  Instead
// every warp takes 32 Gaussians, parks their RowCull states in shared memory and deals the
// (Gaussian, tile row) pairs of all 32 out to its lanes, 32 rows at a time.  The count and the
// emission are the same template, so both run the same fp64 code on the same inputs and agree
// exactly on the number of instances.  Within a Gaussian the emitted tiles are in no particular
// order: the tile sort is stable across Gaussians (their slots are disjoint and in depth order),
// and a Gaussian appears at most once per tile.
//
// hist != nullptr: the digit histograms of the (<= 2) radix passes over the emitted tile ids are
// accumulated on the fly (shared-memory REDs, flushed per block), which saves the separate
// histogram pass over the keys (radix_histogram).
struct EmitHist {
  uint32_t* hist;          // [RS_MAX_PASSES][256] in the radix workspace, or nullptr
  int shift1;              // pass 0 digit = key & mask0, pass 1 digit = (key >> shift1) & mask1
  uint32_t mask0, mask1;
};

// Rects of at most this many tiles keep all their tiles.  With depth slices (surfel path) the far part of the list is
// never emitted, so a coarse count costs nothing there and the count pass shrinks (bench view: 0.215 ms at 32,
// 0.127 at 96, 0.082 at 256; step 3.65 / 3.61 / 3.57 ms); without slices (3DGS path) every emitted instance is
// sorted and masked, and the finer threshold wins (c3 through the 3DGS rasterizer: 3.67 ms at 32, 4.34 at 256).
constexpr int SMALL_RECT_SLICED = 256, SMALL_RECT_UNSLICED = 32;

struct KeptArgs {
  int P, gx;
  int g0, g1;                     // emit: the Gaussians [g0, g1) of the depth order (one depth slice)
  int lpg_shift;                  // emit: log2 of the lanes per Gaussian (0: 32 Gaussians per warp ... 5: one per warp)
  int nchunks;                    // groups of 256 threads' worth of work; the (capped) grid strides over them
  const uint8_t* done;            // emit, slices > 0: tiles whose pixels all saturated in an earlier slice ...
  uint32_t sentinel;              // ... get this key instead (sorts behind every tile, never read)
  const uint32_t* live;           // emit: number of unfinished tiles; 0 -> this slice is not needed at all
  RowCache rows;                  // rows.base == nullptr: no cache
  int small_rect;                 // rects of at most this many tiles are not tested row by row
  const uint32_t* tiles_touched;  // count: [P] in id order
  const uint32_t* order;          // emit: ids in depth order
  const uint32_t* offsets;        // emit: inclusive scan of the kept counts, depth order
  const uint2* rects;
  const CullRec* cull;
  uint32_t* kept;                 // count: out [P]
  uint32_t* keys;                 // emit: out
  uint32_t* vals;
  EmitHist eh;
};

template <bool EMIT>
__global__ void __launch_bounds__(256) kept_rows_kernel(const KeptArgs a) {
  __shared__ RowCull s_rc[256];
  __shared__ int4 s_box[256];            // x0, x1, y0, rows
  __shared__ uint32_t s_pref[256];       // inclusive prefix of rows within the warp
  __shared__ uint32_t s_run[256];        // count: kept tiles so far; emit: next free slot
  __shared__ uint2 s_ie[EMIT ? 256 : 1]; // emit: id, end of the Gaussian's slots
  __shared__ uint32_t s_rb[256];         // first cached row of the Gaussian (ROW_NONE: compute)
  __shared__ uint32_t s_h[2][EMIT ? RS_RADIX : 1];
  const int t = threadIdx.x, lane = t & 31, wb = t & ~31;
  // The front depth slices hold few Gaussians with very many rows each (splats close to the camera): with 32 of them
  // per warp only a handful of SMs would work, one row stream after the other.  There the Gaussians are spread out,
  // one per 2^lpg_shift lanes; the lanes in between own nothing and only help with the rows and tiles.
  if (EMIT && a.live != nullptr && *a.live == 0u) return;
  const bool hist = EMIT && a.eh.hist != nullptr;
  if (hist) { s_h[0][t] = 0u; s_h[1][t] = 0u; __syncthreads(); }
  // the grid is capped (slice_grid): a slice nobody needs is one small launch; everything below is warp-private
  for (int chunk = blockIdx.x; chunk < a.nchunks; chunk += gridDim.x) {
  __syncwarp();
  const int gt = chunk * 256 + t;
  const bool owner = !EMIT || (gt & ((1 << a.lpg_shift) - 1)) == 0;
  const int i = EMIT ? a.g0 + (gt >> a.lpg_shift) : gt;

  uint32_t id = 0, off = 0, end = 0;
  bool vis = false;
  if (owner && i < (EMIT ? a.g1 : a.P)) {
    if (EMIT) {
      off = i == 0 ? 0u : a.offsets[i - 1]; end = a.offsets[i];
      vis = end > off;
      if (vis) id = a.order[i];
    } else {
      id = (uint32_t)i;
      vis = a.tiles_touched[i] > 0u;
    }
  }
  int rows = 0;
  const bool cached = a.rows.base != nullptr;
  uint32_t rb = ROW_NONE;
  uint32_t run0 = off;                    // 0 when counting
  if (vis) {
    const uint2 rc = a.rects[id];
    const int x0 = rc.x & 0xffff, x1 = rc.x >> 16, y0 = rc.y & 0xffff, y1 = rc.y >> 16;
    // Small rects keep all their tiles: the row test costs ~1300 fp64-heavy instructions per Gaussian and takes an
    // ordinary splat of 5 x 5 tiles down to ~19; the emitted list only has to be a superset of the reachable
    // instances (tile_hits_kernel drops the rest) and, with depth slices, the far part of it is never emitted anyway.
    // What the test is for are the splats that cross the camera plane and would land in every tile of the screen.
    if ((x1 - x0) * (y1 - y0) <= a.small_rect) {
      if (EMIT) { s_rc[t].mode = 1; rows = y1 - y0; }
      else run0 = (uint32_t)((x1 - x0) * (y1 - y0));
    } else {
      if (EMIT && cached) rb = a.rows.base[id];
      if (rb == ROW_NONE) {                    // count, or a Gaussian whose rows did not fit the cache
        rowcull_init(a.cull[id], s_rc[t]);
        rows = (s_rc[t].mode == 0) ? 0 : (y1 - y0);
      } else {
        rows = rb == ROW_EMPTY ? 0 : (y1 - y0);
      }
    }
    s_box[t] = make_int4(x0, x1, y0, rows);
  }
  s_run[t] = run0;
  if (EMIT) s_ie[t] = make_uint2(id, end);
  uint32_t pref = (uint32_t)rows;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, pref, d);
    if (lane >= d) pref += v;
  }
  s_pref[t] = pref;
  const uint32_t total = __shfl_sync(0xffffffffu, pref, 31);
  // row cache: the count hands out `total` consecutive rows to the warp and notes where each Gaussian's rows start
  if (cached && !EMIT) {
    uint32_t wbase = 0u;
    if (lane == 0 && total) wbase = atomicAdd(a.rows.cursor, total);
    wbase = __shfl_sync(0xffffffffu, wbase, 0);
    if (total && wbase + total <= a.rows.capacity) rb = wbase + (pref - (uint32_t)rows);
    if (i < a.P) a.rows.base[i] = rows ? rb : ROW_EMPTY;
  }
  s_rb[t] = rows ? rb : ROW_NONE;
  __syncwarp();

  auto put = [&](uint32_t o, uint32_t key, uint32_t v) {
    if (a.done != nullptr && a.done[key]) key = a.sentinel;
    a.keys[o] = key; a.vals[o] = v;
    if (hist) { atomicAdd(&s_h[0][key & a.eh.mask0], 1u); atomicAdd(&s_h[1][(key >> a.eh.shift1) & a.eh.mask1], 1u); }
  };

  for (uint32_t k0 = 0; k0 < total; k0 += 32) {
    const uint32_t k = k0 + lane;
    const bool valid = k < total;
    int g = 0;                            // owner of row k: the first g with s_pref[g] > k
    int a1 = 0, b1 = -1, a2 = 0, b2 = -1, d0 = 0, d1 = -1, y = 0;
    uint32_t n = 0;
    if (valid) {
#pragma unroll
      for (int st = 16; st > 0; st >>= 1)
        if (s_pref[wb + g + st - 1] <= k) g += st;
      const int4 bx = s_box[wb + g];
      const uint32_t r = k - (s_pref[wb + g] - (uint32_t)bx.w);         // row of the Gaussian's rect
      y = bx.z + (int)r;
      const uint32_t grb = s_rb[wb + g];
      if (EMIT && grb != ROW_NONE) {                                    // the count pass already did this row
        const uint4 c = a.rows.rows[grb + r];
        a1 = (int)(c.x & 0xffffu); b1 = (int)(c.x >> 16) - 1; a2 = (int)(c.y & 0xffffu); b2 = (int)(c.y >> 16) - 1;
        d0 = (int)(c.z & 0xffffu); d1 = (int)(c.z >> 16) - 1;
        n = c.w;
      } else {
        rowcull_ranges(s_rc[wb + g], y, bx.x, bx.y, a1, b1, a2, b2, d0, d1);
        n = (uint32_t)rowcull_count(a1, b1, a2, b2, d0, d1);
        if (!EMIT && grb != ROW_NONE)
          a.rows.rows[grb + r] = make_uint4(((uint32_t)(a1 & 0xffff)) | ((uint32_t)((b1 + 1) & 0xffff) << 16),
                                            ((uint32_t)(a2 & 0xffff)) | ((uint32_t)((b2 + 1) & 0xffff) << 16),
                                            ((uint32_t)(d0 & 0xffff)) | ((uint32_t)((d1 + 1) & 0xffff) << 16), n);
      }
    }
    uint32_t o = 0;
    if (n) o = atomicAdd(&s_run[wb + g], n);
    if (EMIT) {
      // deal the tiles of these 32 rows out to the lanes: instance j of the iteration belongs to the
      // first row whose inclusive tile prefix exceeds j; consecutive lanes write consecutive slots
      uint32_t tp = n;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, tp, d);
        if (lane >= d) tp += v;
      }
      const uint32_t T = __shfl_sync(0xffffffffu, tp, 31);
      const uint32_t r1 = ((uint32_t)(a1 & 0xffff)) | ((uint32_t)((b1 + 1) & 0xffff) << 16);   // tile columns are 16-bit (rects)
      const uint32_t r2 = ((uint32_t)(a2 & 0xffff)) | ((uint32_t)((b2 + 1) & 0xffff) << 16);
      const uint32_t rd = ((uint32_t)(d0 & 0xffff)) | ((uint32_t)((d1 + 1) & 0xffff) << 16);
      const uint32_t rowbase = (uint32_t)(y * a.gx);
      for (uint32_t j0 = 0; j0 < T; j0 += 32) {
        const uint32_t j = j0 + lane;
        int ow = 0;
#pragma unroll
        for (int st = 16; st > 0; st >>= 1) {
          const uint32_t v = __shfl_sync(0xffffffffu, tp, ow + st - 1);
          if (v <= j) ow += st;
        }
        ow = min(ow, 31);
        const uint32_t o_tp = __shfl_sync(0xffffffffu, tp, ow), o_n = __shfl_sync(0xffffffffu, n, ow);
        const uint32_t o_o = __shfl_sync(0xffffffffu, o, ow);
        const uint32_t o_r1 = __shfl_sync(0xffffffffu, r1, ow), o_r2 = __shfl_sync(0xffffffffu, r2, ow);
        const uint32_t o_rd = __shfl_sync(0xffffffffu, rd, ow), o_base = __shfl_sync(0xffffffffu, rowbase, ow);
        const int o_g = __shfl_sync(0xffffffffu, g, ow);
        if (j < T) {
          int r = (int)(j - (o_tp - o_n));                       // index within the row's tiles
          const int ra1 = o_r1 & 0xffff, rb1 = (int)(o_r1 >> 16) - 1, ra2 = o_r2 & 0xffff, rb2 = (int)(o_r2 >> 16) - 1;
          const int len1 = max(0, rb1 - ra1 + 1), len2 = max(0, rb2 - ra2 + 1);
          int x;
          if (r < len1) x = ra1 + r;
          else if (r < len1 + len2) x = ra2 + (r - len1);
          else {                                                 // the r-th tile of the disk's range outside both conic ranges
            r -= len1 + len2;
            const int rd0 = o_rd & 0xffff, rd1 = (int)(o_rd >> 16) - 1;
            x = rd0;
            for (; x <= rd1; ++x)
              if (!((x >= ra1 && x <= rb1) || (x >= ra2 && x <= rb2))) { if (r == 0) break; --r; }
          }
          const uint2 ie = s_ie[wb + o_g];
          const uint32_t slot = o_o + (j - (o_tp - o_n));
          if (slot < ie.y) put(slot, o_base + (uint32_t)x, ie.x);
        }
      }
    }
  }
  __syncwarp();
  if (!EMIT) {
    if (i < a.P) a.kept[i] = s_run[t];
  } else {
    // Count and emission are one template on the same inputs, so s_run[t] == end here.  Should two
    // compilations ever disagree, no slot may stay uninitialised: repeat the Gaussian on its first
    // tile (the `o < end` guards above cover the opposite case).
    if (vis) {
      const int4 bx = s_box[t];
      for (uint32_t o = s_run[t]; o < end; ++o) put(o, (uint32_t)(bx.z * a.gx + bx.x), id);
    }
  }
  }   // next chunk
  if (hist) {
    __syncthreads();
    const uint32_t h0 = s_h[0][t], h1 = s_h[1][t];
    if (h0) atomicAdd(a.eh.hist + t, h0);
    if (h1) atomicAdd(a.eh.hist + RS_RADIX + t, h1);
  }
}

int surfel_count_kept(int P, const uint32_t* tiles_touched, const uint2* rects, const CullRec* cull, uint32_t* kept,
                      const RowCache* rows, cudaStream_t s, bool debug) {
  if (P <= 0) return B200GS_OK;
  StageScope _t(ST_COUNT_KEPT, s);
  KeptArgs a{};
  a.P = P; a.tiles_touched = tiles_touched; a.rects = rects; a.cull = cull; a.kept = kept;
  if (rows) a.rows = *rows;
  a.small_rect = rows ? SMALL_RECT_SLICED : SMALL_RECT_UNSLICED;     // the surfel path hands in its row cache
  a.nchunks = (P + 255) / 256;
  kept_rows_kernel<false><<<a.nchunks, 256, 0, s>>>(a);
  B200GS_LAUNCHED("kept_rows_kernel<count>", debug, s);
  return B200GS_OK;
}

// identifyTileRanges (rasterizer_impl.cu:116-138) over the sorted keys [first, first + L) of one depth slice
// (the whole list when there is one slice); ranges hold absolute positions.
__global__ void __launch_bounds__(256) tile_ranges_kernel(int first, int L, const uint32_t* __restrict__ keys, uint2* ranges,
                                                          const uint32_t* __restrict__ live) {
  if (live != nullptr && *live == 0u) return;
  // grid-stride: the grid is capped (slice_grid) so that a slice nobody needs costs one small launch, not one CTA per
  // 256 instances of it
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < L; idx += gridDim.x * blockDim.x) {
    const uint32_t cur = keys[first + idx];
    if (idx == 0) {
      ranges[cur].x = first;
    } else {
      const uint32_t prev = keys[first + idx - 1];
      if (cur != prev) { ranges[prev].y = first + idx; ranges[cur].x = first + idx; }
    }
    if (idx == L - 1) ranges[cur].y = first + L;
  }
}

// Launch width of the per-slice kernels: enough CTAs to fill the GPU several times over, never one per work item.
static inline int slice_grid(long long work_items) {
  const long long cap = 148 * 16;
  return (int)(work_items < cap ? (work_items > 0 ? work_items : 1) : cap);
}

// ------------------------------------------------------------------ tile hits
// For every sorted (tile, Gaussian) instance: which of the tile's eight 8x4 warp blocks (bit w:
// column w&1, row w>>1, the layout of tile_pixel()) can the Gaussian reach alpha >= 1/255 in?
// Block = pixel centres [8c, 8c+7] x [4r, 4r+3] relative to the tile origin, tested as the
// rectangle dilated by half a pixel: (a) the exact minimum of the CullRec quadratic form f over
// the rectangle -- for ANY quadratic it is attained at a corner, at the vertex of f along an edge
// (if f is convex along it and the vertex lies on the edge) or at the interior critical point (if
// f is convex and the point lies inside) -- compared with a tolerance 100x the fp32 evaluation
// error; (b) the distance of the low-pass disk centre to the rectangle.
// Instances with an empty mask are dropped from the COMPACTED list the compositing kernels walk
// (single-pass stream compaction, chained-scan look-back); the reference's list itself is
// untouched and every survivor carries its position in it (that is what n_contrib counts).
constexpr int TH_THREADS = 256;
constexpr int TH_ITEMS = 4;
constexpr int TH_TILE = TH_THREADS * TH_ITEMS;

// The CullRec form translated to the corner (-0.5, -0.5) of the dilated tile: the tile is [0,16]^2
// and block (r, c) is [8c, 8c+8] x [4r, 4r+4] in this frame.
struct TileForm {
  float A, B, C, D, E, F, tol;   // f(x,y) = A x^2 + 2 B x y + C y^2 + 2 D x + 2 E y + F; hit if min f <= tol
  float dxc, dyc, r2;            // low-pass disk: centre in this frame, squared radius
};

// -> 0: fixed answer "nothing", 1: fixed answer "everything", 2: use the form
__device__ __forceinline__ int load_tile_form(const CullRec* __restrict__ cull, uint32_t id, int tx0, int ty0, TileForm& f) {
  const float4 c2 = __ldg(reinterpret_cast<const float4*>(&cull[id].cx));    // cx cy r2 flag
  if (c2.w != 0.0f) return (c2.w == CULL_NOTHING) ? 0 : 1;
  const double4 c0 = *reinterpret_cast<const double4*>(&cull[id].A);         // A B C D
  const double2 c1 = *reinterpret_cast<const double2*>(&cull[id].E);         // E F
  // translation in fp64: the pixel coordinates are large against the extent of small far-off ellipses
  const double ox = (double)tx0 - 0.5, oy = (double)ty0 - 0.5;
  const double Dd = c0.x * ox + c0.y * oy + c0.w;
  const double Ed = c0.y * ox + c0.z * oy + c1.x;
  const double Fd = (c0.x * ox + 2.0 * c0.y * oy + 2.0 * c0.w) * ox + (c0.z * oy + 2.0 * c1.x) * oy + c1.y;
  f.A = (float)c0.x; f.B = (float)c0.y; f.C = (float)c0.z; f.D = (float)Dd; f.E = (float)Ed; f.F = (float)Fd;
  if (!(fabsf(f.D) < 1e30f) || !(fabsf(f.E) < 1e30f) || !(fabsf(f.F) < 1e30f)) return 1;
  f.tol = 1e-5f * (256.0f * (fabsf(f.A) + 2.0f * fabsf(f.B) + fabsf(f.C)) + 32.0f * (fabsf(f.D) + fabsf(f.E)) + fabsf(f.F));
  f.dxc = c2.x - ((float)tx0 - 0.5f);
  f.dyc = c2.y - ((float)ty0 - 0.5f);
  f.r2 = c2.z;
  return 2;
}

// Can the Gaussian reach the tile at all?  Same test as block_masks() on the whole tile rectangle
// (a superset of the union of its blocks): most instances of a scene with near-camera splats fail
// it, so only the survivors pay for the per-block refinement.
__device__ __forceinline__ bool tile_reachable(const TileForm& f) {
  const float A = f.A, B = f.B, C = f.C, D = f.D, E = f.E, F = f.F;
  const float S = 16.0f;
  float mn = fminf(fminf(F, (A * S + 2.0f * D) * S + F),
                   fminf((C * S + 2.0f * E) * S + F, (A * S + 2.0f * (B * S + D)) * S + (C * S + 2.0f * E) * S + F));
  if (A > 0.0f) {   // horizontal edges y = 0, 16
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const float y = S * j, bx = B * y + D, gy = (C * y + 2.0f * E) * y + F;
      const float xs = -bx / A;
      if (xs > 0.0f && xs < S) mn = fminf(mn, gy - bx * bx / A);
    }
  }
  if (C > 0.0f) {   // vertical edges x = 0, 16
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const float x = S * i, by = B * x + E, gx = (A * x + 2.0f * D) * x + F;
      const float ys = -by / C;
      if (ys > 0.0f && ys < S) mn = fminf(mn, gx - by * by / C);
    }
  }
  const float det = A * C - B * B;
  if (A > 0.0f && det > 0.0f) {
    const float xs = (B * E - C * D) / det, ys = (B * D - A * E) / det;
    if (xs > 0.0f && xs < S && ys > 0.0f && ys < S) mn = fminf(mn, D * xs + E * ys + F);
  }
  const float ddx = fmaxf(fmaxf(-f.dxc, f.dxc - S), 0.0f), ddy = fmaxf(fmaxf(-f.dyc, f.dyc - S), 0.0f);
  return mn <= f.tol || ddx * ddx + ddy * ddy <= f.r2;
}

// Bits of the blocks in tile rows [r0, r1] (clamped to 0..3), both columns.
__host__ __device__ constexpr unsigned rows_bits(int r0, int r1) {
  unsigned m = 0;
  for (int r = (r0 < 0 ? 0 : r0); r <= (r1 > 3 ? 3 : r1); ++r) m |= 3u << (2 * r);
  return m;
}
// Bits of the blocks in columns [c0, c1] (clamped to 0..1), all rows.
__host__ __device__ constexpr unsigned cols_bits(int c0, int c1) {
  unsigned m = 0;
  for (int c = (c0 < 0 ? 0 : c0); c <= (c1 > 1 ? 1 : c1); ++c) m |= 0x55u << c;
  return m;
}

// A block is hit when ANY candidate point of its rectangle has f <= tol, so instead of tracking the
// eight minima every candidate ORs the bits of the blocks it belongs to.
__device__ __forceinline__ unsigned block_masks(const TileForm& f) {
  const float A = f.A, B = f.B, C = f.C, D = f.D, E = f.E, F = f.F, tol = f.tol;
  // one reciprocal per direction instead of a division per grid line (the vertex positions only
  // select a segment and the values are compared with a tolerance 100x the fp32 error)
  const float invA = A > 0.0f ? 1.0f / A : 0.0f, invC = C > 0.0f ? 1.0f / C : 0.0f;
  unsigned mask = 0u;
  // grid lines x in {0,8,16}, y in {0,4,..,16}: corners, and the vertex of f along each horizontal line
#pragma unroll
  for (int j = 0; j < 5; ++j) {
    const float y = 4.0f * j;
    const float gy = (C * y + 2.0f * E) * y + F;      // terms without x
    const float bx = B * y + D;                         // f = A x^2 + 2 bx x + gy
    constexpr unsigned NONE = 0u;
    const unsigned rb = rows_bits(j - 1, j);            // blocks touching this line
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const float x = 8.0f * i;
      const float v = (A * x + 2.0f * bx) * x + gy;
      mask |= (v <= tol) ? (rb & cols_bits(i - 1, i)) : NONE;
    }
    if (A > 0.0f) {
      const float xs = -bx * invA;
      const float val = gy - bx * bx * invA;
      if (val <= tol && xs > 0.0f && xs < 16.0f) mask |= rb & ((xs <= 8.0f ? 0x55u : 0u) | (xs >= 8.0f ? 0xAAu : 0u));
    }
  }
  // vertices along the three vertical lines
  if (C > 0.0f) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const float x = 8.0f * i;
      const float by = B * x + E;                        // f = C y^2 + 2 by y + gx
      const float gx = (A * x + 2.0f * D) * x + F;
      const float ys = -by * invC;
      const float val = gx - by * by * invC;
      if (val <= tol && ys > 0.0f && ys < 16.0f) {
        unsigned rsel = 0u;
#pragma unroll
        for (int r = 0; r < 4; ++r)
          if (ys >= 4.0f * r && ys <= 4.0f * (r + 1)) rsel |= 3u << (2 * r);
        mask |= rsel & cols_bits(i - 1, i);
      }
    }
  }
  // interior critical point
  const float det = A * C - B * B;
  if (A > 0.0f && det > 0.0f) {
    const float idet = 1.0f / det;
    const float xs = (B * E - C * D) * idet, ys = (B * D - A * E) * idet;
    if (xs > 0.0f && xs < 16.0f && ys > 0.0f && ys < 16.0f && D * xs + E * ys + F <= tol) {   // f at the critical point
      unsigned rsel = 0u;
#pragma unroll
      for (int r = 0; r < 4; ++r)
        if (ys >= 4.0f * r && ys <= 4.0f * (r + 1)) rsel |= 3u << (2 * r);
      mask |= rsel & ((xs <= 8.0f ? 0x55u : 0u) | (xs >= 8.0f ? 0xAAu : 0u));
    }
  }
  // low-pass disk: distance of its centre to each block rectangle, separable in x and y
  {
    const float dx0 = fmaxf(fmaxf(0.0f - f.dxc, f.dxc - 8.0f), 0.0f), dx1 = fmaxf(fmaxf(8.0f - f.dxc, f.dxc - 16.0f), 0.0f);
    const float ex0 = dx0 * dx0, ex1 = dx1 * dx1;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const float ddy = fmaxf(fmaxf(4.0f * r - f.dyc, f.dyc - 4.0f * (r + 1)), 0.0f);
      const float ey = ddy * ddy;
      mask |= (ex0 + ey <= f.r2) ? (1u << (2 * r)) : 0u;
      mask |= (ex1 + ey <= f.r2) ? (2u << (2 * r)) : 0u;
    }
  }
  return mask;
}

// state[0] = ticket; look-back words (flag << 32 | count) start at state + 2.
// PREFILTER: whole-tile test first, per-block refinement only for the survivors (exact-lists mode,
// where most instances miss their tile); without it every instance goes straight to the per-block
// test (culled emission already dropped the instances that miss their tile).
// One depth slice = the sorted instances [first, first + R); its survivors go to comp_*[first ...] (absolute
// positions, like ranges / cranges).  Instances carrying the sentinel key `ntiles` (their tile was finished by an
// earlier slice) are dropped.
template <bool PREFILTER>
__global__ void __launch_bounds__(TH_THREADS) tile_hits_kernel(int first, int R, const uint32_t* __restrict__ keys,
                                                               const uint32_t* __restrict__ point_list,
                                                               const CullRec* __restrict__ cull, int gx, uint32_t ntiles,
                                                               const uint2* __restrict__ ranges, uint2* __restrict__ cranges,
                                                               uint32_t* __restrict__ comp_id, uint32_t* __restrict__ comp_pos,
                                                               uint8_t* __restrict__ comp_hit, uint32_t* state,
                                                               int keep_all, const uint32_t* __restrict__ live) {
  if (live != nullptr && *live == 0u) return;
  keys += first; point_list += first;
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_cnt[TH_ITEMS][TH_THREADS / 32];
  __shared__ uint32_t s_base;
  __shared__ uint8_t s_mask[TH_TILE];    // per item of the block: hit mask
  __shared__ uint16_t s_queue[TH_TILE];  // items that reach their tile and need the per-block refinement
  __shared__ uint32_t s_qn;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const uint32_t nblocks = (uint32_t)((R + TH_TILE - 1) / TH_TILE);
  // persistent CTAs: blocks of TH_TILE items are taken in ticket order until none is left (the look-back below only
  // ever waits for tickets that were handed out before, i.e. for CTAs that are running)
  while (true) {
  __syncthreads();                                   // the previous block's shared state is no longer read
  if (t == 0) { s_tile = atomicAdd(state, 1u); s_qn = 0u; }
  __syncthreads();
  const uint32_t blk = s_tile;
  if (blk >= nblocks) return;
  const int base = (int)blk * TH_TILE;

  unsigned mask[TH_ITEMS];
  uint32_t id[TH_ITEMS], tile[TH_ITEMS], wexcl[TH_ITEMS];
  bool keep[TH_ITEMS];
  // phase 1: whole-tile test for every item; survivors are queued
#pragma unroll
  for (int k = 0; k < TH_ITEMS; ++k) {
    const int i = base + k * TH_THREADS + t;
    id[k] = 0u; tile[k] = 0u;
    unsigned m = 0u;
    bool refine = false;
    if (i < R && keys[i] < ntiles) {
      tile[k] = keys[i];
      id[k] = point_list[i];
      TileForm f;
      const int kind = load_tile_form(cull, id[k], (int)(tile[k] % (uint32_t)gx) * TILE_X, (int)(tile[k] / (uint32_t)gx) * TILE_Y, f);
      if (kind == 1) m = 0xFFu;
      else if (kind == 2) {
        if (PREFILTER) refine = tile_reachable(f);
        else m = block_masks(f);
      }
    }
    s_mask[k * TH_THREADS + t] = (uint8_t)m;
    const unsigned bal = PREFILTER ? __ballot_sync(0xffffffffu, refine) : 0u;
    if (bal) {
      uint32_t qb = 0u;
      if (lane == 0) qb = atomicAdd(&s_qn, (uint32_t)__popc(bal));
      qb = __shfl_sync(0xffffffffu, qb, 0);
      if (refine) s_queue[qb + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)(k * TH_THREADS + t);
    }
  }
  __syncthreads();
  // phase 2: per-block refinement of the queued items, densely packed over the threads
  for (uint32_t q = t; q < s_qn; q += TH_THREADS) {
    const int li = s_queue[q];
    const int i = base + li;
    const uint32_t tl = keys[i];
    TileForm f;
    load_tile_form(cull, point_list[i], (int)(tl % (uint32_t)gx) * TILE_X, (int)(tl / (uint32_t)gx) * TILE_Y, f);
    s_mask[li] = (uint8_t)block_masks(f);
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < TH_ITEMS; ++k) {
    const int i = base + k * TH_THREADS + t;
    mask[k] = s_mask[k * TH_THREADS + t];
    if (i < R) tile[k] = keys[i];
    keep[k] = i < R && tile[k] < ntiles && (mask[k] != 0u || keep_all);   // keep_all: cull audit walks the full lists
    const unsigned bal = __ballot_sync(0xffffffffu, keep[k]);
    wexcl[k] = __popc(bal & ((1u << lane) - 1u));
    if (lane == 0) s_cnt[k][warp] = __popc(bal);
  }
  __syncthreads();
  // exclusive offsets of every (round, warp) group and the block total
  uint32_t total = 0, mine[TH_ITEMS] = {0u, 0u, 0u, 0u};
#pragma unroll
  for (int k = 0; k < TH_ITEMS; ++k)
#pragma unroll
    for (int w = 0; w < TH_THREADS / 32; ++w) {
      if (w == warp) mine[k] = total;
      total += s_cnt[k][w];
    }
  if (warp == 0) {
    volatile unsigned long long* st = reinterpret_cast<volatile unsigned long long*>(state + 2);
    if (lane == 0) st[blk] = ((blk == 0 ? 2ull : 1ull) << 32) | total;
    uint32_t excl = 0;
    if (blk > 0) {
      int look = (int)blk - 1;
      while (true) {
        const int q = look - lane;
        unsigned long long w = 2ull << 32;
        if (q >= 0) { do { w = st[q]; } while ((w >> 32) == 0ull); }
        const bool full = (w >> 32) == 2ull;
        const unsigned ballot = __ballot_sync(0xffffffffu, full);
        const int first = __ffs(ballot) - 1;
        uint32_t val = (first < 0 || lane <= first) ? (uint32_t)w : 0u;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
        excl += val;
        if (first >= 0) break;
        look -= 32;
      }
      if (lane == 0) st[blk] = (2ull << 32) | (unsigned long long)(excl + total);
    }
    if (lane == 0) s_base = excl;
  }
  __syncthreads();
  const uint32_t gbase = s_base;
#pragma unroll
  for (int k = 0; k < TH_ITEMS; ++k) {
    const int i = base + k * TH_THREADS + t;
    if (i >= R || tile[k] >= ntiles) continue;
    const uint32_t off = (uint32_t)first + gbase + mine[k] + wexcl[k];
    const uint2 rg = ranges[tile[k]];
    const uint32_t ia = (uint32_t)(first + i);        // absolute position in the sorted list
    if (keep[k]) {
      comp_id[off] = id[k];
      comp_pos[off] = ia - rg.x;
      comp_hit[off] = (uint8_t)mask[k];
    }
    if (ia == rg.x) cranges[tile[k]].x = off;
    if (ia + 1u == rg.y) cranges[tile[k]].y = off + (keep[k] ? 1u : 0u);
  }
  }   // next ticket
}

// ------------------------------------------------------------------ slice plan
// Cuts the depth-ordered Gaussians into slices at given fractions of the cumulated instance count.  One warp per
// boundary: 32-ary search for the first Gaussian whose inclusive offset reaches the target.  Thread 0 then
// assembles the plan (common.cuh: PLAN_G / PLAN_OFF), dropping empty slices.
struct SlicePlanArgs {
  const uint32_t* offsets;
  int P, n_bounds;
  float frac[MAX_SLICES - 1];
  uint32_t min_instances;        // fewer instances than this: one slice
  int* plan;
};

__global__ void __launch_bounds__(32 * (MAX_SLICES - 1)) slice_plan_kernel(const SlicePlanArgs a) {
  __shared__ int s_g[MAX_SLICES - 1];
  __shared__ uint32_t s_off[MAX_SLICES - 1];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t R = a.offsets[a.P - 1];
  if (w < a.n_bounds) {
    const uint32_t target = (uint32_t)fminf((float)R * a.frac[w], (float)R);
    // first i with offsets[i] >= target; invariant: the answer lies in [lo, hi] (hi == P: none)
    int lo = 0, hi = a.P;
    while (hi - lo > 32) {
      const int step = (hi - lo) / 32;
      const int idx = lo + (lane + 1) * step - 1;            // <= hi - 1
      const unsigned m = __ballot_sync(0xffffffffu, a.offsets[idx] >= target);
      if (m == 0u) { lo += 32 * step; continue; }
      const int f = __ffs(m) - 1;
      hi = lo + (f + 1) * step - 1;
      lo = lo + f * step;                                     // probe f-1 (index lo + f*step - 1) was still below
    }
    {
      const int idx = lo + lane;
      const unsigned m = __ballot_sync(0xffffffffu, idx < hi && a.offsets[idx] >= target);
      lo = m ? lo + __ffs(m) - 1 : hi;
    }
    if (lane == 0) { s_g[w] = min(lo + 1, a.P); s_off[w] = lo < a.P ? a.offsets[lo] : R; }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int S = 0;
    a.plan[0] = (int)R;
    a.plan[PLAN_G] = 0; a.plan[PLAN_OFF] = 0;
    if (R >= a.min_instances) {
      for (int k = 0; k < a.n_bounds; ++k) {
        const int g = s_g[k]; const uint32_t off = s_off[k];
        if (g > a.plan[PLAN_G + S] && g < a.P && off > (uint32_t)a.plan[PLAN_OFF + S] && off < R) {
          ++S;
          a.plan[PLAN_G + S] = g; a.plan[PLAN_OFF + S] = (int)off;
        }
      }
    }
    ++S;
    a.plan[PLAN_G + S] = a.P; a.plan[PLAN_OFF + S] = (int)R;
    a.plan[1] = S;
  }
}

int slice_plan(const uint32_t* point_offsets, int P, int max_slices, const float* frac, int* plan_dev, cudaStream_t s, bool debug) {
  SlicePlanArgs a{};
  a.offsets = point_offsets; a.P = P; a.plan = plan_dev;
  a.min_instances = 1u << 21;
  // measured on the 1080p scenes (scripts/gpu_slices.py, profiles/r2d_slices.jsonl): the first 5 % of the instances finish
  // all but a few tiles, 25 % finish all; finer schedules pay more launches than they save
  static const float sched[MAX_SLICES - 1] = {0.05f, 0.25f, 0, 0, 0, 0, 0};
  int n = max_slices <= 0 ? 3 : (max_slices > MAX_SLICES ? MAX_SLICES : max_slices);
  const bool own = frac != nullptr && frac[0] > 0.0f;
  if (max_slices > 1 && own) a.min_instances = 0;         // an explicit schedule is taken literally
  a.n_bounds = 0;
  for (int k = 0; k < n - 1; ++k) {
    const float f = own ? frac[k] : sched[k];
    if (!(f > 0.0f && f < 1.0f)) break;
    a.frac[a.n_bounds++] = f;
  }
  slice_plan_kernel<<<1, 32 * (MAX_SLICES - 1), 0, s>>>(a);
  B200GS_LAUNCHED("slice_plan_kernel", debug, s);
  return B200GS_OK;
}

size_t tile_hits_state_words(size_t R) { return 2 * ((R + TH_TILE - 1) / TH_TILE + 2); }

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

static inline size_t al256(size_t b) { return (b + 255) / 256 * 256; }
constexpr int TILE_SORT_MAX_PASSES = 3;   // tile ids stay below 2^24

Binning Binning::carve(void* base, size_t R, size_t* bytes) {
  Carver c(base);
  Binning b;
  b.point_list = c.take<uint32_t>(R);
  b.tile_ids = c.take<uint32_t>(R);
  b.tmp_vals = c.take<uint32_t>(R);
  b.tmp_keys = c.take<uint32_t>(R);
  b.comp_id = c.take<uint32_t>(R);
  b.comp_pos = c.take<uint32_t>(R);
  b.comp_hit = c.take<uint8_t>(R);
  // Every depth slice sorts and compacts its own range with its own look-back words; all of them live in one region
  // that is cleared once per forward.  Bound over any split of R into <= MAX_SLICES ranges.
  b.zero_bytes = al256(radix_workspace_bytes(R, TILE_SORT_MAX_PASSES)) + al256(tile_hits_state_words(R) * sizeof(uint32_t)) +
                 (size_t)MAX_SLICES * (al256(radix_workspace_bytes(RS_TILE, TILE_SORT_MAX_PASSES)) +
                                       al256(tile_hits_state_words(TH_TILE) * sizeof(uint32_t)));
  b.zero = c.take<char>(b.zero_bytes);
  if (bytes) *bytes = c.used + 128;
  return b;
}

int bin_clear(const Binning& b, const BinJob& j, cudaStream_t s) {
  const size_t tiles1 = (size_t)j.tiles() + 1;
  B200GS_CUDA(cudaMemsetAsync(j.ranges, 0, (size_t)j.slices * tiles1 * sizeof(uint2), s));
  B200GS_CUDA(cudaMemsetAsync(j.cranges, 0, (size_t)j.slices * tiles1 * sizeof(uint2), s));
  if (j.R > 0) B200GS_CUDA(cudaMemsetAsync(b.zero, 0, b.zero_bytes, s));
  return B200GS_OK;
}

// One depth slice k of the instance list: emission of the Gaussians [g[k], g[k+1]) of the depth order into the slots
// [off[k], off[k+1]), stable sort of that range by tile id, its tile ranges, its block masks + compaction.
// `zcursor`: running offset into b.zero (the slice's sort workspace and compaction words are carved from it).
int bin_slice(const BinJob& j, int k, const Binning& b, size_t& zcursor, cudaStream_t s) {
  const int gx = (j.W + TILE_X - 1) / TILE_X;
  const int tiles = j.tiles();
  const int first = j.off[k], n = j.off[k + 1] - j.off[k];
  if (n <= 0) return B200GS_OK;
  const bool sliced = j.slices > 1;
  const uint32_t* live = sliced && k > 0 ? j.live : nullptr;     // nothing is finished before the first slice ran
  uint2* ranges = j.ranges + (size_t)k * (tiles + 1);
  uint2* cranges = j.cranges + (size_t)k * (tiles + 1);
  // same bit count as the reference's sort range above the depth word (rasterizer_impl.cu:306); one more value
  // (the sentinel key `tiles`) when later slices skip finished tiles
  const RadixPlan plan = radix_plan((int)higher_msb((uint32_t)(sliced ? tiles + 1 : tiles)));
  if (plan.passes > TILE_SORT_MAX_PASSES) { set_error("too many tiles (%d)", tiles); return B200GS_EINVAL; }
  void* ws = b.zero + zcursor;
  zcursor += al256(radix_workspace_bytes((size_t)n, plan.passes));
  uint32_t* hit_state = reinterpret_cast<uint32_t*>(b.zero + zcursor);
  zcursor += al256(tile_hits_state_words((size_t)n) * sizeof(uint32_t));
  if (zcursor > b.zero_bytes) { set_error("binning scratch: slice workspaces exceed the carved region"); return B200GS_ESIZE; }
  // the sorted pairs must end up in (tile_ids, point_list): start there for an even pass count
  const bool odd = (plan.passes & 1) != 0;
  uint32_t* ka = (odd ? b.tmp_keys : b.tile_ids);
  uint32_t* va = (odd ? b.tmp_vals : b.point_list);
  uint32_t* kb = (odd ? b.tile_ids : b.tmp_keys);
  uint32_t* vb = (odd ? b.point_list : b.tmp_vals);
  int rc;
  const bool fused_hist = j.cull != nullptr && plan.passes <= 2;   // kept_rows_kernel<emit> accumulates the digit histograms
  {
    StageScope _t(ST_DUPLICATE, s);
    if (j.cull != nullptr) {
      EmitHist eh;
      eh.hist = fused_hist ? radix_hist_ptr(ws) : nullptr;
      eh.mask0 = (1u << plan.bits[0]) - 1u;
      eh.shift1 = plan.passes > 1 ? plan.shift[1] : 0;
      eh.mask1 = plan.passes > 1 ? (1u << plan.bits[1]) - 1u : 0u;
      KeptArgs ea{};
      ea.P = j.P; ea.gx = gx; ea.order = j.order; ea.offsets = j.offsets; ea.rects = j.rects; ea.cull = j.cull;
      ea.g0 = j.g[k]; ea.g1 = j.g[k + 1];
      ea.done = sliced && k > 0 ? j.tile_done : nullptr; ea.sentinel = (uint32_t)tiles; ea.live = live;
      ea.keys = ka; ea.vals = va; ea.eh = eh;     // slots are absolute positions (offsets is the global prefix sum)
      if (j.rows) ea.rows = *j.rows;
      ea.small_rect = j.rows ? SMALL_RECT_SLICED : SMALL_RECT_UNSLICED;
      const int ng = ea.g1 - ea.g0;
      if (ng > 0) {
        const long long per = (long long)n / ng;                 // instances per Gaussian of this slice
        // spread only where Gaussians are few: with hundreds of thousands of them every SM is busy anyway and the
        // idle helper lanes would just be more warps to schedule
        ea.lpg_shift = (per >= 64 && ng < (1 << 17)) ? 5 : ((per >= 24 && ng < (1 << 19)) ? 3 : 0);
        const long long threads = (long long)ng << ea.lpg_shift;
        ea.nchunks = (int)((threads + 255) / 256);
        kept_rows_kernel<true><<<slice_grid(ea.nchunks), 256, 0, s>>>(ea);
        B200GS_LAUNCHED("kept_rows_kernel<emit>", j.debug, s);
      }
    } else {
      duplicate_kernel<<<(j.P + 255) / 256, 256, 0, s>>>(j.P, j.order, j.rects, j.offsets, ka, va, gx);
      B200GS_LAUNCHED("duplicate_kernel", j.debug, s);
    }
  }
  {
    StageScope _t(ST_SORT, s);
    if (!fused_hist) {
      rc = radix_histogram(ka + first, n, plan, ws, s, j.debug);
      if (rc) return rc;
    }
    rc = radix_sort_pairs_u32(ka + first, va + first, kb + first, vb + first, n, plan, ws, s, j.debug, live);
    if (rc) return rc;
  }
  {
    StageScope _t(ST_RANGES, s);
    tile_ranges_kernel<<<slice_grid((n + 255) / 256), 256, 0, s>>>(first, n, b.tile_ids, ranges, live);
    B200GS_LAUNCHED("tile_ranges_kernel", j.debug, s);
  }
  StageScope _t(ST_TILE_HITS, s);
  const int keep_all = j.audit ? 1 : 0;   // the cull audit walks the full lists
  const int blocks = (n + TH_TILE - 1) / TH_TILE;
  if (j.cull_for_hits == nullptr) { set_error("bin_slice: no cull records"); return B200GS_EINVAL; }
  if (j.cull == nullptr)     // exact lists: most instances miss their tile, test the whole tile first
    tile_hits_kernel<true><<<slice_grid(blocks), TH_THREADS, 0, s>>>(first, n, b.tile_ids, b.point_list, j.cull_for_hits, gx, (uint32_t)tiles, ranges,
                                                         cranges, b.comp_id, b.comp_pos, b.comp_hit, hit_state, keep_all, live);
  else
    tile_hits_kernel<false><<<slice_grid(blocks), TH_THREADS, 0, s>>>(first, n, b.tile_ids, b.point_list, j.cull_for_hits, gx, (uint32_t)tiles, ranges,
                                                          cranges, b.comp_id, b.comp_pos, b.comp_hit, hit_state, keep_all, live);
  B200GS_LAUNCHED("tile_hits_kernel", j.debug, s);
  return B200GS_OK;
}

}  // namespace b200gs
