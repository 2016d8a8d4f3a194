// radix_sort.cu -- hand-written onesweep LSD radix sort of (u32 key, u32 value) pairs (sm_100a).
// See radix_sort.cuh for what it replaces and why 32-bit keys suffice.
//
// One pass = one kernel, one read and one write of every pair:
//   * blocks take their tile of RS_TILE pairs in ticket order (so a block only ever waits for
//     blocks that already run);
//   * every warp ranks its 512 keys with match.any: lanes holding the same digit find each other,
//     the lowest one bumps the warp's private digit counter in shared memory, and every lane's
//     rank is the old counter value plus the number of equal-digit lanes below it -- stable in
//     (iteration, lane) order, which is memory order;
//   * thread d owns digit d: it turns the global digit histogram into the digit's bucket base (a block-wide
//     exclusive scan every block repeats for itself: cheaper than a separate launch), the per-warp counters
//     into warp offsets, publishes the
//     block's count of digit d as a 64-bit (flag | count) word and walks back over the preceding
//     blocks' words (aggregate -> keep walking, inclusive prefix -> stop) -- the chained scan that
//     makes a single pass possible;
//   * pairs are reordered through shared memory so that each digit leaves the block as one
//     contiguous, coalesced run.
// The digit histograms that give every digit's global base come from radix_histogram (all passes
// in one read of the keys, match.any-aggregated shared-memory atomics).
#include "common.cuh"
#include "radix_sort.cuh"

namespace b200gs {

RadixPlan radix_plan(int total_bits) {
  RadixPlan p;
  if (total_bits < 1) total_bits = 1;
  if (total_bits > 32) total_bits = 32;
  p.passes = (total_bits + 7) / 8;
  int shift = 0;
  for (int i = 0; i < RS_MAX_PASSES; ++i) { p.shift[i] = 0; p.bits[i] = 0; }
  for (int i = 0; i < p.passes; ++i) {
    const int left = total_bits - shift, todo = p.passes - i;
    p.bits[i] = (left + todo - 1) / todo;
    p.shift[i] = shift;
    shift += p.bits[i];
  }
  return p;
}

// workspace layout (u32 words): hist[RS_MAX_PASSES][256] | tickets[RS_MAX_PASSES] | pad | u64 state[passes][blocks][256]
static constexpr size_t WS_HIST_WORDS = (size_t)RS_MAX_PASSES * RS_RADIX;
static constexpr size_t WS_HEAD_WORDS = WS_HIST_WORDS + 32;
static inline size_t ws_blocks(size_t n) { return (n + RS_TILE - 1) / RS_TILE; }

size_t radix_workspace_bytes(size_t n, int passes) {
  return WS_HEAD_WORDS * sizeof(uint32_t) + (size_t)passes * ws_blocks(n) * RS_RADIX * sizeof(unsigned long long) + 256;
}
static inline uint32_t* ws_hist(void* ws) { return reinterpret_cast<uint32_t*>(ws); }
static inline uint32_t* ws_tickets(void* ws) { return reinterpret_cast<uint32_t*>(ws) + WS_HIST_WORDS; }
static inline unsigned long long* ws_state(void* ws, size_t n, int pass) {
  return reinterpret_cast<unsigned long long*>(reinterpret_cast<uint32_t*>(ws) + WS_HEAD_WORDS) + (size_t)pass * ws_blocks(n) * RS_RADIX;
}

uint32_t* radix_hist_ptr(void* workspace) { return ws_hist(workspace); }

int radix_workspace_clear(void* workspace, size_t n, int passes, cudaStream_t s) {
  B200GS_CUDA(cudaMemsetAsync(workspace, 0, radix_workspace_bytes(n, passes), s));
  return B200GS_OK;
}

// ------------------------------------------------------------------ histogram
struct HistArgs {
  const uint32_t* keys;
  int n, passes;
  int shift[RS_MAX_PASSES];
  uint32_t mask[RS_MAX_PASSES];
  uint32_t* hist;  // [RS_MAX_PASSES][256]
};

__global__ void __launch_bounds__(256) radix_hist_kernel(const HistArgs a) {
  __shared__ uint32_t s_h[RS_MAX_PASSES][RS_RADIX];
  const int t = threadIdx.x, lane = t & 31;
#pragma unroll
  for (int p = 0; p < RS_MAX_PASSES; ++p) s_h[p][t] = 0u;
  __syncthreads();
  // 16-B loads, whole warps iterate together (match.any needs the full warp)
  const int nv = a.n >> 2;                       // number of complete uint4
  const int nv32 = (nv + 31) & ~31;
  const uint4* k4 = reinterpret_cast<const uint4*>(a.keys);
  for (int v = blockIdx.x * blockDim.x + t; v < nv32; v += gridDim.x * blockDim.x) {
    const bool valid = v < nv;
    const uint4 q = valid ? __ldg(k4 + v) : make_uint4(0u, 0u, 0u, 0u);
    const uint32_t key[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int c = 0; c < 4; ++c)
      for (int p = 0; p < a.passes; ++p) {
        const uint32_t d = valid ? ((key[c] >> a.shift[p]) & a.mask[p]) : (uint32_t)RS_RADIX;
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        if (valid && lane == __ffs(peers) - 1) atomicAdd(&s_h[p][d], (uint32_t)__popc(peers));
      }
  }
  if (blockIdx.x == 0 && t < 32) {               // the last n % 4 keys
    const int i = (nv << 2) + t;
    const bool valid = i < a.n;
    const uint32_t key = valid ? a.keys[i] : 0u;
    for (int p = 0; p < a.passes; ++p) {
      const uint32_t d = valid ? ((key >> a.shift[p]) & a.mask[p]) : (uint32_t)RS_RADIX;
      const unsigned peers = __match_any_sync(0xffffffffu, d);
      if (valid && lane == __ffs(peers) - 1) atomicAdd(&s_h[p][d], (uint32_t)__popc(peers));
    }
  }
  __syncthreads();
  for (int p = 0; p < a.passes; ++p) {
    const uint32_t v = s_h[p][t];
    if (v) atomicAdd(&a.hist[p * RS_RADIX + t], v);
  }
}

int radix_histogram(const uint32_t* keys, int n, const RadixPlan& plan, void* workspace, cudaStream_t s, bool debug) {
  if (n <= 0) return B200GS_OK;
  HistArgs a;
  a.keys = keys; a.n = n; a.passes = plan.passes; a.hist = ws_hist(workspace);
  for (int p = 0; p < RS_MAX_PASSES; ++p) { a.shift[p] = plan.shift[p]; a.mask[p] = (1u << plan.bits[p]) - 1u; }
  const int blocks = (int)((((size_t)n + 255) / 256) < 148 * 8 ? (((size_t)n + 255) / 256) : 148 * 8);
  radix_hist_kernel<<<blocks, 256, 0, s>>>(a);
  B200GS_LAUNCHED("radix_hist_kernel", debug, s);
  return B200GS_OK;
}

// ------------------------------------------------------------------ one pass
struct PassArgs {
  const uint32_t* keys_in;
  const uint32_t* vals_in;
  uint32_t* keys_out;
  uint32_t* vals_out;
  int n, shift, nbits;
  const uint32_t* hist;         // [256] digit counts of this pass over all n keys
  unsigned long long* state;    // [blocks][1 << nbits], zeroed: flag << 62 | count  (1 = aggregate, 2 = inclusive)
  uint32_t* ticket;
  const uint32_t* live;         // optional: *live == 0 -> nothing to do
};

// MODE 0: match.any per key inside the counter loop, values fetched during the reorder.
// MODE 1: the RS_ITEMS match.any of a thread issued back to back (independent), values fetched with the keys.
// MODE 2: as 1, with the peer masks built from one ballot per digit bit instead of match.any.
template <int MODE>
__device__ __forceinline__ unsigned digit_peers(uint32_t d, bool valid, int nbits) {
  if (MODE != 2) return __match_any_sync(0xffffffffu, valid ? d : (uint32_t)RS_RADIX);
  unsigned peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
  for (int b = 0; b < 8; ++b) {
    if (b < nbits) {
      const bool bit = (d >> b) & 1u;
      const unsigned vote = __ballot_sync(0xffffffffu, bit);
      peers &= bit ? vote : ~vote;
    }
  }
  return peers;   // meaningful on valid lanes only
}

template <int MODE, int THREADS, int ITEMS, int LB>
__global__ void __launch_bounds__(THREADS, 3) radix_pass_kernel(const PassArgs a) {
  static_assert(THREADS * ITEMS == RS_TILE && THREADS >= RS_RADIX, "tile shape");
  constexpr int WARPS = THREADS / 32;
  __shared__ uint32_t s_keys[RS_TILE];
  __shared__ uint32_t s_vals[RS_TILE];
  using cnt_t = uint32_t;   // per-warp digit counters, then offsets inside the block's digit runs
  __shared__ cnt_t s_cnt[WARPS][RS_RADIX];
  __shared__ uint32_t s_excl[RS_RADIX];    // exclusive prefix of the block's digit counts
  __shared__ uint32_t s_gbase[RS_RADIX];   // global position of local position 0 of each digit's run (mod 2^32)
  __shared__ uint32_t s_wsum[RS_RADIX / 32];
  __shared__ uint32_t s_blk;

  __shared__ uint32_t s_hsum[RS_RADIX / 32];

  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  if (a.live != nullptr && *a.live == 0u) return;
  const uint32_t dmask = (1u << a.nbits) - 1u;
  // exclusive bucket base of digit t over all n keys (the histogram is small and L2-resident)
  uint32_t dbase = 0u;
  {
    const uint32_t hv = t < RS_RADIX ? a.hist[t] : 0u;
    uint32_t inc = hv;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31 && t < RS_RADIX) s_hsum[warp] = inc;
    dbase = inc - hv;
  }
  __syncthreads();
  if (t < RS_RADIX) {
#pragma unroll
    for (int w = 0; w < RS_RADIX / 32; ++w) if (w < warp) dbase += s_hsum[w];
  }
  // Persistent CTAs: tiles of RS_TILE pairs are taken in ticket order until none is left.  The look-back only waits
  // for tickets handed out earlier, i.e. for CTAs that are running, so a capped grid cannot deadlock -- and a pass
  // nobody needs (`live`) is one small launch instead of one CTA per tile.
  const uint32_t nblocks = (uint32_t)((a.n + RS_TILE - 1) / RS_TILE);
  while (true) {
  __syncthreads();                                   // the previous tile's shared state is no longer read
  if (t == 0) s_blk = atomicAdd(a.ticket, 1u);
  for (int e = t; e < WARPS * RS_RADIX; e += THREADS) (&s_cnt[0][0])[e] = (cnt_t)0;
  __syncthreads();
  const uint32_t blk = s_blk;
  if (blk >= nblocks) return;
  const int blk_base = (int)blk * RS_TILE;
  const int wbase = blk_base + warp * (32 * ITEMS);

  uint32_t key[ITEMS], rank[ITEMS], val[MODE ? ITEMS : 1];
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int i = wbase + k * 32 + lane;
    key[k] = i < a.n ? a.keys_in[i] : 0xffffffffu;
  }
  if (MODE) {
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      const int i = wbase + k * 32 + lane;
      val[k] = i < a.n ? a.vals_in[i] : 0u;
    }
  }
  cnt_t* cnt = s_cnt[warp];
  if (MODE) {
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      const bool valid = wbase + k * 32 + lane < a.n;
      rank[k] = digit_peers<MODE>((key[k] >> a.shift) & dmask, valid, a.nbits);   // peer mask for now
    }
  }
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const bool valid = wbase + k * 32 + lane < a.n;
    const uint32_t d = (key[k] >> a.shift) & dmask;
    const unsigned peers = MODE ? rank[k] : digit_peers<0>(d, valid, a.nbits);
    const int leader = __ffs(peers) - 1;
    uint32_t old = 0u;
    if (lane == leader && valid) { old = cnt[d]; cnt[d] = (cnt_t)(old + (uint32_t)__popc(peers)); }
    old = __shfl_sync(0xffffffffu, old, valid ? leader : lane);
    rank[k] = old + (uint32_t)__popc(peers & ((1u << lane) - 1u));
    __syncwarp();
  }
  __syncthreads();

  // thread t < 256 owns digit t
  const int nb = 1 << a.nbits;
  volatile unsigned long long* st = a.state;
  uint32_t total = 0u, excl = 0u;
  if (t < RS_RADIX) {
#pragma unroll
    for (int w = 0; w < WARPS; ++w) {
      const uint32_t c = s_cnt[w][t];
      s_cnt[w][t] = (cnt_t)total;   // offset of warp w's keys inside the digit's run
      total += c;
    }
    if (t < nb) st[(size_t)blk * nb + t] = ((blk == 0 ? 2ull : 1ull) << 62) | total;
    uint32_t inc = total;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) s_wsum[warp] = inc;
    excl = inc - total;
  }
  __syncthreads();
  if (t < RS_RADIX) {
    uint32_t woff = 0u;
#pragma unroll
    for (int w = 0; w < RS_RADIX / 32; ++w) if (w < warp) woff += s_wsum[w];
    excl += woff;
    s_excl[t] = excl;
  }

  if (t < nb) {
    uint32_t prev = 0u;
    if (blk > 0) {
      // Walk back over the preceding blocks' words, LB of them per step: the loads of one step are
      // independent, so a walk over k aggregate-only predecessors costs k / LB memory round trips.
      long long look = (long long)blk - 1;
      bool done = false;
      while (!done) {
        unsigned long long w[LB];
#pragma unroll
        for (int j = 0; j < LB; ++j) w[j] = look - j >= 0 ? st[(size_t)(look - j) * nb + t] : (2ull << 62);
#pragma unroll
        for (int j = 0; j < LB; ++j) {
          if (!done) {
            while ((w[j] >> 62) == 0ull) w[j] = st[(size_t)(look - j) * nb + t];
            prev += (uint32_t)w[j];
            done = (w[j] >> 62) == 2ull;
          }
        }
        look -= LB;
      }
      st[(size_t)blk * nb + t] = (2ull << 62) | (unsigned long long)(prev + total);
    }
    s_gbase[t] = dbase + prev - excl;
  }
  __syncthreads();

  // reorder through shared memory
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    const int i = wbase + k * 32 + lane;
    if (i < a.n) {
      const uint32_t d = (key[k] >> a.shift) & dmask;
      const uint32_t pos = s_excl[d] + cnt[d] + rank[k];
      s_keys[pos] = key[k];
      s_vals[pos] = MODE ? val[k] : a.vals_in[i];
    }
  }
  __syncthreads();
  const int nvalid = min(RS_TILE, a.n - blk_base);
  for (int p = t; p < nvalid; p += THREADS) {
    const uint32_t k = s_keys[p];
    const uint32_t g = s_gbase[(k >> a.shift) & dmask] + (uint32_t)p;
    a.keys_out[g] = k;
    a.vals_out[g] = s_vals[p];
  }
  }   // next ticket
}

int radix_sort_pairs_u32(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b, int n,
                         const RadixPlan& plan, void* workspace, cudaStream_t s, bool debug, const uint32_t* live) {
  if (n <= 0) return B200GS_OK;
  const int blocks = (int)ws_blocks((size_t)n);
  uint32_t *ki = keys_a, *vi = vals_a, *ko = keys_b, *vo = vals_b;
  for (int p = 0; p < plan.passes; ++p) {
    PassArgs a;
    a.keys_in = ki; a.vals_in = vi; a.keys_out = ko; a.vals_out = vo;
    a.n = n; a.shift = plan.shift[p]; a.nbits = plan.bits[p];
    a.hist = ws_hist(workspace) + p * RS_RADIX;
    a.state = ws_state(workspace, (size_t)n, p);
    a.ticket = ws_tickets(workspace) + p;
    a.live = live;
    // Measured on the bench workload (round 1; tile sort of 33.8M pairs, two passes): ranking inside the counter loop
    // 0.76 ms, independent match.any 0.64, ballots per digit bit 0.64; look-back batch 1 0.66, 4 0.61, 8 0.65;
    // 512 threads x 8 keys 0.70-0.85.  Kept: independent match.any, 4-wide look-back.
    radix_pass_kernel<1, 256, 16, 4><<<blocks < 148 * 6 ? blocks : 148 * 6, 256, 0, s>>>(a);
    B200GS_LAUNCHED("radix_pass_kernel", debug, s);
    uint32_t* x;
    x = ki; ki = ko; ko = x;
    x = vi; vi = vo; vo = x;
  }
  return B200GS_OK;
}

}  // namespace b200gs
