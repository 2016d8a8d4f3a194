// radix_sort.cuh -- hand-written stable LSD radix sort of (u32 key, u32 value)
// pairs for sm_100a: one "onesweep" pass per digit (single read of the pairs,
// chained-scan look-back between blocks, shared-memory reorder so every digit
// leaves the block as one contiguous run).
//
// Replaces cub::DeviceRadixSort::SortPairs at S/cuda_rasterizer/rasterizer_impl.cu:308-313
// (and the same call in G/).  The reference sorts 64-bit (tile << 32 | depth bits) keys
// over 32 + ceil(log2(tiles)) bits in one go (6 passes over 12-B pairs at 1080p).  A
// stable LSD sort is the composition of stable sorts on its digit groups, so the same
// ORDER is produced by
//   1. a stable sort of the P Gaussians by the 32 depth bits   (4 passes over P pairs), then
//   2. emission of the (tile, id) instances in that order, and
//   3. a stable sort of the R instances by tile id alone       (2 passes up to 65 536 tiles),
// which moves 8-B pairs twice instead of 12-B pairs six times.  binning.cu does that.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

namespace b200gs {

constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // pairs per block
constexpr int RS_RADIX = 256;                   // bins per pass (digits of <= 8 bits)
constexpr int RS_MAX_PASSES = 4;

// How a key range of `total_bits` bits is cut into passes of <= 8 bits.
struct RadixPlan {
  int passes;
  int shift[RS_MAX_PASSES];
  int bits[RS_MAX_PASSES];
};
RadixPlan radix_plan(int total_bits);

// Workspace: [RS_MAX_PASSES][256] digit histograms, RS_MAX_PASSES block tickets, and one 64-bit look-back word
// per (pass, block, digit) -- every pass has its own words, so ONE clear before the first pass suffices (a caller
// that sorts several independent ranges carves one workspace per range out of a single zeroed region).
size_t radix_workspace_bytes(size_t n, int passes);

// Zeroes a workspace of radix_workspace_bytes(n, passes).  Must precede radix_histogram / radix_sort_pairs_u32.
int radix_workspace_clear(void* workspace, size_t n, int passes, cudaStream_t s);

// The [RS_MAX_PASSES][256] histogram block of a workspace, for producers that count digits themselves.
uint32_t* radix_hist_ptr(void* workspace);

// Accumulates the digit histograms of all passes of `plan` over keys[0, n).
int radix_histogram(const uint32_t* keys, int n, const RadixPlan& plan, void* workspace, cudaStream_t s, bool debug);

// Sorts n pairs stably on key bits [0, total_bits).  (keys_a, vals_a) hold the input and are
// overwritten; the result is in (keys_a, vals_a) when plan.passes is even and in (keys_b, vals_b)
// when it is odd.  The histograms must already be in the workspace (radix_histogram).
// `live` (optional, device): when *live == 0 at launch every block returns at once (depth-sliced binning: all
// tiles were finished by an earlier slice, nothing downstream reads this slice).
int radix_sort_pairs_u32(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b, int n,
                         const RadixPlan& plan, void* workspace, cudaStream_t s, bool debug,
                         const uint32_t* live = nullptr);

}  // namespace b200gs
