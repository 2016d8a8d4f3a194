// common.cuh -- shared device helpers, scratch layouts and error plumbing of
// libb200gs (sm_100a).  Not part of the public ABI (that is include/b200gs.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <atomic>
#include "../../include/b200gs.h"

namespace b200gs {

// ---------------------------------------------------------------- constants
// Tile = 16x16 pixels (S/cuda_rasterizer/config.h:15-17).
constexpr int TILE_X = 16;
constexpr int TILE_Y = 16;
constexpr int TILE_PIX = TILE_X * TILE_Y;
// S/cuda_rasterizer/auxiliary.h:41-44
constexpr float NEAR_N = 0.2f;
constexpr float FAR_N = 100.0f;
constexpr float FILTER_INV_SQUARE = 2.0f;

// ---------------------------------------------------------------- depth slices
// Occlusion-aware binning (culled-emission mode): the Gaussians are cut, in depth order, into up to MAX_SLICES
// slices; slice k is emitted, sorted and composited before slice k+1 is emitted, and tiles whose pixels have all
// saturated (T < 1e-4 reached) receive no further instances.  On scenes with occlusion the compositing only ever
// visits the first few per cent of a tile's list (measured: 1.4-3.3 % at 1080p / 3M Gaussians), so almost all of the
// emission / sort / block-mask work on the later slices disappears.  Plan layout (host ints, b200gs.h):
//   [0] R   [1] S   [2 .. 2+MAX_SLICES] first Gaussian (depth order) of slice k   [11 .. 11+MAX_SLICES] first instance
constexpr int MAX_SLICES = B200GS_MAX_SLICES;
constexpr int PLAN_G = 2, PLAN_OFF = 2 + MAX_SLICES + 1, PLAN_INTS = B200GS_PLAN_INTS;
static_assert(PLAN_OFF + MAX_SLICES + 1 <= PLAN_INTS, "plan layout");

// ---------------------------------------------------------------- errors
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);
extern std::atomic<unsigned long long> g_launches;

#define B200GS_CUDA(expr)                                          \
  do {                                                             \
    cudaError_t _e = (expr);                                       \
    if (_e != cudaSuccess) return ::b200gs::cuda_fail(_e, #expr);  \
  } while (0)

// After a kernel launch: count it, pick up launch errors; in debug mode also
// drain the stream (the reference's CHECK_CUDA, S/cuda_rasterizer/auxiliary.h:300-307).
#define B200GS_LAUNCHED(what, debug, stream)                                   \
  do {                                                                         \
    ::b200gs::g_launches.fetch_add(1, std::memory_order_relaxed);              \
    cudaError_t _e = cudaGetLastError();                                       \
    if (_e == cudaSuccess && (debug)) _e = cudaStreamSynchronize(stream);      \
    if (_e != cudaSuccess) return ::b200gs::cuda_fail(_e, what);               \
  } while (0)

// ---------------------------------------------------------------- stage timing
// Optional per-stage device timing (b200gs_timing_enable / b200gs_timing_read):
// a CUDA event pair is recorded on the LAUNCHING stream around every stage, so
// bench.py can attribute the step time to kernels without a profiler attached.
enum Stage {
  ST_PREPROCESS = 0, ST_SCAN, ST_DUPLICATE, ST_SORT, ST_RANGES, ST_COMPOSITE_FWD,
  ST_COMPOSITE_BWD, ST_PREPROCESS_BWD, ST_CUBEMAP_FWD, ST_CUBEMAP_BWD, ST_TILE_HITS, ST_DEPTH_SORT, ST_COUNT_KEPT,
  ST_ADAM, ST_DENSIFY_STATS, ST_SSIM_FWD, ST_SSIM_BWD, ST_SHADE_FWD, ST_SHADE_BWD, ST_SURFACE_FWD, ST_SURFACE_BWD, ST_SCENE_IO, ST_COUNT
};
extern std::atomic<int> g_timing;        // 0 off, 1 every stage, 2 + k: stage k only
void stage_mark(int stage, bool begin, cudaStream_t s);
struct StageScope {
  int st; cudaStream_t s; bool on;
  StageScope(int stage, cudaStream_t stream) : st(stage), s(stream), on(false) {
    const int m = g_timing.load(std::memory_order_relaxed);
    on = m == 1 || m == 2 + stage;
    if (on) stage_mark(st, true, s);
  }
  ~StageScope() { if (on) stage_mark(st, false, s); }
};

// ---------------------------------------------------------------- scratch carving
// Same idea as the reference's obtain() (S/cuda_rasterizer/rasterizer_impl.h:21-27):
// carve typed arrays out of one byte buffer, each start aligned.
struct Carver {
  char* p;
  size_t used;
  explicit Carver(void* base) : p(reinterpret_cast<char*>(base)), used(0) {}
  template <typename T>
  T* take(size_t count, size_t align = 128) {
    used = (used + align - 1) / align * align;
    T* r = p ? reinterpret_cast<T*>(p + used) : nullptr;
    used += count * sizeof(T);
    return r;
  }
};

// Per-Gaussian packed record written by preprocess and staged (5 x 16 B) by the
// compositing kernels.  80 B, 16-B aligned.
//   q0 = Tu.x Tu.y Tu.z Tv.x      q1 = Tv.y Tv.z Tw.x Tw.y
//   q2 = Tw.z xy.x xy.y opacity   q3 = n.x  n.y  n.z  refl
//   q4 = r    g    b    mask(float 0/1)
struct __align__(16) SurfelRec {
  float4 q[5];
};
static_assert(sizeof(SurfelRec) == 80, "record must be 80 B");

// Per-Gaussian cull record (64 B) written by the surfel preprocess: the point conic of
// {rho3d <= c2} in pixel coordinates, f(x,y) = A x^2 + 2B xy + C y^2 + 2D x + 2E y + F <= 0
// (valid for ellipses AND for the hyperbolas of splats that cross the camera plane), in fp64
// and normalised to max(|A|,|B|,|C|) = 1, plus the low-pass disk {|p - (cx,cy)|^2 <= r2}.
// flag: 0 = use the test, CULL_EVERYTHING / CULL_NOTHING = fixed answer.  See cull_record().
struct __align__(16) CullRec {
  double A, B, C, D, E, F;
  float cx, cy, r2, flag;
};
static_assert(sizeof(CullRec) == 64, "cull record must be 64 B");
constexpr float CULL_EVERYTHING = -1.0f;
constexpr float CULL_NOTHING = -2.0f;

// Depth order of the Gaussians: ping-pong (depth bits, id) pairs of the P-sized radix sort.
struct DepthOrder {
  uint32_t *keys_a, *vals_a, *keys_b, *vals_b;   // result: vals_a = ids in (depth bits, id) order
  void* ws;
  size_t ws_bytes;
  static DepthOrder carve(Carver& c, size_t P);
};

// Row cache: the count pass stores, per (Gaussian, tile row of its rect), the reachable column ranges it computed
// (cull_rows.cuh, fp64) so that the emission of the same forward reads 16 B per row instead of repeating the
// arithmetic.  Rows of the 32 Gaussians of a warp are allocated with one atomicAdd on `cursor`; a Gaussian whose rows
// did not fit (capacity is a fixed multiple of P) is marked and recomputed by the emission.
struct RowCache {
  uint4* rows;            // [capacity] {a1 | (b1+1) << 16, a2 | (b2+1) << 16, d0 | (d1+1) << 16, tiles in the row}
  uint32_t* base;         // [P] first row of Gaussian id, or ROW_NONE / ROW_EMPTY
  uint32_t* cursor;       // rows handed out so far (zeroed with the flags by the preprocess)
  uint32_t capacity;
};
constexpr uint32_t ROW_NONE = 0xffffffffu;    // not cached: compute
constexpr uint32_t ROW_EMPTY = 0xfffffffeu;   // the Gaussian reaches no tile at all
constexpr int ROWS_PER_GAUSSIAN = 4;   // capacity = ROWS_PER_GAUSSIAN * P + 2^20

struct SurfelGeom {
  float* depths;            // [P]
  uint32_t* tiles_touched;  // [P] the reference's count (rect area)
  uint32_t* tiles_kept;     // [P] tiles of the rect the Gaussian can reach (culled emission)
  uint32_t* point_offsets;  // [P] inclusive prefix sum of tiles_kept / tiles_touched IN DEPTH ORDER (order.vals_a)
  DepthOrder order;
  uint8_t* clamped;         // [P] bit c = colour channel c clamped at 0
  uint2* rects;             // [P] tile rect: x = x0 | x1<<16, y = y0 | y1<<16 (only valid if tiles_touched>0)
  SurfelRec* rec;           // [P]
  CullRec* cull;            // [P] region where alpha >= 1/255 is reachable
  uint32_t* scan_state;     // decoupled look-back tile states
  int* flags;               // [4] error flags (prefiltered violation, ...)
  int* plan;                // [PLAN_INTS] the slice plan of this forward (device copy; the backward reads S from it)
  RowCache rows;            // reachable tile columns per (Gaussian, tile row): written by the count, read by the emission
  static SurfelGeom carve(void* base, size_t P, size_t* bytes);
};

constexpr int STATE_PLANES = 16;   // per-pixel compositing state carried from one depth slice to the next
struct SurfelImage {
  float* accum;        // [3,N]: final T, M1, M2
  uint32_t* n_contrib; // [2,N]: last contributor, median contributor
  uint2* ranges;       // [MAX_SLICES][tiles+1] absolute ranges of every slice's sorted list (+1: the sentinel key)
  uint2* cranges;      // [MAX_SLICES][tiles+1] ranges into the compacted instance list (Binning::comp_*)
  uint32_t* tile_base; // [MAX_SLICES+1][tiles] entries of the tile's (uncompacted) lists in the slices before k
  uint8_t* tile_done;  // [tiles] every pixel of the tile saturated in an earlier slice
  uint32_t* live;      // [4] [0] = tiles not yet done
  float* state;        // [STATE_PLANES,N] (only touched when a forward has more than one slice)
  static SurfelImage carve(void* base, int W, int H, size_t* bytes);
};

struct Binning {
  uint32_t* point_list;           // [R] Gaussian ids sorted by (tile, depth bits, id)
  uint32_t* tile_ids;             // [R] tile id of every sorted instance
  uint32_t* tmp_vals;             // [R] ping-pong partners of the two above
  uint32_t* tmp_keys;             // [R]
  // Compacted instance list: the sorted instances that can reach alpha >= 1/255 in at least one
  // 8x4 warp block of their tile (tile_hits_kernel), in list order, with their list position.
  uint32_t* comp_id;              // [<=R] Gaussian id
  uint32_t* comp_pos;             // [<=R] 0-based position in the tile's (uncompacted) list
  uint8_t* comp_hit;              // [<=R] bit w = warp block w of the tile can be reached
  char* zero;                     // radix workspaces + compaction look-back words of all depth slices:
  size_t zero_bytes;              //   cleared once per forward (bin_clear), carved slice by slice (bin_slice)
  static Binning carve(void* base, size_t R, size_t* bytes);
};

// What one forward bins: the depth-ordered Gaussians, their instance offsets and the slice plan.
struct BinJob {
  const uint32_t* order;          // Gaussian ids in depth order
  const uint2* rects;
  const uint32_t* offsets;        // inclusive prefix sum of the per-Gaussian instance counts, depth order
  const CullRec* cull;            // culled emission: the cull records; nullptr: the reference's full lists
  const CullRec* cull_for_hits;   // the cull records (block masks are computed in both modes)
  const RowCache* rows;           // culled emission: the count pass's per-row ranges, or nullptr
  int P, W, H, R;
  int slices;
  const int* g;                   // [slices+1] host
  const int* off;                 // [slices+1] host
  uint2* ranges;                  // [slices][tiles+1]
  uint2* cranges;
  const uint8_t* tile_done;       // device, read by the emission of slices > 0
  const uint32_t* live;           // device
  bool debug, audit;
  int tiles() const { return ((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y); }
};
int bin_clear(const Binning& b, const BinJob& j, cudaStream_t s);
int bin_slice(const BinJob& j, int k, const Binning& b, size_t& zcursor, cudaStream_t s);
int slice_plan(const uint32_t* point_offsets, int P, int max_slices, const float* frac, int* plan_dev, cudaStream_t s, bool debug);

// Backward accumulator: 20 floats per Gaussian, one contiguous 80-B segment so a
// (tile, Gaussian) partial is flushed with one vector of RED.ADD.F32.
//   0..8 dL_dtransMat   9,10 dL_dmean2D.xy   11 dL_dopacity
//   12..14 dL_dcolor    15 dL_drefl          16..18 dL_dnormal   19 pad
constexpr int ACC_STRIDE = 20;

// ---------------------------------------------------------------- 3DGS (EWA) path
// Packed per-Gaussian record of the 3DGS path, 64 B:
//   q0 = xy.x xy.y conic.x conic.y     q1 = conic.z opacity*h 1/depth refl
//   q2 = r g b depth                   q3 = n.x n.y n.z 0
struct __align__(16) GaussRec {
  float4 q[4];
};
static_assert(sizeof(GaussRec) == 64, "record must be 64 B");

struct GaussGeom {
  float* depths;            // [P]
  uint32_t* tiles_touched;  // [P]
  uint32_t* tiles_kept;     // [P] reachable tiles (culled emission)
  uint32_t* point_offsets;  // [P] in depth order
  DepthOrder order;
  uint8_t* clamped;         // [P] bit c
  uint2* rects;             // [P]
  GaussRec* rec;            // [P]
  float* cov3D;             // [P,6] (computed from scale/rotation; unused with cov3D_precomp)
  CullRec* cull;            // [P] {alpha >= 1/255} as a point conic (an ellipse here), same record as the surfel path
  uint32_t* scan_state;
  int* flags;
  int* plan;                // [PLAN_INTS] slice plan of this forward, as in SurfelGeom
  RowCache rows;            // as in SurfelGeom
  static GaussGeom carve(void* base, size_t P, size_t* bytes);
};

constexpr int GSTATE_PLANES = 10;  // T (sign = done), colour, normal, refl, inverse depth, last contributor
struct GaussImage {
  float* accum;         // [N] final T
  uint32_t* n_contrib;  // [N]
  uint2* ranges;        // [MAX_SLICES][tiles+1]   (depth slices: same bookkeeping as SurfelImage)
  uint2* cranges;       // [MAX_SLICES][tiles+1] ranges into the compacted instance list
  uint32_t* tile_base;  // [MAX_SLICES+1][tiles]
  uint8_t* tile_done;   // [tiles]
  uint32_t* live;       // [4]
  float* state;         // [GSTATE_PLANES,N]
  static GaussImage carve(void* base, int W, int H, size_t* bytes);
};

// Backward accumulator of the 3DGS path: 16 floats (64 B) per Gaussian.
//   0,1 dL_dmean2D.xy   2,3 dL_dmean2D_pixels.xy   4,5,6 dL_dconic x,y,w   7 dL_dopacity
//   8..10 dL_dcolor     11..13 dL_dnormal          14 dL_drefl             15 dL_dinvdepth
constexpr int GACC_STRIDE = 16;

int gauss_preprocess(const b200gs_gauss_fwd& a, const GaussGeom& g, cudaStream_t s);
int gauss_composite_forward(const b200gs_gauss_fwd& a, const GaussGeom& g, const GaussImage& im, const Binning& b,
                            int k, int S, cudaStream_t s);
int gauss_composite_backward(const b200gs_gauss_bwd& a, const GaussGeom& g, const GaussImage& im, const Binning& b,
                             float* acc, cudaStream_t s);
int gauss_preprocess_backward(const b200gs_gauss_bwd& a, const GaussGeom& g, const float* acc, cudaStream_t s);

// ---------------------------------------------------------------- device helpers
__device__ __forceinline__ int imin(int a, int b) { return a < b ? a : b; }
__device__ __forceinline__ int imax(int a, int b) { return a > b ? a : b; }

// getRect (S/cuda_rasterizer/auxiliary.h:71-81), with the exact fp32 operation
// order nvcc emits for it: (p - r) * (1/16) and ((p + r) + 16) + (-1)) * (1/16),
// truncation toward zero, clamp to [0, grid].
__device__ __forceinline__ void get_rect(float px, float py, int radius, int gx, int gy,
                                         int& x0, int& y0, int& x1, int& y1) {
  const float r = (float)radius;
  x0 = imin(gx, imax(0, (int)__fmul_rn(__fsub_rn(px, r), 0.0625f)));
  y0 = imin(gy, imax(0, (int)__fmul_rn(__fsub_rn(py, r), 0.0625f)));
  x1 = imin(gx, imax(0, (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(px, r), 16.0f), -1.0f), 0.0625f)));
  y1 = imin(gy, imax(0, (int)__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(py, r), 16.0f), -1.0f), 0.0625f)));
}

// transformPoint4x3 / transformVec4x3 (S/cuda_rasterizer/auxiliary.h:83-112) with
// the contraction the reference build has:  m12 + fma(z,m8, fma(x,m0, y*m4)).
__device__ __forceinline__ float dot3_ref(float x, float m0, float y, float m4, float z, float m8) {
  return __fmaf_rn(z, m8, __fmaf_rn(x, m0, __fmul_rn(y, m4)));
}

// SH basis constants (S/cuda_rasterizer/auxiliary.h:46-64).
__device__ constexpr float SH_C0 = 0.28209479177387814f;
__device__ constexpr float SH_C1 = 0.4886025119029199f;
__device__ constexpr float SH_C2_0 = 1.0925484305920792f;
__device__ constexpr float SH_C2_1 = -1.0925484305920792f;
__device__ constexpr float SH_C2_2 = 0.31539156525252005f;
__device__ constexpr float SH_C2_3 = -1.0925484305920792f;
__device__ constexpr float SH_C2_4 = 0.5462742152960396f;
__device__ constexpr float SH_C3_0 = -0.5900435899266435f;
__device__ constexpr float SH_C3_1 = 2.890611442640554f;
__device__ constexpr float SH_C3_2 = -0.4570457994644658f;
__device__ constexpr float SH_C3_3 = 0.3731763325901154f;
__device__ constexpr float SH_C3_4 = -0.4570457994644658f;
__device__ constexpr float SH_C3_5 = 1.445305721320277f;
__device__ constexpr float SH_C3_6 = -0.5900435899266435f;

// ---------------------------------------------------------------- stage launchers (internal)
int surfel_preprocess(const b200gs_surfel_fwd& a, const SurfelGeom& g, cudaStream_t s);
int inclusive_scan_u32(const uint32_t* in, uint32_t* out, int n, uint32_t* state, cudaStream_t s, bool debug,
                       const uint32_t* gather = nullptr);
int depth_order_and_scan(const float* depths, const uint32_t* tiles_touched, int P, const DepthOrder& o,
                         uint32_t* point_offsets, uint32_t* scan_state, cudaStream_t s, bool debug);
int surfel_count_kept(int P, const uint32_t* tiles_touched, const uint2* rects, const CullRec* cull, uint32_t* kept,
                      const RowCache* rows, cudaStream_t s, bool debug);
size_t tile_hits_state_words(size_t R);
// slice k of S (S == 1: the whole list in one go, no state carried)
int surfel_composite_forward(const b200gs_surfel_fwd& a, const SurfelGeom& g, const SurfelImage& im,
                             const Binning& b, int k, int S, cudaStream_t s);
int surfel_composite_backward(const b200gs_surfel_bwd& a, const SurfelGeom& g, const SurfelImage& im,
                              const Binning& b, float* acc, cudaStream_t s);
int surfel_preprocess_backward(const b200gs_surfel_bwd& a, const SurfelGeom& g, const float* acc, cudaStream_t s);

}  // namespace b200gs
