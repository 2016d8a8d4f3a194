// api.cu -- the extern "C" surface of libb200gs.so (see include/b200gs.h):
// argument checks, scratch carving, stage sequencing.  No torch, no CPU fallback.
#include "common.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

namespace b200gs {

static thread_local char t_err[512] = "";
std::atomic<unsigned long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error in %s: %s", what, cudaGetErrorString(e));
  return B200GS_ECUDA;
}

size_t scan_state_words(int n);  // binning.cu

// ---- stage timing -------------------------------------------------------
std::atomic<int> g_timing{0};
namespace {
struct Mark { cudaEvent_t a, b; int stage; };
std::mutex g_tmu;
std::vector<Mark> g_marks;        // completed or in-flight pairs
std::vector<cudaEvent_t> g_free;  // recycled events
cudaEvent_t get_event() {
  if (!g_free.empty()) { cudaEvent_t e = g_free.back(); g_free.pop_back(); return e; }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
}  // namespace
void stage_mark(int stage, bool begin, cudaStream_t s) {
  std::lock_guard<std::mutex> lk(g_tmu);
  if (begin) {
    Mark m{get_event(), nullptr, stage};
    cudaEventRecord(m.a, s);
    g_marks.push_back(m);
  } else {
    for (size_t i = g_marks.size(); i-- > 0;)
      if (g_marks[i].stage == stage && g_marks[i].b == nullptr) {
        g_marks[i].b = get_event();
        cudaEventRecord(g_marks[i].b, s);
        break;
      }
  }
}

SurfelGeom SurfelGeom::carve(void* base, size_t P, size_t* bytes) {
  Carver c(base);
  SurfelGeom g;
  g.depths = c.take<float>(P);
  g.tiles_touched = c.take<uint32_t>(P);
  g.tiles_kept = c.take<uint32_t>(P);
  g.point_offsets = c.take<uint32_t>(P);
  g.order = DepthOrder::carve(c, P);
  g.clamped = c.take<uint8_t>(P);
  g.rects = c.take<uint2>(P);
  g.rec = c.take<SurfelRec>(P);
  g.cull = c.take<CullRec>(P);
  g.scan_state = c.take<uint32_t>(scan_state_words((int)P));
  g.flags = c.take<int>(4);
  g.plan = c.take<int>(PLAN_INTS);
  g.rows.capacity = (uint32_t)(ROWS_PER_GAUSSIAN * P + (1u << 20));
  g.rows.rows = c.take<uint4>(g.rows.capacity);
  g.rows.base = c.take<uint32_t>(P);
  g.rows.cursor = g.flags ? reinterpret_cast<uint32_t*>(g.flags + 1) : nullptr;
  if (bytes) *bytes = c.used + 128;
  return g;
}

SurfelImage SurfelImage::carve(void* base, int W, int H, size_t* bytes) {
  Carver c(base);
  SurfelImage im;
  const size_t N = (size_t)W * H;
  const size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
  im.accum = c.take<float>(3 * N);
  im.n_contrib = c.take<uint32_t>(2 * N);
  im.ranges = c.take<uint2>((size_t)MAX_SLICES * (tiles + 1));
  im.cranges = c.take<uint2>((size_t)MAX_SLICES * (tiles + 1));
  im.tile_base = c.take<uint32_t>((size_t)(MAX_SLICES + 1) * tiles);
  im.tile_done = c.take<uint8_t>(tiles);
  im.live = c.take<uint32_t>(4);
  im.state = c.take<float>((size_t)STATE_PLANES * N);
  if (bytes) *bytes = c.used + 128;
  return im;
}

// Scratch handed in by the caller may start at any address; carve from the next
// 128-B boundary (the carve sizes include that slack).
static inline void* align128(const void* p) {
  return reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(p) + 127) & ~uintptr_t(127));
}

static int check_fwd(const b200gs_surfel_fwd* a) {
  if (!a) { set_error("null argument struct"); return B200GS_EINVAL; }
  if (a->P < 0 || a->width <= 0 || a->height <= 0) { set_error("bad sizes P=%d W=%d H=%d", a->P, a->width, a->height); return B200GS_EINVAL; }
  if ((a->shs == nullptr) == (a->colors_precomp == nullptr)) {
    set_error("provide exactly one of shs / colors_precomp"); return B200GS_EINVAL; }
  const bool sr = a->scales && a->rotations;
  if (sr == (a->transMat_precomp != nullptr) || (!sr && (a->scales || a->rotations))) {
    set_error("provide exactly one of scale/rotation pair or transMat_precomp"); return B200GS_EINVAL; }
  if (a->shs && (a->sh_degree < 0 || a->sh_degree > 3 || (a->sh_degree + 1) * (a->sh_degree + 1) > a->sh_coeffs)) {
    set_error("sh_degree %d needs %d coefficients, shs has %d", a->sh_degree, (a->sh_degree + 1) * (a->sh_degree + 1), a->sh_coeffs);
    return B200GS_EINVAL; }
  if (!a->means3D || !a->opacities || !a->viewmatrix || !a->projmatrix || !a->campos || !a->background) {
    set_error("missing required input pointer"); return B200GS_EINVAL; }
  if (!a->out_color || !a->out_others || !a->out_refl || !a->radii || !a->gaussian_weights || !a->geom || !a->img) {
    set_error("missing output / scratch pointer"); return B200GS_EINVAL; }
  if (a->geom_bytes < b200gs_surfel_geom_bytes(a->P) || a->img_bytes < b200gs_surfel_image_bytes(a->width, a->height)) {
    set_error("geom/img scratch too small"); return B200GS_ESIZE; }
  return B200GS_OK;
}

}  // namespace b200gs

using namespace b200gs;

extern "C" {

const char* b200gs_last_error(void) { return t_err; }
int b200gs_version(void) { return 1; }
int b200gs_sm_arch(void) { return 100; }
unsigned long long b200gs_launch_count(void) { return g_launches.load(); }
size_t b200gs_struct_bytes(int which) {
  switch (which) {
    case 0: return sizeof(b200gs_surfel_fwd);
    case 1: return sizeof(b200gs_surfel_bwd);
    case 2: return sizeof(b200gs_gauss_fwd);
    case 3: return sizeof(b200gs_gauss_bwd);
    case 4: return sizeof(b200gs_adam);
    case 5: return sizeof(b200gs_ply_layout);
    default: return 0;
  }
}

void b200gs_timing_enable(int on) { g_timing.store(on ? 1 : 0); }
void b200gs_timing_only(int stage) { g_timing.store((stage >= 0 && stage < ST_COUNT) ? 2 + stage : 0); }
int b200gs_timing_stages(void) { return ST_COUNT; }
const char* b200gs_timing_stage_name(int stage) {
  static const char* names[ST_COUNT] = {"preprocess", "scan", "duplicate", "sort", "tile_ranges", "composite_fwd",
                                        "composite_bwd", "preprocess_bwd", "cubemap_fwd", "cubemap_bwd", "tile_hits", "depth_sort", "count_kept",
                                        "adam", "densify_stats", "ssim_fwd", "ssim_bwd", "shade_fwd", "shade_bwd", "surface_fwd", "surface_bwd", "scene_io"};
  return (stage >= 0 && stage < ST_COUNT) ? names[stage] : "";
}
int b200gs_timing_read(float* ms, unsigned long long* counts, int n) {
  std::lock_guard<std::mutex> lk(g_tmu);
  for (int i = 0; i < n; ++i) { if (ms) ms[i] = 0.f; if (counts) counts[i] = 0; }
  for (auto& m : g_marks) {
    if (m.b) {
      float t = 0.f;
      if (cudaEventSynchronize(m.b) == cudaSuccess && cudaEventElapsedTime(&t, m.a, m.b) == cudaSuccess && m.stage < n) {
        if (ms) ms[m.stage] += t;
        if (counts) counts[m.stage] += 1;
      }
      g_free.push_back(m.b);
    }
    g_free.push_back(m.a);
  }
  g_marks.clear();
  return B200GS_OK;
}

size_t b200gs_surfel_geom_bytes(int P) {
  size_t b = 0;
  SurfelGeom::carve(nullptr, (size_t)(P > 0 ? P : 0), &b);
  return b;
}
size_t b200gs_surfel_image_bytes(int width, int height) {
  size_t b = 0;
  SurfelImage::carve(nullptr, width, height, &b);
  return b;
}
size_t b200gs_binning_bytes(int num_rendered) {
  size_t b = 0;
  Binning::carve(nullptr, (size_t)(num_rendered > 0 ? num_rendered : 0), &b);
  return b;
}
size_t b200gs_surfel_accum_bytes(int P) { return (size_t)(P > 0 ? P : 0) * ACC_STRIDE * sizeof(float) + 128; }

int b200gs_surfel_forward_stage1(const b200gs_surfel_fwd* a, int* plan, void* stream) {
  int rc = check_fwd(a);
  if (rc) return rc;
  if (!plan) { set_error("plan is null"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  for (int i = 0; i < PLAN_INTS; ++i) plan[i] = 0;
  plan[1] = 1;
  if (a->P == 0) return B200GS_OK;
  SurfelGeom g = SurfelGeom::carve(align128(a->geom), a->P, nullptr);
  rc = surfel_preprocess(*a, g, s);
  if (rc) return rc;
  const bool exact = a->exact_lists != 0;
  const uint32_t* counts = g.tiles_touched;
  if (!exact) {
    rc = surfel_count_kept(a->P, g.tiles_touched, g.rects, g.cull, g.tiles_kept, &g.rows, s, a->debug != 0);
    if (rc) return rc;
    counts = g.tiles_kept;
  }
  rc = depth_order_and_scan(g.depths, counts, a->P, g.order, g.point_offsets, g.scan_state, s, a->debug != 0);
  if (rc) return rc;
  // depth slices only make sense for the culled emission (the exact lists are the reference's, in one piece)
  rc = slice_plan(g.point_offsets, a->P, exact ? 1 : a->depth_slices, a->slice_frac, g.plan, s, a->debug != 0);
  if (rc) return rc;
  int flag = 0;
  B200GS_CUDA(cudaMemcpyAsync(plan, g.plan, PLAN_INTS * sizeof(int), cudaMemcpyDeviceToHost, s));
  B200GS_CUDA(cudaMemcpyAsync(&flag, g.flags, sizeof(int), cudaMemcpyDeviceToHost, s));
  B200GS_CUDA(cudaStreamSynchronize(s));
  if (flag) { set_error("Point is filtered although prefiltered is set. This shouldn't happen!"); return B200GS_EINVAL; }
  return B200GS_OK;
}

int b200gs_surfel_forward_stage2(const b200gs_surfel_fwd* a, const int* plan, void* stream) {
  int rc = check_fwd(a);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (a->P == 0) return B200GS_OK;
  if (!plan) { set_error("plan is null"); return B200GS_EINVAL; }
  const int R = plan[0], S = plan[1];
  if (R < 0 || S < 1 || S > MAX_SLICES || plan[PLAN_G] != 0 || plan[PLAN_OFF] != 0 || plan[PLAN_G + S] != a->P || plan[PLAN_OFF + S] != R) {
    set_error("bad slice plan (num_rendered %d, %d slices)", R, S); return B200GS_EINVAL; }
  for (int k = 0; k < S; ++k)
    if (plan[PLAN_G + k + 1] < plan[PLAN_G + k] || plan[PLAN_OFF + k + 1] < plan[PLAN_OFF + k]) { set_error("slice plan not ascending"); return B200GS_EINVAL; }
  if (S > 1 && (a->exact_lists || a->cull_audit)) { set_error("depth slices need the culled emission"); return B200GS_EINVAL; }
  if (R > 0 && (!a->binning || a->binning_bytes < b200gs_binning_bytes(R))) { set_error("binning scratch too small"); return B200GS_ESIZE; }
  SurfelGeom g = SurfelGeom::carve(align128(a->geom), a->P, nullptr);
  SurfelImage im = SurfelImage::carve(align128(a->img), a->width, a->height, nullptr);
  Binning b = Binning::carve(R > 0 ? align128(a->binning) : nullptr, R, nullptr);
  BinJob j{};
  j.order = g.order.vals_a; j.rects = g.rects; j.offsets = g.point_offsets;
  j.cull = a->exact_lists != 0 ? nullptr : g.cull; j.cull_for_hits = g.cull; j.rows = &g.rows;
  j.P = a->P; j.W = a->width; j.H = a->height; j.R = R;
  j.slices = S; j.g = plan + PLAN_G; j.off = plan + PLAN_OFF;
  j.ranges = im.ranges; j.cranges = im.cranges; j.tile_done = im.tile_done; j.live = im.live;
  j.debug = a->debug != 0; j.audit = a->cull_audit != 0;
  rc = bin_clear(b, j, s);
  if (rc) return rc;
  if (S > 1) {
    const uint32_t tiles = (uint32_t)j.tiles();
    B200GS_CUDA(cudaMemsetAsync(im.tile_done, 0, tiles, s));
    const uint32_t live0[4] = {tiles, 0u, 0u, 0u};
    B200GS_CUDA(cudaMemcpyAsync(im.live, live0, sizeof(live0), cudaMemcpyHostToDevice, s));
  }
  size_t zcursor = 0;
  for (int k = 0; k < S; ++k) {
    if (R > 0) {
      rc = bin_slice(j, k, b, zcursor, s);
      if (rc) return rc;
    }
    rc = surfel_composite_forward(*a, g, im, b, k, S, s);
    if (rc) return rc;
  }
  return B200GS_OK;
}

int b200gs_surfel_backward(const b200gs_surfel_bwd* a, void* stream) {
  if (!a) { set_error("null argument struct"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  if (a->P <= 0) return B200GS_OK;
  if (!a->geom || !a->img || !a->accum || a->accum_bytes < b200gs_surfel_accum_bytes(a->P)) {
    set_error("missing scratch / accum too small"); return B200GS_ESIZE; }
  if (a->width <= 0 || a->height <= 0 || a->num_rendered < 0) { set_error("bad sizes W=%d H=%d R=%d", a->width, a->height, a->num_rendered); return B200GS_EINVAL; }
  if (a->geom_bytes < b200gs_surfel_geom_bytes(a->P) || a->img_bytes < b200gs_surfel_image_bytes(a->width, a->height) ||
      (a->num_rendered > 0 && (!a->binning || a->binning_bytes < b200gs_binning_bytes(a->num_rendered)))) {
    set_error("backward: geom / img / binning scratch smaller than the forward's (P=%d, %dx%d, num_rendered=%d)", a->P, a->width,
              a->height, a->num_rendered);
    return B200GS_ESIZE; }
  if (!a->dL_dout_color || !a->dL_dout_others || !a->dL_dmeans2D || !a->dL_dnormal || !a->dL_dopacity || !a->dL_dcolors ||
      !a->dL_dmeans3D || !a->dL_dtransMat || !a->dL_dscales || !a->dL_drotations || !a->radii || !a->means3D) {
    set_error("missing gradient pointer"); return B200GS_EINVAL; }
  SurfelGeom g = SurfelGeom::carve(align128(a->geom), a->P, nullptr);
  SurfelImage im = SurfelImage::carve(align128(a->img), a->width, a->height, nullptr);
  Binning b = Binning::carve(a->num_rendered > 0 ? align128(a->binning) : nullptr, a->num_rendered > 0 ? a->num_rendered : 0, nullptr);
  float* acc = reinterpret_cast<float*>(align128(a->accum));
  int rc = surfel_composite_backward(*a, g, im, b, acc, s);
  if (rc) return rc;
  return surfel_preprocess_backward(*a, g, acc, s);
}

}  // extern "C"
