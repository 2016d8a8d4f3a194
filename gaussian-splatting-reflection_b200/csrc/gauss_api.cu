// gauss_api.cu -- extern "C" surface of the 3DGS (EWA) path (include/b200gs.h): argument
// checks, scratch carving, stage sequencing and the parity unpackers.
#include "common.cuh"

namespace b200gs {

size_t scan_state_words(int n);  // binning.cu

GaussGeom GaussGeom::carve(void* base, size_t P, size_t* bytes) {
  Carver c(base);
  GaussGeom g;
  g.depths = c.take<float>(P);
  g.tiles_touched = c.take<uint32_t>(P);
  g.tiles_kept = c.take<uint32_t>(P);
  g.point_offsets = c.take<uint32_t>(P);
  g.order = DepthOrder::carve(c, P);
  g.clamped = c.take<uint8_t>(P);
  g.rects = c.take<uint2>(P);
  g.rec = c.take<GaussRec>(P);
  g.cov3D = c.take<float>(6 * P);
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

GaussImage GaussImage::carve(void* base, int W, int H, size_t* bytes) {
  Carver c(base);
  GaussImage im;
  const size_t N = (size_t)W * H;
  const size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
  im.accum = c.take<float>(N);
  im.n_contrib = c.take<uint32_t>(N);
  im.ranges = c.take<uint2>((size_t)MAX_SLICES * (tiles + 1));
  im.cranges = c.take<uint2>((size_t)MAX_SLICES * (tiles + 1));
  im.tile_base = c.take<uint32_t>((size_t)(MAX_SLICES + 1) * tiles);
  im.tile_done = c.take<uint8_t>(tiles);
  im.live = c.take<uint32_t>(4);
  im.state = c.take<float>((size_t)GSTATE_PLANES * N);
  if (bytes) *bytes = c.used + 128;
  return im;
}

static inline void* align128(const void* p) {
  return reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(p) + 127) & ~uintptr_t(127));
}

static int check_fwd(const b200gs_gauss_fwd* a) {
  if (!a) { set_error("null argument struct"); return B200GS_EINVAL; }
  if (a->P < 0 || a->width <= 0 || a->height <= 0) { set_error("bad sizes P=%d W=%d H=%d", a->P, a->width, a->height); return B200GS_EINVAL; }
  if (a->P == 0) return B200GS_OK;
  if (!a->shs && !a->colors_precomp) {   // G/cuda_rasterizer/rasterizer_impl.cu:249-252
    set_error("For non-RGB, provide precomputed Gaussian colors!"); return B200GS_EINVAL; }
  if (a->shs && a->colors_precomp) { set_error("provide exactly one of shs / colors_precomp"); return B200GS_EINVAL; }
  const bool sr = a->scales && a->rotations;
  if (sr == (a->cov3D_precomp != nullptr)) { set_error("provide exactly one of scale/rotation pair or cov3D_precomp"); return B200GS_EINVAL; }
  if (a->shs && (a->sh_degree < 0 || a->sh_degree > 3 || (a->sh_degree + 1) * (a->sh_degree + 1) > a->sh_coeffs)) {
    set_error("sh_degree %d needs %d coefficients, shs has %d", a->sh_degree, (a->sh_degree + 1) * (a->sh_degree + 1), a->sh_coeffs);
    return B200GS_EINVAL; }
  if (!a->means3D || !a->opacities || !a->viewmatrix || !a->projmatrix || !a->campos || !a->background) {
    set_error("missing required input pointer"); return B200GS_EINVAL; }
  if (!a->out_color || !a->out_normal || !a->out_refl || !a->radii || !a->geom || !a->img) {
    set_error("missing output / scratch pointer"); return B200GS_EINVAL; }
  if (a->geom_bytes < b200gs_gauss_geom_bytes(a->P) || a->img_bytes < b200gs_gauss_image_bytes(a->width, a->height)) {
    set_error("geom/img scratch too small"); return B200GS_ESIZE; }
  return B200GS_OK;
}

__global__ void unpack_gauss_kernel(int P, const GaussRec* rec, const uint32_t* tiles, const uint8_t* clamped_bits,
                                    const float* depths_in, const float* cov3D_in, float* depths, float* means2D,
                                    float* cov3D, float* conic_opacity, float* rgb, uint8_t* clamped) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const bool vis = tiles[idx] > 0;
  float4 q[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) q[k] = vis ? rec[idx].q[k] : make_float4(0, 0, 0, 0);
  if (depths) depths[idx] = vis ? depths_in[idx] : 0.0f;
  if (means2D) { means2D[2 * idx] = q[0].x; means2D[2 * idx + 1] = q[0].y; }
  if (cov3D) for (int k = 0; k < 6; ++k) cov3D[6 * (size_t)idx + k] = vis ? cov3D_in[6 * (size_t)idx + k] : 0.0f;
  if (conic_opacity) {
    float* o = conic_opacity + 4 * (size_t)idx;
    o[0] = q[0].z; o[1] = q[0].w; o[2] = q[1].x; o[3] = q[1].y;
  }
  if (rgb) { rgb[3 * idx] = q[2].x; rgb[3 * idx + 1] = q[2].y; rgb[3 * idx + 2] = q[2].z; }
  if (clamped) {
    const unsigned cb = vis ? clamped_bits[idx] : 0u;
    clamped[3 * idx] = cb & 1u; clamped[3 * idx + 1] = (cb >> 1) & 1u; clamped[3 * idx + 2] = (cb >> 2) & 1u;
  }
}

}  // namespace b200gs

using namespace b200gs;

extern "C" {

size_t b200gs_gauss_geom_bytes(int P) {
  size_t b = 0;
  GaussGeom::carve(nullptr, (size_t)(P > 0 ? P : 0), &b);
  return b;
}
size_t b200gs_gauss_image_bytes(int width, int height) {
  size_t b = 0;
  GaussImage::carve(nullptr, width, height, &b);
  return b;
}
size_t b200gs_gauss_accum_bytes(int P) { return (size_t)(P > 0 ? P : 0) * GACC_STRIDE * sizeof(float) + 128; }

int b200gs_gauss_forward_stage1(const b200gs_gauss_fwd* a, int* plan, void* stream) {
  int rc = check_fwd(a);
  if (rc) return rc;
  if (!plan) { set_error("plan is null"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  for (int i = 0; i < PLAN_INTS; ++i) plan[i] = 0;
  plan[1] = 1;
  if (a->P == 0) return B200GS_OK;
  GaussGeom g = GaussGeom::carve(align128(a->geom), a->P, nullptr);
  rc = gauss_preprocess(*a, g, s);
  if (rc) return rc;
  const bool exact = a->exact_lists != 0;
  const uint32_t* counts = g.tiles_touched;
  if (!exact) {   // culled emission (include/b200gs.h: exact_lists)
    rc = surfel_count_kept(a->P, g.tiles_touched, g.rects, g.cull, g.tiles_kept, &g.rows, s, a->debug != 0);
    if (rc) return rc;
    counts = g.tiles_kept;
  }
  rc = depth_order_and_scan(g.depths, counts, a->P, g.order, g.point_offsets, g.scan_state, s, a->debug != 0);
  if (rc) return rc;
  rc = slice_plan(g.point_offsets, a->P, exact ? 1 : a->depth_slices, a->slice_frac, g.plan, s, a->debug != 0);
  if (rc) return rc;
  int flag = 0;
  B200GS_CUDA(cudaMemcpyAsync(plan, g.plan, PLAN_INTS * sizeof(int), cudaMemcpyDeviceToHost, s));
  B200GS_CUDA(cudaMemcpyAsync(&flag, g.flags, sizeof(int), cudaMemcpyDeviceToHost, s));
  B200GS_CUDA(cudaStreamSynchronize(s));
  if (flag) { set_error("Point is filtered although prefiltered is set. This shouldn't happen!"); return B200GS_EINVAL; }
  return B200GS_OK;
}

int b200gs_gauss_forward_stage2(const b200gs_gauss_fwd* a, const int* plan, void* stream) {
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
  if (S > 1 && a->exact_lists) { set_error("depth slices need the culled emission"); return B200GS_EINVAL; }
  if (R > 0 && (!a->binning || a->binning_bytes < b200gs_binning_bytes(R))) { set_error("binning scratch too small"); return B200GS_ESIZE; }
  GaussGeom g = GaussGeom::carve(align128(a->geom), a->P, nullptr);
  GaussImage im = GaussImage::carve(align128(a->img), a->width, a->height, nullptr);
  Binning b = Binning::carve(R > 0 ? align128(a->binning) : nullptr, R, nullptr);
  BinJob j{};
  j.order = g.order.vals_a; j.rects = g.rects; j.offsets = g.point_offsets;
  j.cull = a->exact_lists != 0 ? nullptr : g.cull; j.cull_for_hits = g.cull; j.rows = &g.rows;
  j.P = a->P; j.W = a->width; j.H = a->height; j.R = R;
  j.slices = S; j.g = plan + PLAN_G; j.off = plan + PLAN_OFF;
  j.ranges = im.ranges; j.cranges = im.cranges; j.tile_done = im.tile_done; j.live = im.live;
  j.debug = a->debug != 0; j.audit = false;
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
    rc = gauss_composite_forward(*a, g, im, b, k, S, s);
    if (rc) return rc;
  }
  return B200GS_OK;
}

int b200gs_gauss_backward(const b200gs_gauss_bwd* a, void* stream) {
  if (!a) { set_error("null argument struct"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  if (a->P <= 0) return B200GS_OK;
  if (!a->geom || !a->img || !a->accum || a->accum_bytes < b200gs_gauss_accum_bytes(a->P)) {
    set_error("missing scratch / accum too small"); return B200GS_ESIZE; }
  if (a->width <= 0 || a->height <= 0 || a->num_rendered < 0) { set_error("bad sizes W=%d H=%d R=%d", a->width, a->height, a->num_rendered); return B200GS_EINVAL; }
  if (a->geom_bytes < b200gs_gauss_geom_bytes(a->P) || a->img_bytes < b200gs_gauss_image_bytes(a->width, a->height) ||
      (a->num_rendered > 0 && (!a->binning || a->binning_bytes < b200gs_binning_bytes(a->num_rendered)))) {
    set_error("backward: geom / img / binning scratch smaller than the forward's (P=%d, %dx%d, num_rendered=%d)", a->P, a->width,
              a->height, a->num_rendered);
    return B200GS_ESIZE; }
  if (!a->dL_dout_color || !a->dL_dout_normal || !a->dL_dmeans2D || !a->dL_dopacity || !a->dL_dcolors || !a->dL_dnormals ||
      !a->dL_dmeans3D || !a->dL_dcov3D || !a->dL_dscales || !a->dL_drotations || !a->radii || !a->means3D || !a->opacities) {
    set_error("missing gradient pointer"); return B200GS_EINVAL; }
  GaussGeom g = GaussGeom::carve(align128(a->geom), a->P, nullptr);
  GaussImage im = GaussImage::carve(align128(a->img), a->width, a->height, nullptr);
  Binning b = Binning::carve(a->num_rendered > 0 ? align128(a->binning) : nullptr, a->num_rendered > 0 ? a->num_rendered : 0, nullptr);
  float* acc = reinterpret_cast<float*>(align128(a->accum));
  int rc = gauss_composite_backward(*a, g, im, b, acc, s);
  if (rc) return rc;
  return gauss_preprocess_backward(*a, g, acc, s);
}

int b200gs_gauss_inspect_geom(const void* geom, int P, float* depths, float* means2D, float* cov3D, float* conic_opacity,
                              float* rgb, uint32_t* tiles_touched, uint32_t* point_offsets, uint8_t* clamped, void* stream) {
  if (P <= 0) return B200GS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  GaussGeom g = GaussGeom::carve(align128(geom), P, nullptr);
  unpack_gauss_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, g.rec, g.tiles_touched, g.clamped, g.depths, g.cov3D, depths, means2D,
                                                      cov3D, conic_opacity, rgb, clamped);
  B200GS_LAUNCHED("unpack_gauss_kernel", false, s);
  if (tiles_touched) B200GS_CUDA(cudaMemcpyAsync(tiles_touched, g.tiles_touched, (size_t)P * 4, cudaMemcpyDeviceToDevice, s));
  // the reference's point_offsets: inclusive prefix of tiles_touched in id order (ours is in depth order)
  if (point_offsets) return inclusive_scan_u32(g.tiles_touched, point_offsets, P, g.scan_state, s, false);
  return B200GS_OK;
}

int b200gs_gauss_inspect_image(const void* img, int width, int height, float* accum_alpha, uint32_t* n_contrib,
                               uint32_t* ranges, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  GaussImage im = GaussImage::carve(align128(img), width, height, nullptr);
  const size_t N = (size_t)width * height;
  const size_t tiles = (size_t)((width + TILE_X - 1) / TILE_X) * ((height + TILE_Y - 1) / TILE_Y);
  if (accum_alpha) B200GS_CUDA(cudaMemcpyAsync(accum_alpha, im.accum, N * 4, cudaMemcpyDeviceToDevice, s));
  if (n_contrib) B200GS_CUDA(cudaMemcpyAsync(n_contrib, im.n_contrib, N * 4, cudaMemcpyDeviceToDevice, s));
  if (ranges) B200GS_CUDA(cudaMemcpyAsync(ranges, im.ranges, tiles * 8, cudaMemcpyDeviceToDevice, s));
  return B200GS_OK;
}

}  // extern "C"
