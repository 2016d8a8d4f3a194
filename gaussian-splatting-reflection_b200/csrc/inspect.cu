// inspect.cu -- markVisible and the parity-check unpackers that turn the private
// scratch layouts back into the reference's SoA state arrays
// (S/cuda_rasterizer/rasterizer_impl.h:33-69).
#include "common.cuh"

namespace b200gs {

// checkFrustum (S/cuda_rasterizer/rasterizer_impl.cu:54-66, auxiliary.h:189-214)
__global__ void mark_visible_kernel(int P, const float* __restrict__ p, const float* __restrict__ vm, uint8_t* present) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const float z = __fadd_rn(vm[14], dot3_ref(p[3 * idx], vm[2], p[3 * idx + 1], vm[6], p[3 * idx + 2], vm[10]));
  present[idx] = z > 0.2f;
}

__global__ void unpack_geom_kernel(int P, const SurfelRec* rec, const uint32_t* tiles, const uint8_t* clamped_bits,
                                   float* means2D, float* transMat, float* normal_opacity, float* rgb, uint8_t* clamped) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= P) return;
  const bool vis = tiles[idx] > 0;
  float4 q[5];
#pragma unroll
  for (int k = 0; k < 5; ++k) q[k] = vis ? rec[idx].q[k] : make_float4(0, 0, 0, 0);
  if (means2D) { means2D[2 * idx] = q[2].y; means2D[2 * idx + 1] = q[2].z; }
  if (transMat) {
    float* t = transMat + 9 * (size_t)idx;
    t[0] = q[0].x; t[1] = q[0].y; t[2] = q[0].z; t[3] = q[0].w; t[4] = q[1].x; t[5] = q[1].y;
    t[6] = q[1].z; t[7] = q[1].w; t[8] = q[2].x;
  }
  if (normal_opacity) {
    float* o = normal_opacity + 4 * (size_t)idx;
    o[0] = q[3].x; o[1] = q[3].y; o[2] = q[3].z; o[3] = q[2].w;
  }
  if (rgb) { rgb[3 * idx] = q[4].x; rgb[3 * idx + 1] = q[4].y; rgb[3 * idx + 2] = q[4].z; }
  if (clamped) {
    const unsigned cb = vis ? clamped_bits[idx] : 0u;
    clamped[3 * idx] = cb & 1u; clamped[3 * idx + 1] = (cb >> 1) & 1u; clamped[3 * idx + 2] = (cb >> 2) & 1u;
  }
}

// The reference's 64-bit sort keys (rasterizer_impl.cu:100-104): tile << 32 | depth bits.
__global__ void rebuild_keys_kernel(int R, const uint32_t* tile_ids, const uint32_t* point_list, const float* depths,
                                    uint64_t* keys) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < R) keys[i] = ((uint64_t)tile_ids[i] << 32) | __float_as_uint(depths[point_list[i]]);
}

__global__ void mask_depth_kernel(int P, const float* depths, const uint32_t* tiles, float* out) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < P) out[idx] = tiles[idx] > 0 ? depths[idx] : 0.0f;
}

}  // namespace b200gs

using namespace b200gs;

extern "C" {

int b200gs_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present,
                        void* stream) {
  (void)projmatrix;  // the reference computes p_hom but only tests the view-space depth
  if (P <= 0) return B200GS_OK;
  if (!means3D || !viewmatrix || !present) { set_error("mark_visible: null pointer"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, viewmatrix, present);
  B200GS_LAUNCHED("mark_visible_kernel", false, s);
  return B200GS_OK;
}

int b200gs_surfel_inspect_geom(const void* geom, int P, float* depths, float* means2D, float* transMat,
                               float* normal_opacity, float* rgb, uint32_t* tiles_touched, uint32_t* point_offsets,
                               uint8_t* clamped, void* stream) {
  if (P <= 0) return B200GS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  void* base = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(geom) + 127) & ~uintptr_t(127));
  SurfelGeom g = SurfelGeom::carve(base, P, nullptr);
  // Entries of culled Gaussians are never written by preprocess; report them as 0.
  if (depths) { mask_depth_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, g.depths, g.tiles_touched, depths); B200GS_LAUNCHED("mask_depth_kernel", false, s); }
  unpack_geom_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, g.rec, g.tiles_touched, g.clamped, means2D, transMat, normal_opacity, rgb, clamped);
  B200GS_LAUNCHED("unpack_geom_kernel", false, s);
  if (tiles_touched) B200GS_CUDA(cudaMemcpyAsync(tiles_touched, g.tiles_touched, (size_t)P * 4, cudaMemcpyDeviceToDevice, s));
  // the reference's point_offsets: inclusive prefix of tiles_touched in id order (ours is in depth order)
  if (point_offsets) return inclusive_scan_u32(g.tiles_touched, point_offsets, P, g.scan_state, s, false);
  return B200GS_OK;
}

int b200gs_inspect_binning(const void* binning, int R, const float* depths, uint64_t* keys, uint32_t* point_list,
                           void* stream) {
  if (R <= 0) return B200GS_OK;
  cudaStream_t s = (cudaStream_t)stream;
  void* base = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(binning) + 127) & ~uintptr_t(127));
  Binning b = Binning::carve(base, R, nullptr);
  if (keys) {
    if (!depths) { set_error("inspect_binning: keys need the per-Gaussian depths (from *_inspect_geom)"); return B200GS_EINVAL; }
    rebuild_keys_kernel<<<(R + 255) / 256, 256, 0, s>>>(R, b.tile_ids, b.point_list, depths, keys);
    B200GS_LAUNCHED("rebuild_keys_kernel", false, s);
  }
  if (point_list) B200GS_CUDA(cudaMemcpyAsync(point_list, b.point_list, (size_t)R * 4, cudaMemcpyDeviceToDevice, s));
  return B200GS_OK;
}

int b200gs_surfel_inspect_image(const void* img, int W, int H, float* accum_alpha, uint32_t* n_contrib, uint32_t* ranges,
                                void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  void* base = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(img) + 127) & ~uintptr_t(127));
  SurfelImage im = SurfelImage::carve(base, W, H, nullptr);
  const size_t N = (size_t)W * H;
  const size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
  if (accum_alpha) B200GS_CUDA(cudaMemcpyAsync(accum_alpha, im.accum, 3 * N * 4, cudaMemcpyDeviceToDevice, s));
  if (n_contrib) B200GS_CUDA(cudaMemcpyAsync(n_contrib, im.n_contrib, 2 * N * 4, cudaMemcpyDeviceToDevice, s));
  if (ranges) B200GS_CUDA(cudaMemcpyAsync(ranges, im.ranges, tiles * 8, cudaMemcpyDeviceToDevice, s));
  return B200GS_OK;
}

int b200gs_surfel_inspect_tiles(const void* img, int W, int H, uint8_t* tile_done, uint32_t* live, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  void* base = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(img) + 127) & ~uintptr_t(127));
  SurfelImage im = SurfelImage::carve(base, W, H, nullptr);
  const size_t tiles = (size_t)((W + TILE_X - 1) / TILE_X) * ((H + TILE_Y - 1) / TILE_Y);
  if (tile_done) B200GS_CUDA(cudaMemcpyAsync(tile_done, im.tile_done, tiles, cudaMemcpyDeviceToDevice, s));
  if (live) B200GS_CUDA(cudaMemcpyAsync(live, im.live, sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
  return B200GS_OK;
}

}  // extern "C"
