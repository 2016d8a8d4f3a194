// scene_io.cu -- the on-disk side of the path ("next" row 4): the vertex records of a
// point_cloud.ply, as GaussianModel.save_ply / load_ply write and read them (scene/gaussian_model.py:225-262,
// :296-373), to and from the per-Gaussian arrays the preprocess kernels read.
//
// A record is `stride` little-endian fp32 properties (x y z nx ny nz f_dc_* f_rest_* opacity refl scale_* rot_*
// in the reference's writer; the reader goes by NAME, so the host hands over a column map).  The reference
// walks the file one property at a time through float64 numpy columns, re-shapes f_rest from channel-major
// [3, K-1] to [K-1, 3] with a transpose + contiguous copy, and uploads seven tensors; here the raw record
// block is uploaded once and ONE kernel scatters it: a CTA stages PLY_VERTS whole records in shared memory
// with 128-bit coalesced loads (the block is contiguous), then writes every destination array with
// consecutive threads on consecutive destination floats.  Saving is the same kernel run the other way
// (properties no array feeds -- the normals -- are written as zeros, as the reference writes them).
//
// HBM-bound byte shuffling: 4 * stride bytes in + 4 * (mapped columns) bytes out per Gaussian (62 -> 59 floats,
// 484 B for the degree-3 surfel model).  No arithmetic: results are the same bits as the reference's.
#include "common.cuh"
#include "../../include/b200gs.h"

namespace b200gs {

constexpr int PLY_VERTS = 64;       // records per CTA
constexpr int PLY_THREADS = 256;

struct PlyFieldDev { float* data; int cols, row_stride, first; };
struct PlyArgs {
  PlyFieldDev field[B200GS_PLY_MAX_FIELDS];
  unsigned char map[B200GS_PLY_MAX_COLUMNS];   // map[field.first + c] = property column of the record
  int n_fields, stride;
  long long P;
  float* records;
};

template <bool PACK>
__global__ void __launch_bounds__(PLY_THREADS) ply_kernel(const __grid_constant__ PlyArgs a) {
  extern __shared__ __align__(16) float s_rec[];          // [PLY_VERTS][stride]
  __shared__ unsigned char s_map[B200GS_PLY_MAX_COLUMNS];
  const long long v0 = (long long)blockIdx.x * PLY_VERTS;
  const int nv = (int)min((long long)PLY_VERTS, a.P - v0);
  const int nfl = nv * a.stride;
  float* blk = a.records + v0 * a.stride;
  // (v0 * stride) is a multiple of 64 floats: the block start is 16-B aligned whenever `records` is
  const bool vec = (reinterpret_cast<uintptr_t>(a.records) & 15) == 0;
  if (!PACK) {
    if (vec) {
      for (int i = threadIdx.x; i < (nfl >> 2); i += PLY_THREADS)
        reinterpret_cast<float4*>(s_rec)[i] = __ldcs(reinterpret_cast<const float4*>(blk) + i);
      for (int i = (nfl & ~3) + threadIdx.x; i < nfl; i += PLY_THREADS) s_rec[i] = __ldcs(blk + i);
    } else {
      for (int i = threadIdx.x; i < nfl; i += PLY_THREADS) s_rec[i] = __ldcs(blk + i);
    }
  } else {
    for (int i = threadIdx.x; i < nfl; i += PLY_THREADS) s_rec[i] = 0.0f;
  }
  // the column map goes to shared memory: lanes index it with different c, which the constant bank would serialise
  for (int i = threadIdx.x; i < B200GS_PLY_MAX_COLUMNS; i += PLY_THREADS) s_map[i] = a.map[i];
  __syncthreads();
  for (int f = 0; f < a.n_fields; ++f) {
    const PlyFieldDev fd = a.field[f];
    float* dst = fd.data + v0 * fd.row_stride;
    const int n = nv * fd.cols;
    // (v, c) of element i = v * cols + c, advanced by PLY_THREADS per trip without dividing again
    const int qs = PLY_THREADS / fd.cols, rs = PLY_THREADS - qs * fd.cols;
    int v = threadIdx.x / fd.cols, c = threadIdx.x - v * fd.cols;
    const unsigned char* mp = s_map + fd.first;
    for (int i = threadIdx.x; i < n; i += PLY_THREADS) {
      float* g = dst + (long long)v * fd.row_stride + c;
      float* s = s_rec + v * a.stride + mp[c];
      if (PACK) *s = __ldcs(g); else __stcs(g, *s);
      v += qs; c += rs;
      if (c >= fd.cols) { c -= fd.cols; ++v; }
    }
  }
  if (PACK) {
    __syncthreads();
    if (vec) {
      for (int i = threadIdx.x; i < (nfl >> 2); i += PLY_THREADS)
        __stcs(reinterpret_cast<float4*>(blk) + i, reinterpret_cast<const float4*>(s_rec)[i]);
      for (int i = (nfl & ~3) + threadIdx.x; i < nfl; i += PLY_THREADS) __stcs(blk + i, s_rec[i]);
    } else {
      for (int i = threadIdx.x; i < nfl; i += PLY_THREADS) __stcs(blk + i, s_rec[i]);
    }
  }
}

static int ply_launch(bool pack, long long P, float* records, const b200gs_ply_layout* L, void* stream) {
  if (!L || P < 0) { set_error("ply: bad arguments"); return B200GS_EINVAL; }
  if (L->stride < 1 || L->stride > B200GS_PLY_MAX_COLUMNS) { set_error("ply: record stride out of range"); return B200GS_EINVAL; }
  if (L->n_fields < 0 || L->n_fields > B200GS_PLY_MAX_FIELDS) { set_error("ply: too many fields"); return B200GS_EINVAL; }
  PlyArgs a{};
  a.n_fields = L->n_fields; a.stride = L->stride; a.P = P; a.records = records;
  unsigned char used[B200GS_PLY_MAX_COLUMNS] = {};
  int next = 0;
  for (int f = 0; f < L->n_fields; ++f) {
    const b200gs_ply_field& s = L->field[f];
    if (s.cols < 1 || s.row_stride < s.cols || s.first != next || s.first + s.cols > B200GS_PLY_MAX_COLUMNS)
      { set_error("ply: bad field description"); return B200GS_EINVAL; }
    if (P > 0 && !s.data) { set_error("ply: null field pointer"); return B200GS_EINVAL; }
    for (int c = 0; c < s.cols; ++c) {
      int col = L->map[s.first + c];
      if (col >= L->stride) { set_error("ply: mapped column outside the record"); return B200GS_EINVAL; }
      if (used[col]++) { set_error("ply: record column mapped twice"); return B200GS_EINVAL; }
      a.map[s.first + c] = (unsigned char)col;
    }
    next += s.cols;
    a.field[f] = PlyFieldDev{s.data, s.cols, s.row_stride, s.first};
  }
  if (P == 0) return B200GS_OK;
  if (!records) { set_error("ply: null record block"); return B200GS_EINVAL; }
  cudaStream_t cs = (cudaStream_t)stream;
  StageScope ts(ST_SCENE_IO, cs);
  const long long grid = (P + PLY_VERTS - 1) / PLY_VERTS;
  if (grid > 0x7fffffffLL) { set_error("ply: too many records"); return B200GS_EINVAL; }
  const size_t smem = (size_t)PLY_VERTS * L->stride * sizeof(float);
  if (pack) ply_kernel<true><<<(unsigned)grid, PLY_THREADS, smem, cs>>>(a);
  else ply_kernel<false><<<(unsigned)grid, PLY_THREADS, smem, cs>>>(a);
  B200GS_LAUNCHED("ply_kernel", false, cs);
  return B200GS_OK;
}

}  // namespace b200gs

extern "C" int b200gs_ply_unpack(long long P, const float* records, const b200gs_ply_layout* layout, void* stream) {
  return b200gs::ply_launch(false, P, const_cast<float*>(records), layout, stream);
}
extern "C" int b200gs_ply_pack(long long P, const b200gs_ply_layout* layout, float* records, void* stream) {
  return b200gs::ply_launch(true, P, records, layout, stream);
}
