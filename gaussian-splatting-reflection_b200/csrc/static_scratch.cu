// static_scratch.cu -- b200gs_surfel_backward_static: the surfel backward (compositing backward K4 + preprocess
// backward K5) for callers that cannot allocate -- code under CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc.  No counterpart in the reference.
//
// The only scratch the backward needs beyond what the forward left behind is the per-Gaussian accumulator K4
// reduces into and K5 reads (ACC_STRIDE = 20 floats per Gaussian, `accum` of b200gs_surfel_bwd).  Here it is a
// file-scope __device__ array of a fixed maximum size (B200GS_STATIC_ACCUM_MAX_POINTS Gaussians, 10 MiB) instead of a
// caller allocation: every GPU that loads the library holds it until the process exits.
//
//  * In host code `g_static_accum` is a host-side stand-in for the symbol, NOT a device address: its device
//    address is looked up once per device, with that device current, by cudaGetSymbolAddress, and that pointer is
//    what the kernels are given.
//  * The entry point calls no cudaMalloc / cudaFree / allocating library routine; it copies the argument struct on
//    the host stack, points `accum` at the array and runs the same two stages as b200gs_surfel_backward.
//  * Every call on a device shares the one array.  It keeps the previous call's contents; K4's launcher clears the
//    P rows it uses (cudaMemsetAsync on the caller's stream) before accumulating, so nothing has to be zeroed by
//    the caller.  Two calls that overlap on the GPU (different streams without an ordering between them) RACE on
//    the array: serialise them.
#include <mutex>

#include "common.cuh"

namespace b200gs {

constexpr int STATIC_ACCUM_MAX_POINTS = B200GS_STATIC_ACCUM_MAX_POINTS;
constexpr size_t STATIC_ACCUM_FLOATS = (size_t)STATIC_ACCUM_MAX_POINTS * ACC_STRIDE + 32;   // + the 128 B of slack accum_bytes asks for

static __device__ __align__(256) float g_static_accum[STATIC_ACCUM_FLOATS];

constexpr int STATIC_MAX_DEVICES = 64;
static std::mutex g_static_mu;
static float* g_static_ptr[STATIC_MAX_DEVICES] = {};

// device address of the array on the current device (looked up at first use on that device)
static int static_accum_ptr(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { set_error("surfel_backward_static: cudaGetDevice: %s", cudaGetErrorString(e)); return B200GS_ECUDA; }
  if (dev < 0 || dev >= STATIC_MAX_DEVICES) { set_error("surfel_backward_static: device ordinal %d out of range", dev); return B200GS_EINVAL; }
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (!g_static_ptr[dev]) {
    float* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_accum);
    if (e != cudaSuccess || !p) {
      set_error("surfel_backward_static: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
      return B200GS_ECUDA;
    }
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return B200GS_OK;
}

}  // namespace b200gs

using namespace b200gs;

extern "C" int b200gs_surfel_backward_static(const b200gs_surfel_bwd* a, void* stream) {
  if (!a) { set_error("null argument struct"); return B200GS_EINVAL; }
  if (a->P <= 0) return B200GS_OK;
  if (a->P > STATIC_ACCUM_MAX_POINTS) {
    set_error("surfel_backward_static: P = %d exceeds the static scratch's maximum of %d Gaussians "
              "(B200GS_STATIC_ACCUM_MAX_POINTS); use b200gs_surfel_backward with a caller-allocated accum",
              a->P, STATIC_ACCUM_MAX_POINTS);
    return B200GS_ESIZE;
  }
  float* acc = nullptr;
  int rc = static_accum_ptr(&acc);
  if (rc) return rc;
  b200gs_surfel_bwd b = *a;                       // host-stack copy: the caller's struct is left alone
  b.accum = acc;
  b.accum_bytes = STATIC_ACCUM_FLOATS * sizeof(float);
  return b200gs_surfel_backward(&b, stream);      // same checks, same two stages; clears the rows of acc it uses
}
