// peer_allreduce.cu -- the one collective of the path: the sum over the ranks of the flat
// [59 P] fp32 gradient arena the backward kernels accumulated a batch of views into (python/b200gs/dp.py), done by
// this library's own kernel over NVLink / NVSwitch peer memory instead of ncclAllReduce.
//
// The arena lives in symmetric memory (one identical allocation per rank, every rank's copy mapped into every
// process, plus -- on NVSwitch systems -- a MULTICAST address that names all copies at once).  Rank r owns the r-th
// 1/n slice of the arena:
//   NVLS    v = multimem.ld_reduce.add.f32(mc + i)   the switch reads element i from all n GPUs and returns the sum
//           multimem.st(mc + i, v)                   the switch writes the sum to all n GPUs
//           -> each GPU sends its slice-complement once and receives the sums once: N (n-1)/n bytes each way, the
//              reduction arithmetic happens in the switch, the SMs only move 16-B vectors;
//   P2P     (no multicast object) v = sum over peers p of load(peer_ptr[p] + i), stores to every peer: the same link
//           traffic, the adds on the SMs.
// A ring all-reduce moves 2 N (n-1)/n bytes each way per GPU, twice as much.
//
// Ordering across GPUs is the caller's: a device-side barrier over all ranks before the launch (every rank's backward
// has finished writing its arena) and one after it (every slice has been broadcast); python/b200gs/dp.py uses the
// symmetric-memory handle's signal-pad barrier on the same stream.  No reference counterpart (the reference is
// single-GPU); the generic fallback is ncclAllReduce on the same memory.
#include "common.cuh"

namespace b200gs {

constexpr int AR_THREADS = 512;
constexpr int AR_UNROLL = 4;

__device__ __forceinline__ float4 mc_ld_reduce(const float4* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st(float4* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               :: "l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// [begin, end) in float4 units: this rank's slice
__global__ void __launch_bounds__(AR_THREADS) nvls_allreduce_kernel(float4* mc, size_t begin, size_t end) {
  const size_t stride = (size_t)gridDim.x * AR_THREADS;
  size_t i = begin + (size_t)blockIdx.x * AR_THREADS + threadIdx.x;
  for (; i + (AR_UNROLL - 1) * stride < end; i += AR_UNROLL * stride) {
    float4 v[AR_UNROLL];
#pragma unroll
    for (int k = 0; k < AR_UNROLL; ++k) v[k] = mc_ld_reduce(mc + i + k * stride);   // AR_UNROLL reductions in flight
#pragma unroll
    for (int k = 0; k < AR_UNROLL; ++k) mc_st(mc + i + k * stride, v[k]);
  }
  for (; i < end; i += stride) mc_st(mc + i, mc_ld_reduce(mc + i));
  __threadfence_system();
}

constexpr int AR_MAX_WORLD = 16;
struct PeerPtrs { float4* p[AR_MAX_WORLD]; };

template <int WORLD>
__global__ void __launch_bounds__(AR_THREADS) p2p_allreduce_kernel(const PeerPtrs peers, int rank, size_t begin, size_t end) {
  const size_t stride = (size_t)gridDim.x * AR_THREADS;
  for (size_t i = begin + (size_t)blockIdx.x * AR_THREADS + threadIdx.x; i < end; i += stride) {
    float4 v[WORLD];
#pragma unroll
    for (int k = 0; k < WORLD; ++k) {                       // start with the rank after this one: spread the link load
      const int p = (rank + 1 + k) % WORLD;
      v[k] = __ldcg(peers.p[p] + i);
    }
    float4 s = v[0];
#pragma unroll
    for (int k = 1; k < WORLD; ++k) { s.x += v[k].x; s.y += v[k].y; s.z += v[k].z; s.w += v[k].w; }
#pragma unroll
    for (int k = 0; k < WORLD; ++k) {
      const int p = (rank + 1 + k) % WORLD;
      __stcg(peers.p[p] + i, s);
    }
  }
  __threadfence_system();
}

}  // namespace b200gs

using namespace b200gs;

extern "C" int b200gs_peer_allreduce_f32(void* multicast_ptr, void* const* peer_ptrs, int rank, int world, size_t n_floats,
                                         int ctas, void* stream) {
  if (world < 1 || world > AR_MAX_WORLD || rank < 0 || rank >= world) { set_error("peer_allreduce: bad rank %d / world %d", rank, world); return B200GS_EINVAL; }
  if (world == 1 || n_floats == 0) return B200GS_OK;
  if (n_floats % 4 != 0) { set_error("peer_allreduce: the buffer must hold a multiple of 4 floats (pad it)"); return B200GS_EINVAL; }
  if (!multicast_ptr && !peer_ptrs) { set_error("peer_allreduce: neither a multicast address nor peer pointers"); return B200GS_EINVAL; }
  cudaStream_t s = (cudaStream_t)stream;
  const size_t nv = n_floats / 4;
  const size_t per = (nv + world - 1) / world;
  const size_t begin = per * (size_t)rank < nv ? per * (size_t)rank : nv;
  const size_t end = begin + per < nv ? begin + per : nv;
  if (end <= begin) return B200GS_OK;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  size_t want = (end - begin + AR_THREADS - 1) / AR_THREADS;
  const size_t cap = ctas > 0 ? (size_t)ctas : (size_t)sms * 4;
  const int blocks = (int)(want < cap ? want : cap);
  if (multicast_ptr) {
    if (reinterpret_cast<uintptr_t>(multicast_ptr) & 15u) { set_error("peer_allreduce: multicast address not 16-B aligned"); return B200GS_EINVAL; }
    nvls_allreduce_kernel<<<blocks, AR_THREADS, 0, s>>>(reinterpret_cast<float4*>(multicast_ptr), begin, end);
    B200GS_LAUNCHED("nvls_allreduce_kernel", false, s);
    return B200GS_OK;
  }
  PeerPtrs pp;
  for (int i = 0; i < world; ++i) {
    if (!peer_ptrs[i] || (reinterpret_cast<uintptr_t>(peer_ptrs[i]) & 15u)) { set_error("peer_allreduce: peer pointer %d missing or not 16-B aligned", i); return B200GS_EINVAL; }
    pp.p[i] = reinterpret_cast<float4*>(peer_ptrs[i]);
  }
  switch (world) {
    case 2: p2p_allreduce_kernel<2><<<blocks, AR_THREADS, 0, s>>>(pp, rank, begin, end); break;
    case 4: p2p_allreduce_kernel<4><<<blocks, AR_THREADS, 0, s>>>(pp, rank, begin, end); break;
    case 8: p2p_allreduce_kernel<8><<<blocks, AR_THREADS, 0, s>>>(pp, rank, begin, end); break;
    default: set_error("peer_allreduce: the peer-to-peer variant is built for 2, 4 or 8 ranks (got %d)", world); return B200GS_EINVAL;
  }
  B200GS_LAUNCHED("p2p_allreduce_kernel", false, s);
  return B200GS_OK;
}
