// radix_sort.cuh -- stable LSD radix sort of (u64 key, u32 value) pairs on key
// bits [0, end_bit); replaces cub::DeviceRadixSort::SortPairs at
// S/cuda_rasterizer/rasterizer_impl.cu:308-313.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

namespace b200gs {
size_t radix_sort_temp_bytes(size_t n);
int radix_sort_pairs(const uint64_t* keys_in, uint64_t* keys_out, const uint32_t* vals_in, uint32_t* vals_out, int n,
                     int end_bit, void* temp, size_t temp_bytes, cudaStream_t s, bool debug);
}  // namespace b200gs
