// gauss_math.cuh -- 3x3 helpers of the 3DGS path with glm's storage (column-major) and glm's
// expression order for products, so fp-contraction matches the reference build.
#pragma once
#include "common.cuh"

namespace b200gs {

struct M3 { float c[3][3]; };  // column-major like glm: c[col][row]

// glm::operator*(mat3, mat3) (glm/detail/type_mat3x3.inl): same expression order.
__device__ __forceinline__ M3 mul(const M3& A, const M3& B) {
  M3 R;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) R.c[c][r] = A.c[0][r] * B.c[c][0] + A.c[1][r] * B.c[c][1] + A.c[2][r] * B.c[c][2];
  return R;
}
__device__ __forceinline__ M3 transpose(const M3& A) {
  M3 R;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) R.c[c][r] = A.c[r][c];
  return R;
}

}  // namespace b200gs
