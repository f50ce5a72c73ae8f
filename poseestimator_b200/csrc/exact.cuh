// The reference's fp64 edge verdict, shared by graph.cu (tile kernels, fix-up) and tail.cu (fused fix-up):
//     a = sqrt((dx*dx + dy*dy) + dz*dz), b likewise, edge <=> |a - b| <= beta, every operation rounded separately (no FMA).
#pragma once
#include <cuda_runtime.h>

namespace b2r {

__device__ __forceinline__ double sq3_ref(double dx, double dy, double dz) {
  // (dx*dx + dy*dy) + dz*dz, each operation rounded separately (reference order, no FMA)
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

__device__ __forceinline__ bool verdict_exact(double A, double B, double beta) {
  const double a = __dsqrt_rn(A), b = __dsqrt_rn(B);
  return fabs(__dsub_rn(a, b)) <= beta;
}

// Exact decision of one listed pair (i, j) of the tensor-core filter.  The filter wrote the bits of every listed pair as 0 and did
// not count them, so the fix-up only SETS bits and increments degrees: reductions without a return value, nothing to wait for.
// A pair of two different 128-row tiles is listed once (by the tile that owns row i) and owns both mirrored bits; inside a
// diagonal tile both (i, j) and (j, i) are listed and each entry owns its own row's bit.
__device__ __forceinline__ void fix_listed_pair(const double* __restrict__ src, const double* __restrict__ dst, uint32_t* adj, int pitch, int* deg,
                                                uint32_t e, double beta) {
  const size_t i = e >> 16, j = e & 0xffffu;
  const double A = sq3_ref(__dsub_rn(src[3 * j], src[3 * i]), __dsub_rn(src[3 * j + 1], src[3 * i + 1]), __dsub_rn(src[3 * j + 2], src[3 * i + 2]));
  const double B = sq3_ref(__dsub_rn(dst[3 * j], dst[3 * i]), __dsub_rn(dst[3 * j + 1], dst[3 * i + 1]), __dsub_rn(dst[3 * j + 2], dst[3 * i + 2]));
  if (!verdict_exact(A, B, beta)) return;
  atomicOr(adj + i * pitch + (j >> 5), 1u << (j & 31));
  atomicAdd(&deg[i], 1);
  if ((i >> 7) != (j >> 7)) {
    atomicOr(adj + j * pitch + (i >> 5), 1u << (i & 31));
    atomicAdd(&deg[j], 1);
  }
}

}  // namespace b2r
