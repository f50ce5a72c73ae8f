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

// Exact decision of one listed pair (i, j) of the tensor-core filter: both symmetric bits are set or cleared atomically and
// every bit that actually flips corrects its row's degree (the filter counted its placeholder bits).
__device__ __forceinline__ void fix_listed_pair(const double* __restrict__ src, const double* __restrict__ dst, uint32_t* adj, int pitch, int* deg,
                                                uint32_t e, double beta) {
  const size_t i = e >> 16, j = e & 0xffffu;
  const double A = sq3_ref(__dsub_rn(src[3 * j], src[3 * i]), __dsub_rn(src[3 * j + 1], src[3 * i + 1]), __dsub_rn(src[3 * j + 2], src[3 * i + 2]));
  const double B = sq3_ref(__dsub_rn(dst[3 * j], dst[3 * i]), __dsub_rn(dst[3 * j + 1], dst[3 * i + 1]), __dsub_rn(dst[3 * j + 2], dst[3 * i + 2]));
  uint32_t* wij = adj + i * pitch + (j >> 5);
  uint32_t* wji = adj + j * pitch + (i >> 5);
  const uint32_t bj_ = 1u << (j & 31), bi_ = 1u << (i & 31);
  if (verdict_exact(A, B, beta)) {
    if (!(atomicOr(wij, bj_) & bj_)) atomicAdd(&deg[i], 1);
    if (!(atomicOr(wji, bi_) & bi_)) atomicAdd(&deg[j], 1);
  } else {
    if (atomicAnd(wij, ~bj_) & bj_) atomicSub(&deg[i], 1);
    if (atomicAnd(wji, ~bi_) & bi_) atomicSub(&deg[j], 1);
  }
}

}  // namespace b2r
