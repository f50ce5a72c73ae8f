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
struct FixPair { double s[6], t[6]; };   // src_i, src_j, dst_i, dst_j
__device__ __forceinline__ FixPair load_listed_pair(const double* __restrict__ src, const double* __restrict__ dst, uint32_t e) {
  const size_t i = e >> 16, j = e & 0xffffu;
  FixPair p;
#pragma unroll
  for (int k = 0; k < 3; ++k) { p.s[k] = src[3 * i + k]; p.s[3 + k] = src[3 * j + k]; p.t[k] = dst[3 * i + k]; p.t[3 + k] = dst[3 * j + k]; }
  return p;
}
__device__ __forceinline__ void decide_listed_pair(const FixPair& p, uint32_t* adj, int pitch, int* deg, uint32_t e, double beta) {
  const size_t i = e >> 16, j = e & 0xffffu;
  const double A = sq3_ref(__dsub_rn(p.s[3], p.s[0]), __dsub_rn(p.s[4], p.s[1]), __dsub_rn(p.s[5], p.s[2]));
  const double B = sq3_ref(__dsub_rn(p.t[3], p.t[0]), __dsub_rn(p.t[4], p.t[1]), __dsub_rn(p.t[5], p.t[2]));
  if (!verdict_exact(A, B, beta)) return;
  atomicOr(adj + i * pitch + (j >> 5), 1u << (j & 31));
  atomicAdd(&deg[i], 1);
  if ((i >> 7) != (j >> 7)) {
    atomicOr(adj + j * pitch + (i >> 5), 1u << (i & 31));
    atomicAdd(&deg[j], 1);
  }
}
__device__ __forceinline__ void fix_listed_pair(const double* __restrict__ src, const double* __restrict__ dst, uint32_t* adj, int pitch, int* deg,
                                                uint32_t e, double beta) {
  decide_listed_pair(load_listed_pair(src, dst, e), adj, pitch, deg, e, beta);
}

}  // namespace b2r
