// Single-pass ordered stream compaction helpers shared by presteps.cu and raster.cu.
#pragma once
#include <cuda_runtime.h>

namespace b2r {
namespace {

#define FULL 0xffffffffu

// ------------------------------------------------------------------------------------------------------------------
// Single-pass ordered compaction: decoupled look-back over tiles taken in ticket order.
// status word: bits 63..62 = 0 empty, 1 tile aggregate, 2 inclusive prefix; low 62 bits = value.  ws: [0] ticket,
// [1..] status words; zeroed before the launch.
// ------------------------------------------------------------------------------------------------------------------
constexpr unsigned long long kLbValue = (1ull << 62) - 1;

__device__ __forceinline__ int take_ticket(unsigned long long* ws, int* s_tile) {
  if (threadIdx.x == 0) *s_tile = (int)atomicAdd(&ws[0], 1ull);
  __syncthreads();
  return *s_tile;
}

// thread 0 only: publish this tile's aggregate, return the exclusive prefix of the tile.  (A warp-wide look-back, 32 predecessors per
// step, was measured in round 2: the 4096 x 4096 back-projection went from 0.286 to 0.329 ms -- the nearest predecessor usually holds
// its inclusive prefix already, so the single-thread walk reads one or two words where the warp reads 32.)
__device__ unsigned long long lookback(unsigned long long* ws, int tile, unsigned long long aggregate) {
  volatile unsigned long long* status = ws + 1;
  if (tile == 0) { status[0] = (2ull << 62) | aggregate; return 0; }
  status[tile] = (1ull << 62) | aggregate;
  unsigned long long excl = 0;
  for (int p = tile - 1; p >= 0; --p) {
    unsigned long long s;
    do { s = status[p]; } while ((s >> 62) == 0);
    excl += s & kLbValue;
    if ((s >> 62) == 2) break;
  }
  status[tile] = (2ull << 62) | (excl + aggregate);
  return excl;
}

// block-wide exclusive scan of one count per thread (256 threads); returns the thread's offset, total in *total
__device__ __forceinline__ unsigned block_scan_256(unsigned c, unsigned* s_warp /*8*/, unsigned* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned x = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned y = __shfl_up_sync(FULL, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) s_warp[warp] = x;
  __syncthreads();
  unsigned wb = 0, t = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) { const unsigned s = s_warp[w]; if (w < warp) wb += s; t += s; }
  *total = t;
  return wb + x - c;
}


}  // namespace
}  // namespace b2r
