// Device code of the clique stage shared by clique.cu (one kernel per step) and tail.cu (one fused kernel per batch).
#pragma once
#include "common.cuh"

namespace b2r {

__device__ __forceinline__ int words_for(int n) { return 4 * ((n + kTile - 1) / kTile); }

// shared memory the seed heuristic needs for a problem of n vertices (P, U, Rm bitsets, scan scratch, member list)
__host__ __device__ inline size_t heuristic_smem_bytes(int n) {
  const size_t W = 4 * (size_t)((n + kTile - 1) / kTile);
  return (3 * W + 64) * 4 + (size_t)2 * (n + 8);
}

// |row_m & mask| for up to 8 adjacency rows at once, one warp: every lane issues its 16-byte loads of ALL rows before the first
// popcount, so a sweep costs about one L2 round trip per 32 x 16 bytes of row instead of one per word (the scalar loop this
// replaces was a chain of dependent loads: 20 round trips per four rows at N = 5000).  W4 = words per row / 4.
template <int R, bool kCoherent>
__device__ __forceinline__ void rows_and_count(const uint32_t* __restrict__ adj, int pitch, const int (&ids)[R], const uint32_t* mask, int W4, int lane,
                                               int (&cnt)[R]) {
#pragma unroll
  for (int q = 0; q < R; ++q) cnt[q] = 0;
  for (int w = lane; w < W4; w += 32) {
    const uint4 m = reinterpret_cast<const uint4*>(mask)[w];
    uint4 x[R];
#pragma unroll
    for (int q = 0; q < R; ++q) {
      x[q] = make_uint4(0u, 0u, 0u, 0u);
      if (ids[q] >= 0) {
        const uint4* row = reinterpret_cast<const uint4*>(adj + (size_t)ids[q] * pitch) + w;
        x[q] = kCoherent ? __ldcg(row) : __ldg(row);
      }
    }
#pragma unroll
    for (int q = 0; q < R; ++q) cnt[q] += __popc(x[q].x & m.x) + __popc(x[q].y & m.y) + __popc(x[q].z & m.z) + __popc(x[q].w & m.w);
  }
#pragma unroll
  for (int q = 0; q < R; ++q) cnt[q] = __reduce_add_sync(0xffffffffu, cnt[q]);
}

// Greedy clique from seed v inside P = N(v), one CTA of NT threads (the rule of the PMC_EXACT path; restated in
// oracle/teaser_oracle.cpp greedy_from_seed).  Every round evaluates degP(w) = |N(w) & P| for ALL members w at once -- one warp per
// member, four independent row loads in flight -- and then
//   * takes every member adjacent to all others (degP == |P| - 1) in one go, else
//   * drops every member adjacent to fewer than half of the others (bulk peel: 16-26 rounds -> 3 on config 3), else
//   * takes the member of largest degP (ties: lowest id) and intersects P with its row.
// Returns the clique size, or 0 when the seed cannot beat *lb_now (another seed's result).  kCoherent: adjacency rows are read
// with ld.global.cg (the fused tail modified them with atomics earlier in the same kernel), else through the read-only path.
template <int NT, bool kCoherent>
__device__ int heuristic_seed_cta(const ProbDesc& d, int v, uint32_t* sh, const int* lb_now) {
  constexpr unsigned kFull = 0xffffffffu;
  constexpr int NW = NT / 32;
  const int n = d.n;
  const int W = words_for(n);
  const int Wb = (n + 31) >> 5;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t* P = sh;                       // W
  uint32_t* U = sh + W;                   // W
  uint32_t* Rm = sh + 2 * W;              // W: members to drop this round
  int* scan = (int*)(sh + 3 * W);         // 40
  unsigned* s_best = (unsigned*)(scan + 40);
  int* s_nU = (int*)(scan + 41);
  int* s_lb = (int*)(scan + 42);
  int* s_nR = (int*)(scan + 43);
  int* s_base = (int*)(scan + 44);
  unsigned short* list = (unsigned short*)(scan + 48);  // n entries
  auto ld = [](const uint32_t* p) -> uint32_t { return kCoherent ? __ldcg(p) : __ldg(p); };
  const uint32_t* row0 = d.adj + (size_t)v * d.pitch;
  __syncthreads();   // the caller may still be reading sh
  for (int w = threadIdx.x; w < W; w += NT) P[w] = ld(row0 + w);
  __syncthreads();
  int size = 1;
  for (;;) {
    // ascending member list + count
    if (threadIdx.x == 0) *s_base = 0;
    __syncthreads();
    for (int w0 = 0; w0 < Wb; w0 += NT) {
      const int w = w0 + threadIdx.x;
      const uint32_t word = (w < Wb) ? P[w] : 0u;
      const int c = __popc(word);
      int incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(kFull, incl, o); if (lane >= o) incl += t; }
      if (lane == 31) scan[warp] = incl;
      __syncthreads();
      if (warp == 0) {
        int x = lane < NW ? scan[lane] : 0, xi = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(kFull, xi, o); if (lane >= o) xi += t; }
        scan[lane] = xi - x;
        if (lane == 31) scan[32] = xi;
      }
      __syncthreads();
      int pos = *s_base + scan[warp] + incl - c;
      uint32_t bits = word;
      while (bits) { list[pos++] = (unsigned short)(32 * w + (__ffs(bits) - 1)); bits &= bits - 1; }
      __syncthreads();
      if (threadIdx.x == 0) *s_base += scan[32];
      __syncthreads();
    }
    const int cnt = *s_base;
    if (cnt == 0) break;
    if (threadIdx.x == 0) { *s_lb = *(volatile const int*)lb_now; *s_best = 0u; *s_nU = 0; *s_nR = 0; }
    for (int w = threadIdx.x; w < W; w += NT) { U[w] = 0u; Rm[w] = 0u; }
    __syncthreads();
    if (size + cnt <= *s_lb) return 0;  // another seed already reached this size: cannot improve
    unsigned wbest = 0;
    constexpr int R = 8;
    for (int m0 = warp; m0 < cnt; m0 += R * NW) {
      int c[R], ids[R];
#pragma unroll
      for (int q = 0; q < R; ++q) {
        const int m = m0 + q * NW;
        ids[q] = m < cnt ? (int)list[m] : -1;
      }
      rows_and_count<R, kCoherent>(d.adj, d.pitch, ids, P, W >> 2, lane, c);
#pragma unroll
      for (int q = 0; q < R; ++q) {
        if (ids[q] < 0) continue;
        const int cq = c[q];
        const unsigned id = (unsigned)ids[q];
        wbest = max(wbest, ((unsigned)cq << 16) | (0xffffu - id));
        if (cq == cnt - 1 && lane == 0) atomicOr(&U[id >> 5], 1u << (id & 31));
        if (2 * cq < cnt - 1 && lane == 0) atomicOr(&Rm[id >> 5], 1u << (id & 31));   // adjacent to less than half of the others
      }
    }
    if (lane == 0 && wbest) atomicMax(s_best, wbest);
    __syncthreads();
    int nu = 0, nr = 0;
    for (int w = threadIdx.x; w < Wb; w += NT) { nu += __popc(U[w]); nr += __popc(Rm[w]); }
    nu = __reduce_add_sync(kFull, nu); nr = __reduce_add_sync(kFull, nr);
    if (lane == 0 && nu) atomicAdd(s_nU, nu);
    if (lane == 0 && nr) atomicAdd(s_nR, nr);
    __syncthreads();
    const int nU = *s_nU, nR = *s_nR;
    if (nU > 0) {
      size += nU;
      for (int w = threadIdx.x; w < W; w += NT) P[w] &= ~U[w];
    } else if (nR > 0 && nR < cnt) {
      for (int w = threadIdx.x; w < W; w += NT) P[w] &= ~Rm[w];
    } else {
      const unsigned u = 0xffffu - (*s_best & 0xffffu);
      size += 1;
      const uint32_t* row = d.adj + (size_t)u * d.pitch;
      for (int w = threadIdx.x; w < W; w += NT) P[w] &= ld(row + w);
    }
    __syncthreads();
  }
  return size;
}

}  // namespace b2r
