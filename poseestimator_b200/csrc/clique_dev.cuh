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

// |row_q & mask| for up to R adjacency rows at once, one warp: every lane issues its 16-byte loads of ALL rows before the first
// popcount, so a sweep costs about one memory round trip per 32 x 16 bytes of row instead of one per word.  rowp[q] may point to
// global memory or to a row staged in shared memory (generic loads), nullptr = no row.  W4 = words per row / 4.
template <int R>
__device__ __forceinline__ void rows_and_count(const uint4* const (&rowp)[R], const uint32_t* mask, int W4, int lane, int (&cnt)[R]) {
#pragma unroll
  for (int q = 0; q < R; ++q) cnt[q] = 0;
  for (int w = lane; w < W4; w += 32) {
    const uint4 m = reinterpret_cast<const uint4*>(mask)[w];
    uint4 x[R];
#pragma unroll
    for (int q = 0; q < R; ++q) x[q] = rowp[q] ? rowp[q][w] : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
    for (int q = 0; q < R; ++q) cnt[q] += __popc(x[q].x & m.x) + __popc(x[q].y & m.y) + __popc(x[q].z & m.z) + __popc(x[q].w & m.w);
  }
#pragma unroll
  for (int q = 0; q < R; ++q) cnt[q] = __reduce_add_sync(0xffffffffu, cnt[q]);
}

// Adjacency rows staged in shared memory (fused tail): the first round of the seed heuristic needs the rows of all |P| members,
// later rounds and the core peel need subsets of them.  One cp.async.bulk per row brings them in with a single memory round trip
// (the 2 x 4 dependent sweeps this replaces cost ~2.7 k cycles each out of DRAM); afterwards a row is found from its vertex id
// through the bitset of requested vertices and its per-word prefix popcount.
struct RowCache {
  uint32_t* rows;        // cap rows of W words, 16-byte aligned
  uint32_t* P1;          // W words: vertices whose rows were requested (ascending: the first n_cached of them are resident)
  unsigned short* pre;   // per word of P1: set bits below it
  uint64_t* mbar;
  int cap, n_cached, W;
};
__device__ __forceinline__ const uint4* cached_row(const RowCache& rc, int id) {
  if (rc.n_cached == 0) return nullptr;
  const uint32_t w = rc.P1[id >> 5], b = 1u << (id & 31);
  if (!(w & b)) return nullptr;
  const int slot = (int)rc.pre[id >> 5] + __popc(w & (b - 1u));
  return slot < rc.n_cached ? reinterpret_cast<const uint4*>(rc.rows + (size_t)slot * rc.W) : nullptr;
}
__device__ __forceinline__ void rc_mbar_init(uint64_t* bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void rc_mbar_expect(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void rc_mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(bar);
  uint32_t done = 0;
  while (!done) {
    asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p;}" : "=r"(done) : "r"(a), "r"(parity) : "memory");
  }
}
__device__ __forceinline__ void rc_bulk_row(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"((uint32_t)__cvta_generic_to_shared(dst)),
               "l"(src), "r"(bytes), "r"((uint32_t)__cvta_generic_to_shared(bar))
               : "memory");
}

// Greedy clique from seed v inside P = N(v), one CTA of NT threads (the rule of the PMC_EXACT path; restated in
// oracle/teaser_oracle.cpp greedy_from_seed).  Every round evaluates degP(w) = |N(w) & P| for ALL members w at once -- one warp per
// member, four independent row loads in flight -- and then
//   * takes every member adjacent to all others (degP == |P| - 1) in one go, else
//   * drops every member adjacent to fewer than half of the others (bulk peel: 16-26 rounds -> 3 on config 3), else
//   * takes the member of largest degP (ties: lowest id) and intersects P with its row.
// Returns the clique size, or 0 when the seed cannot beat *lb_now (another seed's result).  kCoherent: adjacency rows are read
// with ld.global.cg (the fused tail modified them with atomics earlier in the same kernel), else through the read-only path.
// kCache: the rows of the first round's members are staged in shared memory (RowCache) and every later row read looks there first.
constexpr int kHeuThreadPerMemberW4 = 64;   // rows of up to 256 words: thread per member (heuristic_seed_cta)

template <int NT, bool kCoherent, bool kCache = false>
__device__ int heuristic_seed_cta(const ProbDesc& d, int v, uint32_t* sh, const int* lb_now, RowCache* rc = nullptr, TailMarks tm = TailMarks()) {
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
  int round = 0;
  tm.mark(8);    // seed row in shared memory
  for (;; ++round) {
    // ascending member list + count
    const bool fill = kCache && rc->n_cached == 0 && rc->cap > 0;   // uniform: n_cached changes between barriers only
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
      if (kCache && fill && w < Wb) { rc->P1[w] = word; rc->pre[w] = (unsigned short)min(pos, 65535); }
      uint32_t bits = word;
      while (bits) { list[pos++] = (unsigned short)(32 * w + (__ffs(bits) - 1)); bits &= bits - 1; }
      __syncthreads();
      if (threadIdx.x == 0) *s_base += scan[32];
      __syncthreads();
    }
    const int cnt = *s_base;
    if (round < 3) tm.mark(9 + 4 * round);    // member list
    if (cnt == 0) break;
    if (threadIdx.x == 0) { *s_lb = *(volatile const int*)lb_now; *s_best = 0u; *s_nU = 0; *s_nR = 0; }
    for (int w = threadIdx.x; w < W; w += NT) { U[w] = 0u; Rm[w] = 0u; }
    __syncthreads();
    if (size + cnt <= *s_lb) return 0;  // another seed already reached this size: cannot improve
    if (kCache && fill) {
      // one bulk copy per member row, all in flight together; the first min(cnt, cap) members of the ascending list become resident
      const int nc = min(cnt, rc->cap);
      const uint32_t row_bytes = (uint32_t)W * 4u;
      if (threadIdx.x == 0) rc_mbar_expect(rc->mbar, (uint32_t)nc * row_bytes);
      asm volatile("fence.proxy.async;" ::: "memory");   // the rows were modified through the generic proxy (fix-up atomics) earlier in this kernel
      for (int m = threadIdx.x; m < nc; m += NT) rc_bulk_row(rc->rows + (size_t)m * W, d.adj + (size_t)list[m] * d.pitch, row_bytes, rc->mbar);
      rc_mbar_wait(rc->mbar, 0);
      __syncthreads();
      if (threadIdx.x == 0) rc->n_cached = nc;
      __syncthreads();
    }
    if (round < 3) tm.mark(10 + 4 * round);   // rows staged
    if (kCache && tm.clk && threadIdx.x == 0 && round == 0) { tm.clk[30] = (unsigned)rc->n_cached; tm.clk[31] = (unsigned)rc->cap; }
    unsigned wbest = 0;
    const int W4 = W >> 2;
    if (W4 <= kHeuThreadPerMemberW4) {
      // Short rows (n <= 8192): ONE THREAD per member walks its row in 16-byte steps.  A warp per member spends ~30 instructions per
      // member on a row of one or two loads (lookup, shuffle reduction, lane-0 bookkeeping) and leaves 24 of 32 lanes idle at
      // n <= 1024: 18 k cycles per round for 560 members (config 4), now 3 k.  Lane l starts at 16-byte column l mod W4, so the eight
      // lanes of a shared-memory phase read different bank groups of the staged rows.  (Sending only the staged rows through this
      // pass and the others through the warp pass was measured: slower than either form alone.)
      const uint4* P4 = reinterpret_cast<const uint4*>(P);
      for (int m0 = 0; m0 < cnt; m0 += NT) {
        const int m = m0 + threadIdx.x;
        if (m < cnt) {
          const int id = (int)list[m];
          const uint4* rp = nullptr;
          if (kCache) rp = cached_row(*rc, id);
          if (!rp) rp = reinterpret_cast<const uint4*>(d.adj + (size_t)id * d.pitch);
          int ww = lane % W4;
          int cq = 0;
          for (int k = 0; k < W4; ++k) {
            const uint4 x = rp[ww], mk = P4[ww];
            cq += __popc(x.x & mk.x) + __popc(x.y & mk.y) + __popc(x.z & mk.z) + __popc(x.w & mk.w);
            ww = ww + 1 == W4 ? 0 : ww + 1;
          }
          wbest = max(wbest, ((unsigned)cq << 16) | (0xffffu - (unsigned)id));
          if (cq == cnt - 1) atomicOr(&U[id >> 5], 1u << (id & 31));
          if (2 * cq < cnt - 1) atomicOr(&Rm[id >> 5], 1u << (id & 31));   // adjacent to less than half of the others
        }
      }
      wbest = __reduce_max_sync(kFull, wbest);
      if (round == 0) { tm.mark(26); tm.mark(27); }
    } else {
      constexpr int R = 8;
      for (int m0 = warp; m0 < cnt; m0 += R * NW) {
        int c[R], ids[R];
        const uint4* rowp[R];
#pragma unroll
        for (int q = 0; q < R; ++q) {
          const int m = m0 + q * NW;
          ids[q] = m < cnt ? (int)list[m] : -1;
          rowp[q] = nullptr;
          if (ids[q] >= 0) {
            if (kCache) rowp[q] = cached_row(*rc, ids[q]);
            if (!rowp[q]) rowp[q] = reinterpret_cast<const uint4*>(d.adj + (size_t)ids[q] * d.pitch);
          }
        }
        if (round == 0 && m0 == warp) tm.mark(26);
        rows_and_count<R>(rowp, P, W >> 2, lane, c);
        if (round == 0 && m0 == warp) tm.mark(27);
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
    }
    if (round == 0) tm.mark(30);
    if (lane == 0 && wbest) atomicMax(s_best, wbest);
    __syncthreads();
    if (round < 3) tm.mark(11 + 4 * round);   // degrees inside P counted
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
    if (round < 3) tm.mark(12 + 4 * round);   // P updated
  }
  return size;
}

}  // namespace b2r
