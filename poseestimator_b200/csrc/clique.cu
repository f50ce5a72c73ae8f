// Maximum-clique inlier selection on the bit-packed consistency graph.
//
// Replaces upstream's MaxCliqueSolver::findMaxClique (PMC: k-core -> heuristic -> exact branch and bound),
// reached from `solver.solve` (HelpersRealtime.py:161,166: PMC_EXACT).  GPU formulation:
//   1. heuristic_kernel   one warp per seed vertex grows a clique greedily inside N(seed) (highest degree first);
//                         the neighbourhood is a register bitset spread over the warp, "P &= N(u)" is one
//                         coalesced row load, the candidate pick one REDUX.  -> lower bound lb
//   2. prune_kernel       (lb-1)-core peeling with a shared-memory alive bitset (ballot / popc), then the
//                         surviving vertices are numbered in ascending original id.
//   3. compact_kernel     gathers the surviving sub-matrix into a small dense bit matrix (ballot per word).
//   4. bnb_kernel         exact bit-parallel branch and bound (greedy colouring bound, BBMC style): one warp per
//                         root subtree, bitsets in registers, AND + popc + REDUX, incumbent shared by atomicMax.
//   5. canon_kernel       deterministic answer: the maximum clique whose sorted id list is lexicographically
//                         smallest (upstream returns "a" maximum clique depending on thread timing, A.5-8).
// All sizes stay on the device between the stages; the host never synchronises inside a solve.
#include <stdio.h>
#include "common.cuh"
#include "clique_dev.cuh"

namespace b2r {

#define FULL 0xffffffffu
constexpr unsigned kNone = 0xffffffffu;

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}


// ---------------------------------------------------------------------------------------------------------
// warp-distributed bitsets: word w lives in lane (w & 31), slot (w >> 5)
// ---------------------------------------------------------------------------------------------------------
template <int WPL>
__device__ __forceinline__ int bs_count(const uint32_t (&X)[WPL]) {
  int c = 0;
#pragma unroll
  for (int s = 0; s < WPL; ++s) c += __popc(X[s]);
  return __reduce_add_sync(FULL, c);
}

template <int WPL>
__device__ __forceinline__ unsigned bs_first(const uint32_t (&X)[WPL], int lane) {
  unsigned id = kNone;
#pragma unroll
  for (int s = WPL - 1; s >= 0; --s)
    if (X[s]) id = ((unsigned)(lane + 32 * s) << 5) + (__ffs(X[s]) - 1);
  return __reduce_min_sync(FULL, id);
}

template <int WPL>
__device__ __forceinline__ void bs_clear(uint32_t (&X)[WPL], unsigned v, int lane) {
  if (lane == (int)((v >> 5) & 31)) {
#pragma unroll
    for (int s = 0; s < WPL; ++s)
      if (s == (int)(v >> 10)) X[s] &= ~(1u << (v & 31));
  }
}

template <int WPL>
__device__ __forceinline__ void bs_and_row(uint32_t (&X)[WPL], const uint32_t* __restrict__ row, int W, int lane) {
#pragma unroll
  for (int s = 0; s < WPL; ++s) {
    const int w = lane + 32 * s;
    X[s] = (w < W) ? (X[s] & row[w]) : 0u;
  }
}

template <int WPL>
__device__ __forceinline__ void bs_andnot_row(uint32_t (&X)[WPL], const uint32_t* __restrict__ row, int W, int lane) {
#pragma unroll
  for (int s = 0; s < WPL; ++s) {
    const int w = lane + 32 * s;
    if (w < W) X[s] &= ~row[w];
  }
}

// keep only ids > v
template <int WPL>
__device__ __forceinline__ void bs_keep_above(uint32_t (&X)[WPL], unsigned v, int lane) {
  const int wv = (int)(v >> 5);
#pragma unroll
  for (int s = 0; s < WPL; ++s) {
    const int w = lane + 32 * s;
    if (w < wv) X[s] = 0u;
    else if (w == wv) X[s] &= ~((2u << (v & 31)) - 1u);
  }
}

// ---------------------------------------------------------------------------------------------------------
// 1. greedy heuristic
// ---------------------------------------------------------------------------------------------------------
// One warp grows a clique from `seed` inside P = N(seed).  Every round it evaluates, for ALL members w of P at once
// (independent row loads -> memory-level parallelism instead of a dependent chain), degP(w) = |N(w) & P|:
//   * members adjacent to every other member (degP == |P|-1) are all taken in one go;
//   * otherwise the member with the largest degP (ties: lowest id) is taken and P &= N(u).
// A dominant clique is therefore swallowed in 2-4 rounds instead of |clique| dependent steps.
// The same rule is restated in oracle/teaser_oracle.cpp (greedy_from_seed) for the PMC_HEU / KCORE_HEU modes.
template <int WPL>
__device__ __forceinline__ int greedy_grow(const ProbDesc& d, int W, int v, int lane, bool strict, uint32_t* sP,
                                           uint32_t (&Cb)[WPL], bool record) {
  constexpr int UN = WPL <= 8 ? 4 : (WPL <= 16 ? 2 : 1);
  uint32_t P[WPL];
  const uint32_t* row0 = d.adj + (size_t)v * d.pitch;
#pragma unroll
  for (int s = 0; s < WPL; ++s) {
    const int w = lane + 32 * s;
    P[s] = (w < W) ? row0[w] : 0u;
    if (w < W) sP[w] = P[s];
    if (record) Cb[s] = 0u;
  }
  if (record && lane == ((v >> 5) & 31)) {
#pragma unroll
    for (int s = 0; s < WPL; ++s)
      if (s == (v >> 10)) Cb[s] |= 1u << (v & 31);
  }
  __syncwarp();
  int size = 1;
  int cnt = bs_count<WPL>(P);
  while (cnt > 0) {
    const int lbnow = *(volatile int*)&d.st->lb;
    if (strict ? (size + cnt <= lbnow) : (size + cnt < lbnow)) return 0;  // cannot improve (strict) / tie (non-strict)
    unsigned best = 0;
    uint32_t U[WPL];
#pragma unroll
    for (int s = 0; s < WPL; ++s) U[s] = 0u;
    int wi = 0;
    uint32_t word = sP[0];
    for (;;) {
      int ids[UN];
#pragma unroll
      for (int q = 0; q < UN; ++q) {
        while (word == 0u && wi + 1 < W) { ++wi; word = sP[wi]; }
        if (word) { ids[q] = 32 * wi + (__ffs(word) - 1); word &= word - 1; }
        else ids[q] = -1;
      }
      if (ids[0] < 0) break;
      int c[UN];
#pragma unroll
      for (int q = 0; q < UN; ++q) {
        c[q] = 0;
        if (ids[q] >= 0) {
          const uint32_t* row = d.adj + (size_t)ids[q] * d.pitch;
#pragma unroll
          for (int s = 0; s < WPL; ++s) {
            const int w = lane + 32 * s;
            if (w < W) c[q] += __popc(P[s] & __ldg(row + w));
          }
        }
      }
#pragma unroll
      for (int q = 0; q < UN; ++q) {
        if (ids[q] < 0) continue;
        const int cq = __reduce_add_sync(FULL, c[q]);
        const unsigned id = (unsigned)ids[q];
        best = max(best, ((unsigned)cq << 16) | (0xffffu - id));
        if (cq == cnt - 1 && lane == (int)((id >> 5) & 31)) {
#pragma unroll
          for (int s = 0; s < WPL; ++s)
            if (s == (int)(id >> 10)) U[s] |= 1u << (id & 31);
        }
      }
    }
    const int nU = bs_count<WPL>(U);
    if (nU > 0) {
      size += nU;
#pragma unroll
      for (int s = 0; s < WPL; ++s) { P[s] &= ~U[s]; if (record) Cb[s] |= U[s]; }
    } else {
      const unsigned u = 0xffffu - (best & 0xffffu);
      ++size;
      if (record && lane == (int)((u >> 5) & 31)) {
#pragma unroll
        for (int s = 0; s < WPL; ++s)
          if (s == (int)(u >> 10)) Cb[s] |= 1u << (u & 31);
      }
      bs_and_row<WPL>(P, d.adj + (size_t)u * d.pitch, W, lane);
    }
    cnt = bs_count<WPL>(P);
    __syncwarp();
#pragma unroll
    for (int s = 0; s < WPL; ++s) {
      const int w = lane + 32 * s;
      if (w < W) sP[w] = P[s];
    }
    __syncwarp();
  }
  return size;
}

// all_seeds == 0 (PMC_EXACT): 8 warps per problem, warp k starts from the highest-degree vertex of id class k
//                             (picked by degree_kernel); the bound only has to be good, the exact search proves it.
// all_seeds == 1 (PMC_HEU / KCORE_HEU): every vertex is a seed; winner = largest clique, ties -> lowest seed id.
template <int WPL>
__global__ void __launch_bounds__(256) heuristic_kernel(const ProbDesc* __restrict__ descs, int all_seeds) {
  extern __shared__ uint32_t sh[];
  const ProbDesc d = descs[blockIdx.y];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int W = words_for(d.n);
  int v;
  if (all_seeds) {
    v = blockIdx.x * 8 + warp;
    if (v >= d.n) return;
    if (d.deg[v] + 1 < *(volatile int*)&d.st->lb) return;
  } else {
    const unsigned long long key = d.st->seed_key[warp];
    if (key == 0ull) return;
    v = (int)(0xffffffffu - (unsigned)(key & 0xffffffffu));
  }
  uint32_t dummy[WPL];
  const int size = greedy_grow<WPL>(d, W, v, lane, !all_seeds, sh + (size_t)warp * W, dummy, false);
  if (size > 0 && lane == 0) {
    atomicMax(&d.st->heu_key, ((unsigned long long)size << 32) | (unsigned long long)(0xffffffffu - (unsigned)v));
    atomicMax(&d.st->lb, size);
  }
}

// CTA-per-seed variant of the same greedy rule (PMC_EXACT path): the degP evaluation of one round is spread over 32
// warps with 4 independent row loads in flight per warp, so a round costs a few L2 round trips instead of |P| of them.
// grid = (seeds per problem, batch); seed k = highest-degree vertex of id class k (degree_kernel).
constexpr int kHeuThreads = 1024;
// Seeds of the 8 id classes that get a CTA.  The bound only has to be good, the exact search proves it; every seed CTA (1024 threads)
// takes the place of one resident graph CTA of another batch in flight: 4 seeds 100.2k reg/s, 2 seeds 102.6k, 1 seed 104.2k (config 3).
#ifndef B2R_HEU_SEEDS
#define B2R_HEU_SEEDS 2
#endif
constexpr int kHeuSeeds = B2R_HEU_SEEDS;
__global__ void __launch_bounds__(kHeuThreads) heuristic_block_kernel(const ProbDesc* __restrict__ descs) {
  extern __shared__ uint32_t sh[];
  const ProbDesc d = descs[blockIdx.y];
  ProbState* st = d.st;
  if (st->tail_done) return;   // the fused tail (tail.cu) finished this problem
  const unsigned long long skey = st->seed_key[blockIdx.x];
  if (skey == 0ull) return;
  const int v = (int)(0xffffffffu - (unsigned)(skey & 0xffffffffu));
  const int size = heuristic_seed_cta<kHeuThreads, false>(d, v, sh, &st->lb);
  if (size > 0 && threadIdx.x == 0) {
    atomicMax(&st->heu_key, ((unsigned long long)size << 32) | (unsigned long long)(0xffffffffu - (unsigned)v));
    atomicMax(&st->lb, size);
  }
}

// Re-runs the winning seed and writes its clique (ascending ids).  PMC_HEU / KCORE_HEU fall-through only.
template <int WPL>
__global__ void __launch_bounds__(32) heuristic_replay_kernel(const ProbDesc* __restrict__ descs) {
  extern __shared__ uint32_t sh[];
  const ProbDesc d = descs[blockIdx.x];
  const int lane = threadIdx.x;
  if (d.st->clique_size >= 0) return;  // already decided (k-core shortcut)
  const int W = words_for(d.n);
  const unsigned long long key = d.st->heu_key;
  const int seed = (int)(0xffffffffu - (unsigned)(key & 0xffffffffu));
  uint32_t Cb[WPL];
  if (lane == 0) d.st->lb = 0;  // replay must not abort against the bound it produced itself
  __syncwarp();
  greedy_grow<WPL>(d, W, seed, lane, false, sh, Cb, true);
  int off = 0;
  for (int w = 0; w < W; ++w) {
    uint32_t word = 0;
#pragma unroll
    for (int s = 0; s < WPL; ++s)
      if (s == (w >> 5)) word = Cb[s];
    word = __shfl_sync(FULL, word, w & 31);
    if ((word >> lane) & 1) d.clique[off + __popc(word & ((1u << lane) - 1u))] = 32 * w + lane;
    off += __popc(word);
  }
  if (lane == 0) { d.st->clique_size = off; d.st->omega = off; d.st->lb = off; }
}

// ---------------------------------------------------------------------------------------------------------
// 2. k-core peeling on the full graph + numbering of the survivors
// ---------------------------------------------------------------------------------------------------------
// Peels `alive` (shared bitset over n vertices) to the T-core of the graph restricted to alive.
__device__ void peel_full(const ProbDesc& d, int W, uint32_t* alive, int T, int* sh_changed) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int n = d.n;
  for (;;) {
    __syncthreads();
    if (threadIdx.x == 0) *sh_changed = 0;
    __syncthreads();
    for (int v = warp; v < n; v += nwarps) {
      if (!((alive[v >> 5] >> (v & 31)) & 1)) continue;
      const uint32_t* row = d.adj + (size_t)v * d.pitch;
      int c = 0;
      for (int w = lane; w < W; w += 32) c += __popc(row[w] & alive[w]);
      c = __reduce_add_sync(FULL, c);
      if (c < T && lane == 0) { atomicAnd(&alive[v >> 5], ~(1u << (v & 31))); *sh_changed = 1; }
    }
    __syncthreads();
    if (!*sh_changed) break;
  }
}

// exclusive prefix of popcounts over the alive words -> ids[] (ascending original id); returns count
__device__ int number_alive(const uint32_t* alive, int Wb, int* ids, int* sh_scan /* >= 33 ints */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  __shared__ int sh_base;
  if (threadIdx.x == 0) sh_base = 0;
  __syncthreads();
  // chunks of blockDim.x words
  for (int w0 = 0; w0 < Wb; w0 += blockDim.x) {
    const int w = w0 + threadIdx.x;
    const uint32_t word = (w < Wb) ? alive[w] : 0u;
    const int c = __popc(word);
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(FULL, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) sh_scan[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int x = (lane < nwarps) ? sh_scan[lane] : 0;
      int xi = x;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(FULL, xi, o);
        if (lane >= o) xi += t;
      }
      sh_scan[lane] = xi - x;              // exclusive warp offsets
      if (lane == 31) sh_scan[32] = xi;    // chunk total
    }
    __syncthreads();
    int pos = sh_base + sh_scan[warp] + incl - c;
    uint32_t bits = word;
    while (bits) {
      const int b = __ffs(bits) - 1;
      bits &= bits - 1;
      ids[pos++] = 32 * w + b;
    }
    __syncthreads();
    if (threadIdx.x == 0) sh_base += sh_scan[32];
    __syncthreads();
  }
  return sh_base;
}

__global__ void __launch_bounds__(1024) prune_kernel(const ProbDesc* __restrict__ descs) {
  extern __shared__ uint32_t sh[];
  const ProbDesc d = descs[blockIdx.x];
  ProbState* st = d.st;
  if (st->tail_done) return;
  const int n = d.n;
  const int W = words_for(n);
  const int Wb = (n + 31) >> 5;
  uint32_t* alive = sh;                 // W words (>= Wb)
  int* sh_scan = (int*)(sh + W);        // 40 ints
  int* sh_changed = sh_scan + 36;
  const int lb = st->lb;
  const int T = max(lb - 1, 0);
  // alive0 = { deg >= T }
  for (int w = threadIdx.x; w < W; w += blockDim.x) alive[w] = 0u;
  __syncthreads();
  for (int v0 = 0; v0 < n; v0 += blockDim.x) {
    const int v = v0 + threadIdx.x;
    const bool a = v < n && d.deg[v] >= T;
    const uint32_t bal = __ballot_sync(FULL, a);
    if ((threadIdx.x & 31) == 0 && (v >> 5) < W) alive[v >> 5] = bal;
  }
  __syncthreads();
  if (T > 0) peel_full(d, W, alive, T, sh_changed);
  __syncthreads();
  const int cnt = number_alive(alive, Wb, d.ids, sh_scan);
  if (threadIdx.x == 0) {
    st->T = T;
    st->n_alive = cnt;
    const int Wc = (((cnt + 31) >> 5) + 3) & ~3;
    st->Wc = Wc;
    st->omega = lb;
    if (Wc > kMaxCompactWords) st->flags |= FLAG_COMPACT_TOO_LARGE;
  }
}

// ---------------------------------------------------------------------------------------------------------
// 3. gather the surviving sub-matrix
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) compact_kernel(const ProbDesc* __restrict__ descs) {
  const ProbDesc d = descs[blockIdx.y];
  const ProbState* st = d.st;
  if (st->tail_done) return;
  const int np = st->n_alive, Wc = st->Wc;
  if (st->flags & FLAG_COMPACT_TOO_LARGE) return;
  const int lane = threadIdx.x & 31;
  const long long tasks = (long long)np * Wc;
  const long long stride = (long long)gridDim.x * 8;
  for (long long task = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); task < tasks; task += stride) {
    const int ip = (int)(task / Wc), wj = (int)(task % Wc);
    const int jp = 32 * wj + lane;
    bool bit = false;
    if (jp < np) {
      const int oi = d.ids[ip], oj = d.ids[jp];
      bit = (d.adj[(size_t)oi * d.pitch + (oj >> 5)] >> (oj & 31)) & 1u;
    }
    const uint32_t word = __ballot_sync(FULL, bit);
    if (lane == 0) d.cadj[(size_t)ip * d.pitch + wj] = word;
  }
}

// ---------------------------------------------------------------------------------------------------------
// 4. exact search: warp-level bit-parallel DFS with greedy-colouring bounds
// ---------------------------------------------------------------------------------------------------------
struct AdjView {
  const uint32_t* base;
  int pitch;
  __device__ __forceinline__ const uint32_t* row(unsigned v) const { return base + (size_t)v * pitch; }
};

// Colours P greedily (colour classes = independent sets taken in id order); writes the candidates whose colour
// is >= kmin as (colour << 16 | vertex), in non-decreasing colour order.  Returns their number, or -1 on overflow.
template <int WPL>
__device__ int colour_candidates(const uint32_t (&P)[WPL], const AdjView& A, int Wc, int kmin, uint32_t* cand, int cap,
                                 int lane) {
  uint32_t Q[WPL];
#pragma unroll
  for (int s = 0; s < WPL; ++s) Q[s] = P[s];
  int k = 0, nc = 0;
  for (;;) {
    unsigned v = bs_first<WPL>(Q, lane);
    if (v == kNone) break;
    ++k;
    uint32_t S[WPL];
#pragma unroll
    for (int s = 0; s < WPL; ++s) S[s] = Q[s];
    while (v != kNone) {
      bs_clear<WPL>(S, v, lane);
      bs_clear<WPL>(Q, v, lane);
      bs_andnot_row<WPL>(S, A.row(v), Wc, lane);
      if (k >= kmin) {
        if (nc >= cap) return -1;
        if (lane == 0) cand[nc] = ((unsigned)k << 16) | v;
        ++nc;
      }
      v = bs_first<WPL>(S, lane);
    }
  }
  return nc;
}

struct DfsCtx {
  AdjView A;
  int Wc;
  uint32_t* ar;                 // this warp's arena
  size_t cap;                   // words
  int* best;                    // shared incumbent (optimisation mode)
  bool decision;                // stop at the first clique of size >= need
  int need;
  const volatile int* abort_below;  // decision mode: give up when *abort_below < my_id
  int my_id;
  ProbState* st;
  unsigned long long deadline;  // globaltimer ns, 0 = none
  unsigned long long nodes;
};

// frame: [0] prev offset  [1] ncand  [2] next candidate index  [3] |R|   [4..4+Wc) P   [..] candidates
template <int WPL>
__device__ bool dfs(const uint32_t (&P0)[WPL], int size0, DfsCtx& c, int lane) {
  const int Wc = c.Wc;
  uint32_t* ar = c.ar;
  size_t top = 0, fp = 0;
  int depth = 0;
  int h_prev = 0, h_idx = -1, h_size = 0;

  auto best_now = [&]() -> int { return c.decision ? c.need - 1 : *(volatile int*)c.best; };

  // pushes a frame for candidate set X with |R| = size; 1 pushed, 0 nothing to explore, -1 overflow
  auto push = [&](const uint32_t (&X)[WPL], int size, int cnt) -> int {
    if (fp + 4 + (size_t)Wc + (size_t)cnt > c.cap) return -1;
    int kmin = best_now() - size + 1;
    if (kmin < 1) kmin = 1;
#pragma unroll
    for (int s = 0; s < WPL; ++s) {
      const int w = lane + 32 * s;
      if (w < Wc) ar[fp + 4 + w] = X[s];
    }
    const int nc = colour_candidates<WPL>(X, c.A, Wc, kmin, ar + fp + 4 + Wc, cnt, lane);
    if (nc < 0) return -1;
    if (nc == 0) return 0;
    if (lane == 0) { ar[fp] = (uint32_t)top; ar[fp + 1] = (uint32_t)nc; ar[fp + 2] = (uint32_t)(nc - 1); ar[fp + 3] = (uint32_t)size; }
    __syncwarp();
    h_prev = (int)top; h_idx = nc - 1; h_size = size;
    top = fp;
    fp = fp + 4 + Wc + nc;
    ++depth;
    return 1;
  };

  {
    const int cnt0 = bs_count<WPL>(P0);
    if (cnt0 == 0) return false;
    const int r = push(P0, size0, cnt0);
    if (r < 0) { if (lane == 0) atomicOr(&c.st->flags, FLAG_ARENA_OVERFLOW); return false; }
    if (r == 0) return false;
  }

  while (depth > 0) {
    if (h_idx < 0) {  // frame exhausted -> pop
      fp = top;
      top = (size_t)h_prev;
      --depth;
      if (depth > 0) {
        __syncwarp();
        h_prev = (int)ar[top]; h_idx = (int)ar[top + 2]; h_size = (int)ar[top + 3];
      }
      continue;
    }
    const uint32_t ce = ar[top + 4 + Wc + h_idx];
    --h_idx;
    const int k = (int)(ce >> 16);
    const unsigned u = ce & 0xffffu;
    const int bnow = best_now();
    if (h_size + k <= bnow) { h_idx = -1; continue; }  // every remaining candidate has colour <= k
    ++c.nodes;
    if ((c.nodes & 255) == 0) {
      if (c.deadline && globaltimer_ns() > c.deadline) { if (lane == 0) atomicOr(&c.st->flags, FLAG_TIME_LIMIT); return false; }
      if (c.abort_below && *c.abort_below < c.my_id) return false;
    }
    uint32_t P[WPL], NP[WPL];
    const uint32_t* row = c.A.row(u);
#pragma unroll
    for (int s = 0; s < WPL; ++s) {
      const int w = lane + 32 * s;
      P[s] = (w < Wc) ? ar[top + 4 + w] : 0u;
      NP[s] = (w < Wc) ? (P[s] & row[w]) : 0u;
    }
    // remove u from this frame's P for the siblings still to come
    if (lane == (int)((u >> 5) & 31)) ar[top + 4 + (u >> 5)] &= ~(1u << (u & 31));
    const int cnt = bs_count<WPL>(NP);
    const int nsize = h_size + 1;
    if (c.decision) {
      if (nsize >= c.need) return true;
      if (cnt == 0 || nsize + cnt < c.need) continue;
    } else {
      if (cnt == 0) {
        if (nsize > bnow && lane == 0) atomicMax(c.best, nsize);
        continue;
      }
      if (nsize + cnt <= bnow) continue;
    }
    if (lane == 0) ar[top + 2] = (uint32_t)h_idx;
    const int r = push(NP, nsize, cnt);
    if (r < 0) { if (lane == 0) atomicOr(&c.st->flags, FLAG_ARENA_OVERFLOW); return false; }
  }
  return false;
}

// shared-memory copy of the compact adjacency when it fits
__device__ AdjView stage_adjacency(const ProbDesc& d, int np, int Wc, uint32_t* smem, size_t smem_words) {
  AdjView A;
  if ((size_t)np * Wc <= smem_words) {
    const int vec = Wc >> 2;  // uint4 per row
    for (int idx = threadIdx.x; idx < np * vec; idx += blockDim.x) {
      const int r = idx / vec, q = idx % vec;
      reinterpret_cast<uint4*>(smem)[(size_t)r * vec + q] = reinterpret_cast<const uint4*>(d.cadj + (size_t)r * d.pitch)[q];
    }
    A.base = smem; A.pitch = Wc;
  } else {
    A.base = d.cadj; A.pitch = d.pitch;
  }
  __syncthreads();
  return A;
}

template <int WPL>
__device__ void bnb_roots(const ProbDesc& d, const AdjView& A, int np, int Wc, uint32_t* ar, size_t cap,
                          unsigned long long deadline, int lane) {
  ProbState* st = d.st;
  DfsCtx c;
  c.A = A; c.Wc = Wc; c.ar = ar; c.cap = cap; c.best = &st->omega; c.decision = false; c.need = 0;
  c.abort_below = nullptr; c.my_id = 0; c.st = st; c.deadline = deadline; c.nodes = 0;
  for (;;) {
    unsigned r = 0;
    if (lane == 0) r = atomicAdd(&st->root_counter, 1u);
    r = __shfl_sync(FULL, r, 0);
    if (r >= (unsigned)np) break;
    if (*(volatile uint32_t*)&st->flags & (FLAG_ARENA_OVERFLOW | FLAG_TIME_LIMIT)) break;
    const unsigned v = r;
    uint32_t P[WPL];
    const uint32_t* row = A.row(v);
#pragma unroll
    for (int s = 0; s < WPL; ++s) {
      const int w = lane + 32 * s;
      P[s] = (w < Wc) ? row[w] : 0u;
    }
    bs_keep_above<WPL>(P, v, lane);
    const int cnt = bs_count<WPL>(P);
    if (1 + cnt <= *(volatile int*)&st->omega) continue;
    dfs<WPL>(P, 1, c, lane);
  }
  if (lane == 0 && c.nodes) atomicAdd(&st->bnb_nodes, c.nodes);
}

__global__ void __launch_bounds__(kBnbWarpsPerBlock * 32) bnb_kernel(const ProbDesc* __restrict__ descs, KernelParams kp,
                                                                    uint32_t* __restrict__ arena, size_t arena_words_per_warp,
                                                                    size_t smem_words) {
  extern __shared__ uint32_t sh[];
  const ProbDesc d = descs[blockIdx.y];
  ProbState* st = d.st;
  if (st->tail_done) return;
  __shared__ int sh_hdr[4];
  if (threadIdx.x == 0) {  // omega / flags change under us (other CTAs of this problem): read them once per CTA
    sh_hdr[0] = st->n_alive; sh_hdr[1] = st->Wc; sh_hdr[2] = *(volatile int*)&st->omega; sh_hdr[3] = (int)*(volatile uint32_t*)&st->flags;
  }
  __syncthreads();
  const int np = sh_hdr[0], Wc = sh_hdr[1];
  if (sh_hdr[3] & FLAG_COMPACT_TOO_LARGE) return;
  if (np <= sh_hdr[2]) return;  // nothing larger than the incumbent can exist
  const AdjView A = stage_adjacency(d, np, Wc, sh, smem_words);
  const int lane = threadIdx.x & 31;
  const size_t wid = ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * kBnbWarpsPerBlock + (threadIdx.x >> 5);
  uint32_t* ar = arena + wid * arena_words_per_warp;
  const unsigned long long deadline = kp.time_limit_ns ? globaltimer_ns() + kp.time_limit_ns : 0ull;
  if (Wc <= 32) bnb_roots<1>(d, A, np, Wc, ar, arena_words_per_warp, deadline, lane);
  else if (Wc <= 64) bnb_roots<2>(d, A, np, Wc, ar, arena_words_per_warp, deadline, lane);
  else if (Wc <= 128) bnb_roots<4>(d, A, np, Wc, ar, arena_words_per_warp, deadline, lane);
  else if (Wc <= 256) bnb_roots<8>(d, A, np, Wc, ar, arena_words_per_warp, deadline, lane);
  else if (Wc <= 512) bnb_roots<16>(d, A, np, Wc, ar, arena_words_per_warp, deadline, lane);
  else bnb_roots<32>(d, A, np, Wc, ar, arena_words_per_warp, deadline, lane);
}

// ---------------------------------------------------------------------------------------------------------
// 5. canonical maximum clique (lexicographically smallest sorted id list among the cliques of size omega)
// ---------------------------------------------------------------------------------------------------------
struct CanonShared {
  int need, outc, changed, cntP, mincnt, found, fail;
  unsigned next;
};

template <int WPL>
__device__ void canon_candidates(const ProbDesc& d, const AdjView& A, int Wc, const uint32_t* Pb, const int* plist,
                                 CanonShared* cs, uint32_t* ar, size_t cap, unsigned long long deadline, int lane) {
  DfsCtx c;
  c.A = A; c.Wc = Wc; c.ar = ar; c.cap = cap; c.best = nullptr; c.decision = true; c.need = cs->need - 1;
  c.abort_below = &cs->found; c.st = d.st; c.deadline = deadline; c.nodes = 0;
  const int cntP = cs->cntP;
  for (;;) {
    unsigned ci = 0;
    if (lane == 0) ci = atomicAdd(&cs->next, 1u);
    ci = __shfl_sync(FULL, ci, 0);
    if (ci >= (unsigned)cntP) break;
    const int v = plist[ci];
    if (v > *(volatile int*)&cs->found) break;
    uint32_t Q[WPL];
    const uint32_t* row = A.row((unsigned)v);
#pragma unroll
    for (int s = 0; s < WPL; ++s) {
      const int w = lane + 32 * s;
      Q[s] = (w < Wc) ? (Pb[w] & row[w]) : 0u;
    }
    bs_keep_above<WPL>(Q, (unsigned)v, lane);
    bool ok;
    if (c.need <= 0) ok = true;
    else if (bs_count<WPL>(Q) < c.need) ok = false;
    else { c.my_id = v; ok = dfs<WPL>(Q, 0, c, lane); }
    if (ok && lane == 0) atomicMin(&cs->found, v);
  }
  if (lane == 0 && c.nodes) atomicAdd(&d.st->bnb_nodes, c.nodes);
}

__global__ void __launch_bounds__(kCanonWarps * 32) canon_kernel(const ProbDesc* __restrict__ descs, KernelParams kp,
                                                                uint32_t* __restrict__ arena, size_t arena_words_per_warp,
                                                                size_t smem_words) {
  extern __shared__ uint32_t sh[];
  __shared__ CanonShared cs;
  const ProbDesc d = descs[blockIdx.x];
  ProbState* st = d.st;
  if (st->tail_done) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (st->flags & (FLAG_COMPACT_TOO_LARGE | FLAG_ARENA_OVERFLOW)) {
    if (threadIdx.x == 0) st->clique_size = 0;
    return;
  }
  const int np = st->n_alive, Wc = st->Wc;
  uint32_t* Pb = sh;        // Wc words: candidate set
  uint32_t* Rb = sh + Wc;   // Wc words: members chosen so far (compact ids)
  const AdjView A = stage_adjacency(d, np, Wc, sh + 2 * Wc, smem_words - 2 * (size_t)Wc);
  int* plist = d.rot_inl;   // scratch, free until the rotation stage
  int* deg_in = d.tls_id;   // scratch (>= n ints): degree of every member inside P
  uint32_t* ar = arena + ((size_t)blockIdx.x * kCanonWarps + warp) * arena_words_per_warp;
  const unsigned long long deadline = kp.time_limit_ns ? globaltimer_ns() + kp.time_limit_ns : 0ull;
  for (int w = threadIdx.x; w < Wc; w += blockDim.x) {
    uint32_t x = 0u;
    if (32 * w < np) x = (np - 32 * w >= 32) ? 0xffffffffu : ((1u << (np - 32 * w)) - 1u);
    Pb[w] = x;
    Rb[w] = 0u;
  }
  if (threadIdx.x == 0) { cs.need = st->omega; cs.outc = 0; cs.fail = 0; }
  __syncthreads();

  // Lexicographically smallest maximum clique of P.  Three moves, repeated:
  //  (a) peel P to its (need-1)-core;
  //  (b) members adjacent to every other member of P belong to every maximum clique of P: take them all at once
  //      (sorted(U + C') < sorted(U + C'') iff sorted(C') < sorted(C''), so removing U keeps the order of the rest);
  //  (c) otherwise the smallest id that still extends to a clique of the required size is the next member
  //      (decision searches, one warp per candidate).
  while (cs.need > 0 && !cs.fail) {
    const int need = cs.need;
    for (;;) {
      __syncthreads();
      if (threadIdx.x == 0) { cs.changed = 0; cs.cntP = 0; cs.mincnt = 0x7fffffff; }
      __syncthreads();
      if (Wc <= 64) {
        // short rows (the cores of real-descriptor frames: a few hundred vertices): one thread per member.  A warp per member leaves
        // 20+ lanes idle and pays a shuffle reduction and three lane-0 atomics per member; the loop runs once or twice per member of
        // the clique and was most of the 0.08 ms this kernel took on a config-2 frame.
        int kept = 0, cmin_t = 0x7fffffff;
        for (int v0 = 0; v0 < np; v0 += kCanonWarps * 32) {
          const int v = v0 + (int)threadIdx.x;
          if (v < np && ((Pb[v >> 5] >> (v & 31)) & 1)) {
            const uint32_t* row = A.row((unsigned)v);
            int c = 0;
            for (int w = 0; w < Wc; ++w) c += __popc(row[w] & Pb[w]);
            if (c < need - 1) { atomicAnd(&Pb[v >> 5], ~(1u << (v & 31))); cs.changed = 1; }
            else { ++kept; cmin_t = min(cmin_t, c); deg_in[v] = c; }
          }
        }
        kept = __reduce_add_sync(FULL, kept);
        cmin_t = __reduce_min_sync(FULL, cmin_t);
        if (lane == 0 && kept) { atomicAdd(&cs.cntP, kept); atomicMin(&cs.mincnt, cmin_t); }
      } else
      for (int v = warp; v < np; v += kCanonWarps) {
        if (!((Pb[v >> 5] >> (v & 31)) & 1)) continue;
        const uint32_t* row = A.row((unsigned)v);
        int c = 0;
        for (int w = lane; w < Wc; w += 32) c += __popc(row[w] & Pb[w]);
        c = __reduce_add_sync(FULL, c);
        if (lane == 0) {
          if (c < need - 1) { atomicAnd(&Pb[v >> 5], ~(1u << (v & 31))); cs.changed = 1; }
          else { atomicAdd(&cs.cntP, 1); atomicMin(&cs.mincnt, c); deg_in[v] = c; }
        }
      }
      __syncthreads();
      if (!cs.changed) break;
    }
    const int cntP = cs.cntP;
    if (cntP < need) { if (threadIdx.x == 0) cs.fail = 1; __syncthreads(); break; }
    if (cs.mincnt == cntP - 1) {  // P is a clique (of exactly `need` vertices, omega being the maximum)
      for (int w = threadIdx.x; w < Wc; w += blockDim.x) Rb[w] |= Pb[w];
      __syncthreads();
      if (threadIdx.x == 0) { cs.outc += cntP; cs.need = (cntP == need) ? 0 : -1; }
      __syncthreads();
      if (cs.need < 0) { if (threadIdx.x == 0) cs.fail = 1; __syncthreads(); }
      break;
    }
    // (b) universal members
    if (threadIdx.x == 0) cs.found = 0;
    __syncthreads();
    {
      int took = 0;
      for (int v = threadIdx.x; v < np; v += blockDim.x) {
        if (((Pb[v >> 5] >> (v & 31)) & 1) && deg_in[v] == cntP - 1) {
          atomicOr(&Rb[v >> 5], 1u << (v & 31));
          atomicAnd(&Pb[v >> 5], ~(1u << (v & 31)));
          ++took;
        }
      }
      if (took) atomicAdd(&cs.found, took);
    }
    __syncthreads();
    if (cs.found > 0) {
      const int took = cs.found;
      __syncthreads();
      if (threadIdx.x == 0) { cs.outc += took; cs.need -= took; }
      __syncthreads();
      continue;
    }
    // (c) ascending list of P's members, then the smallest feasible one
    if (warp == 0) {
      int off = 0;
      for (int w0 = 0; w0 < Wc; w0 += 32) {
        const int w = w0 + lane;
        const uint32_t word = (w < Wc) ? Pb[w] : 0u;
        const int c = __popc(word);
        int incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int t = __shfl_up_sync(FULL, incl, o);
          if (lane >= o) incl += t;
        }
        int pos = off + incl - c;
        uint32_t bits = word;
        while (bits) { plist[pos++] = 32 * w + (__ffs(bits) - 1); bits &= bits - 1; }
        off += __shfl_sync(FULL, incl, 31);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) { cs.found = 0x7fffffff; cs.next = 0u; }
    __syncthreads();
    if (Wc <= 32) canon_candidates<1>(d, A, Wc, Pb, plist, &cs, ar, arena_words_per_warp, deadline, lane);
    else if (Wc <= 64) canon_candidates<2>(d, A, Wc, Pb, plist, &cs, ar, arena_words_per_warp, deadline, lane);
    else if (Wc <= 128) canon_candidates<4>(d, A, Wc, Pb, plist, &cs, ar, arena_words_per_warp, deadline, lane);
    else if (Wc <= 256) canon_candidates<8>(d, A, Wc, Pb, plist, &cs, ar, arena_words_per_warp, deadline, lane);
    else if (Wc <= 512) canon_candidates<16>(d, A, Wc, Pb, plist, &cs, ar, arena_words_per_warp, deadline, lane);
    else canon_candidates<32>(d, A, Wc, Pb, plist, &cs, ar, arena_words_per_warp, deadline, lane);
    __syncthreads();
    const int vs = cs.found;
    if (vs == 0x7fffffff) { if (threadIdx.x == 0) cs.fail = 1; __syncthreads(); break; }
    const uint32_t* row = A.row((unsigned)vs);
    for (int w = threadIdx.x; w < Wc; w += blockDim.x) {
      uint32_t x = Pb[w] & row[w];
      if (w < (vs >> 5)) x = 0u;
      else if (w == (vs >> 5)) x &= ~((2u << (vs & 31)) - 1u);
      Pb[w] = x;
    }
    if (threadIdx.x == 0) { Rb[vs >> 5] |= 1u << (vs & 31); cs.outc += 1; cs.need -= 1; }
    __syncthreads();
  }
  __syncthreads();
  // emit the members in ascending compact id == ascending original id
  if (!cs.fail && warp == 0) {
    int off = 0;
    for (int w0 = 0; w0 < Wc; w0 += 32) {
      const int w = w0 + lane;
      const uint32_t word = (w < Wc) ? Rb[w] : 0u;
      const int c = __popc(word);
      int incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(FULL, incl, o);
        if (lane >= o) incl += t;
      }
      int pos = off + incl - c;
      uint32_t bits = word;
      while (bits) { d.clique[pos++] = d.ids[32 * w + (__ffs(bits) - 1)]; bits &= bits - 1; }
      off += __shfl_sync(FULL, incl, 31);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (cs.fail) { st->flags |= FLAG_CANON_FAILED; st->clique_size = 0; }
    else st->clique_size = cs.outc;
  }
}

// inlier_selection_mode == NONE: every correspondence is kept (A.5-7)
__global__ void __launch_bounds__(256) clique_all_kernel(const ProbDesc* __restrict__ descs) {
  const ProbDesc d = descs[blockIdx.y];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < d.n) d.clique[i] = i;
  if (i == 0) { d.st->clique_size = d.n; d.st->omega = d.n; }
}

// KCORE_HEU (A.5-3): max core by bisection on the peel threshold; if max_core > threshold * n the answer is the
// set of vertices of the innermost core, otherwise clique_size stays -1 and the heuristic replay decides.
__global__ void __launch_bounds__(1024) kcore_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  extern __shared__ uint32_t sh[];
  const ProbDesc d = descs[blockIdx.x];
  ProbState* st = d.st;
  const int n = d.n;
  const int W = words_for(n);
  const int Wb = (n + 31) >> 5;
  uint32_t* alive = sh;
  int* sh_scan = (int*)(sh + W);
  int* sh_changed = sh_scan + 36;
  int* sh_any = sh_scan + 37;
  auto core_nonempty = [&](int k) -> bool {  // is the k-core non-empty?  (leaves it in `alive`)
    __syncthreads();
    for (int w = threadIdx.x; w < W; w += blockDim.x) alive[w] = 0u;
    __syncthreads();
    for (int v0 = 0; v0 < n; v0 += blockDim.x) {
      const int v = v0 + threadIdx.x;
      const bool a = v < n && d.deg[v] >= k;
      const uint32_t bal = __ballot_sync(FULL, a);
      if ((threadIdx.x & 31) == 0 && (v >> 5) < W) alive[v >> 5] = bal;
    }
    __syncthreads();
    if (k > 0) peel_full(d, W, alive, k, sh_changed);
    __syncthreads();
    if (threadIdx.x == 0) *sh_any = 0;
    __syncthreads();
    for (int w = threadIdx.x; w < Wb; w += blockDim.x) if (alive[w]) *sh_any = 1;
    __syncthreads();
    return *sh_any != 0;
  };
  int lo = 0, hi = st->max_deg;  // invariant: lo-core non-empty
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (core_nonempty(mid)) lo = mid; else hi = mid - 1;
  }
  const int max_core = lo;
  if (kp.kcore_threshold != 1.0 && max_core > (int)(kp.kcore_threshold * (double)n)) {
    core_nonempty(max_core);
    const int cnt = number_alive(alive, Wb, d.clique, sh_scan);
    if (threadIdx.x == 0) { st->clique_size = cnt; st->omega = cnt; }
  } else if (threadIdx.x == 0) {
    st->clique_size = -1;
  }
}

__global__ void mark_undecided_kernel(const ProbDesc* __restrict__ descs, int batch) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < batch) descs[b].st->clique_size = -1;
}

// ---------------------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------------------
static int wpl_for(int n_max) {
  const int W = 4 * ((n_max + kTile - 1) / kTile);
  const int need = (W + 31) / 32;
  int w = 1;
  while (w < need) w <<= 1;
  return w;
}

template <template <int> class Launcher>
static void dispatch_wpl(int wpl, const ProbDesc* d_desc, dim3 grid, int block, size_t smem, int arg, cudaStream_t st) {
  switch (wpl) {
    case 1: Launcher<1>::go(d_desc, grid, block, smem, arg, st); break;
    case 2: Launcher<2>::go(d_desc, grid, block, smem, arg, st); break;
    case 4: Launcher<4>::go(d_desc, grid, block, smem, arg, st); break;
    case 8: Launcher<8>::go(d_desc, grid, block, smem, arg, st); break;
    case 16: Launcher<16>::go(d_desc, grid, block, smem, arg, st); break;
    case 32: Launcher<32>::go(d_desc, grid, block, smem, arg, st); break;
    default: Launcher<64>::go(d_desc, grid, block, smem, arg, st); break;
  }
}
template <int WPL> struct HeuLaunch {
  static void go(const ProbDesc* d, dim3 g, int b, size_t smem, int all_seeds, cudaStream_t s) {
    static PerDeviceOnce attr;
    attr.run([] { return cudaFuncSetAttribute(heuristic_kernel<WPL>, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024); });
    heuristic_kernel<WPL><<<g, b, smem, s>>>(d, all_seeds);
  }
};
template <int WPL> struct ReplayLaunch {
  static void go(const ProbDesc* d, dim3 g, int b, size_t smem, int, cudaStream_t s) { heuristic_replay_kernel<WPL><<<g, b, smem, s>>>(d); }
};

// opt-in shared memory of the current device; the kernels' attributes are set once PER DEVICE before the value is used
static size_t smem_optin_bytes() {
  static PerDeviceOnce attr;
  int dev = 0, v = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  const size_t optin = (size_t)v;
  attr.run([&] {
    const size_t dyn = optin - 2048;
    cudaError_t e = cudaFuncSetAttribute(bnb_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(canon_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(prune_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(kcore_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(heuristic_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    return e;
  });
  return optin;
}

int launch_clique(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, uint32_t* arena,
                  size_t arena_words_per_warp, int bnb_blocks, cudaStream_t st, cudaEvent_t* ev, bool bound_known) {
  int launches = 0;
  const int wpl = wpl_for(n_max);
  const int W = 4 * ((n_max + kTile - 1) / kTile);
  const size_t optin = smem_optin_bytes();
  const size_t prune_smem = (size_t)(W + 64) * 4;
  dim3 gseed((n_max + 7) / 8, batch);
  if (kp.inlier_mode == B2R_INLIER_NONE) {
    clique_all_kernel<<<dim3((n_max + 255) / 256, batch), 256, 0, st>>>(d_desc);
    ++launches;
    if (ev) for (int i = 0; i < 4; ++i) cudaEventRecord(ev[i], st);
    return launches;
  }
  if (kp.inlier_mode == B2R_INLIER_KCORE_HEU) {
    kcore_kernel<<<batch, 1024, prune_smem, st>>>(d_desc, kp);
    ++launches;
  } else if (kp.inlier_mode == B2R_INLIER_PMC_HEU) {
    mark_undecided_kernel<<<(batch + 127) / 128, 128, 0, st>>>(d_desc, batch);
    ++launches;
  }
  const bool all_seeds = kp.inlier_mode != B2R_INLIER_PMC_EXACT;
  if (all_seeds) {
    dispatch_wpl<HeuLaunch>(wpl, d_desc, gseed, 256, (size_t)8 * W * 4, 1, st);
  } else if (!bound_known) {   // bound_known: every open problem comes from the fused tail, which ran the same seeds (st->lb, st->heu_key)
    const size_t hs = (size_t)(3 * W + 64) * 4 + (size_t)2 * (n_max + 8);
    heuristic_block_kernel<<<dim3(kHeuSeeds, batch), kHeuThreads, hs, st>>>(d_desc);
  }
  if (all_seeds || !bound_known) ++launches;
  if (ev) cudaEventRecord(ev[0], st);
  if (kp.inlier_mode != B2R_INLIER_PMC_EXACT) {
    dispatch_wpl<ReplayLaunch>(wpl, d_desc, dim3(batch), 32, (size_t)W * 4, 0, st);
    ++launches;
    if (ev) for (int i = 1; i < 4; ++i) cudaEventRecord(ev[i], st);
    return launches;
  }
  prune_kernel<<<batch, 1024, prune_smem, st>>>(d_desc);
  const int cblocks = batch >= 64 ? 16 : (batch >= 8 ? 64 : 296);
  compact_kernel<<<dim3(cblocks, batch), 256, 0, st>>>(d_desc);
  launches += 2;
  if (ev) cudaEventRecord(ev[1], st);
  const size_t dyn = optin - 2048;
  bnb_kernel<<<dim3(bnb_blocks, batch), kBnbWarpsPerBlock * 32, dyn, st>>>(d_desc, kp, arena, arena_words_per_warp, dyn / 4);
  ++launches;
  if (ev) cudaEventRecord(ev[2], st);
  // the canonicalisation reuses the first batch*kCanonWarps arenas
  canon_kernel<<<batch, kCanonWarps * 32, dyn, st>>>(d_desc, kp, arena, arena_words_per_warp, dyn / 4);
  ++launches;
  if (ev) cudaEventRecord(ev[3], st);
  return launches;
}

}  // namespace b2r
