// Fused tail: everything `solver.solve` (HelpersRealtime.py:166, registration_utils.py:261) does after the consistency graph, for one
// registration per CTA, in ONE launch -- exact fix-up of the pairs the tensor-core filter listed, vertex-degree maxima and seeds,
// greedy clique bound, (lb-1)-core pruning, and when the core IS the clique (the proof of maximality upstream's PMC reaches through
// max_core + 1 == lb, A.5-4) the chain TIMs + GNC-TLS rotation (A.6) and the adaptive-voting translation (A.7).
//
// Why: the stages are tiny and strictly dependent (SURVEY 8a-4 .. a-6: "latency only -> batch across detections"); as eight
// grid-wide launches of ~64 CTAs they cost 0.29 ms per 64 x 5000 batch, three quarters of it launch gaps and cold first touches.
// Problems are independent, so a CTA carries one from the graph to the pose without any grid-wide synchronisation.
//
// A problem the fused path cannot finish is HANDED OVER untouched to the multi-kernel path (clique.cu / rt.cu), which b2r_sync
// launches for the batch when it finds ProbState::tail_done == 0:
//   tail_why 1  the graph itself is not final (fp32 path refused the problem or its pair list overflowed): fp64 filter first
//   tail_why 2  the (lb-1)-core is larger than the greedy clique: the exact branch and bound has to decide
// Results are identical on both paths: same heuristic rule, same peel, same rotation / translation bodies (rt.cuh).
#include <stdlib.h>

#include "clique_dev.cuh"
#include "exact.cuh"
#include "rt.cuh"

namespace b2r {

constexpr int kTailThreads = 3 * kRtThreads;   // 768: the translation stage needs exactly three groups of kRtThreads
constexpr int kTailWarps = kTailThreads / 32;
#ifndef B2R_HEU_SEEDS
#define B2R_HEU_SEEDS 2
#endif

constexpr int kTailMaxExtra = 8;                               // |core| - lb up to which the fused tail decides the clique itself
constexpr int kTailMaxPairs = kTailMaxExtra * kTailMaxExtra;   // a core of lb + e vertices with omega >= lb has at most e * e non-adjacent pairs

struct TailShared {
  unsigned long long key[8];
  unsigned long long deg_sum;
  int lb, changed, cnt, base;
  int cmin;          // smallest degree inside the surviving set seen by the current peel pass
  int scan[40];
  // near-clique core: the non-adjacent pairs inside the core and the vertices to drop (tail_small_cover)
  int n_pairs, n_cover;
  unsigned short pu[kTailMaxPairs], pv[kTailMaxPairs], cover[kTailMaxExtra];
  unsigned long long mu[kTailMaxPairs], mv[kTailMaxPairs];   // tail_cover_masks
};

// ascending list of the set bits of `bits` (Wb words) as 16-bit ids; returns the count (uniform)
__device__ int tail_list_bits(const uint32_t* bits, int Wb, unsigned short* list, TailShared* ts) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (threadIdx.x == 0) ts->base = 0;
  __syncthreads();
  for (int w0 = 0; w0 < Wb; w0 += kTailThreads) {
    const int w = w0 + threadIdx.x;
    const uint32_t word = (w < Wb) ? bits[w] : 0u;
    const int c = __popc(word);
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) ts->scan[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      int x = lane < kTailWarps ? ts->scan[lane] : 0, xi = x;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(FULL, xi, o); if (lane >= o) xi += t; }
      ts->scan[lane] = xi - x;
      if (lane == 31) ts->scan[32] = xi;
    }
    __syncthreads();
    int pos = ts->base + ts->scan[warp] + incl - c;
    uint32_t b = word;
    while (b) { list[pos++] = (unsigned short)(32 * w + (__ffs(b) - 1)); b &= b - 1; }
    __syncthreads();
    if (threadIdx.x == 0) ts->base += ts->scan[32];
    __syncthreads();
  }
  return ts->base;
}

// Exact maximum clique of a core that is a clique up to a few vertices (|core| = lb + e, e <= kTailMaxExtra): every core vertex has at
// least lb - 1 neighbours inside the core, i.e. at most e non-neighbours, so the COMPLEMENT of the core's graph is tiny; a maximum
// clique is the core minus a minimum vertex cover of those non-adjacent pairs (size k* = |core| - omega <= e).  One thread enumerates
// all covers of minimum size by branching on the two ends of an uncovered pair (depth <= e, <= 2^e leaves) and keeps the one whose
// sorted id list is lexicographically LARGEST: dropping it leaves the lexicographically smallest maximum clique, the canonical
// answer of the oracle and of canon_kernel.  Returns the cover in ts->cover / ts->n_cover (-1: no cover of size <= e exists).
//
// Pairs are bits of a 64-bit word (np <= kTailMaxPairs = 64): ts->mu[i] / ts->mv[i] = the pairs that the first / second end of pair i
// covers (tail_cover_masks, one thread per pair), so a node of the search is an OR, a find-first-set and a popcount instead of a scan
// of the pair list in shared memory (172 k cycles for 11 pairs, e = 4, before; the popcount bound -- a vertex covers at most e pairs
// -- also cuts branches that cannot finish within k).
__device__ void tail_cover_masks(TailShared* ts) {
  const int np = ts->n_pairs;
  const int i = threadIdx.x;
  if (i < np) {
    const unsigned short u = ts->pu[i], v = ts->pv[i];
    unsigned long long mu = 0ull, mv = 0ull;
    for (int j = 0; j < np; ++j) {
      const unsigned short a = ts->pu[j], b = ts->pv[j];
      if (a == u || b == u) mu |= 1ull << j;
      if (a == v || b == v) mv |= 1ull << j;
    }
    ts->mu[i] = mu; ts->mv[i] = mv;
  }
}

__device__ void tail_small_cover(TailShared* ts, int e) {
  const int np = ts->n_pairs;
  const unsigned long long all = np >= 64 ? ~0ull : ((1ull << np) - 1ull);
  unsigned long long cov[kTailMaxExtra + 1];
  int sel[kTailMaxExtra], eidx[kTailMaxExtra];
  int best[kTailMaxExtra], cand[kTailMaxExtra];
  unsigned second = 0u;   // bit d: depth d is on the second end of its pair
  int best_n = -1;
  for (int k = 0; k <= e && best_n < 0; ++k) {
    int depth = 0;
    cov[0] = 0ull;
    for (;;) {
      const unsigned long long un = all & ~cov[depth];
      if (!un) {   // sel[0 .. depth) covers every pair
        for (int q = 0; q < depth; ++q) cand[q] = sel[q];
        for (int a = 1; a < depth; ++a) { const int x = cand[a]; int b = a - 1; while (b >= 0 && cand[b] > x) { cand[b + 1] = cand[b]; --b; } cand[b + 1] = x; }
        bool better = best_n < 0;
        if (!better && depth == best_n)
          for (int q = 0; q < depth; ++q) if (cand[q] != best[q]) { better = cand[q] > best[q]; break; }
        if (better) { best_n = depth; for (int q = 0; q < depth; ++q) best[q] = cand[q]; }
      } else if (depth < k && __popcll(un) <= (k - depth) * e) {
        const int ei = __ffsll((long long)un) - 1;
        eidx[depth] = ei; second &= ~(1u << depth); sel[depth] = (int)ts->pu[ei];
        cov[depth + 1] = cov[depth] | ts->mu[ei];
        ++depth;
        continue;
      }
      // backtrack: the other end of the deepest pair that still has one
      while (depth > 0) {
        const int dd = depth - 1;
        if (!((second >> dd) & 1u)) {
          second |= 1u << dd; sel[dd] = (int)ts->pv[eidx[dd]];
          cov[depth] = cov[dd] | ts->mv[eidx[dd]];
          break;
        }
        --depth;
      }
      if (depth == 0) break;
    }
  }
  ts->n_cover = best_n;
  for (int q = 0; q < best_n; ++q) ts->cover[q] = (unsigned short)best[q];
}

// mode 0: the graph comes from the tensor-core filter of this batch (pair list to fix, degrees accumulated by the filter);
// mode 1: the graph is final and degree_kernel has run (other filters, estimate_scaling, b2r_solve_from_adjacency)
__global__ void __launch_bounds__(kTailThreads, 1) tail_kernel(const ProbDesc* __restrict__ descs, KernelParams kp, int mode, int dyn_smem_bytes) {
  extern __shared__ __align__(16) unsigned char tsm[];
  __shared__ TailShared ts;
  __shared__ RowCache rcache;
  __shared__ __align__(8) uint64_t rc_bar;
  const ProbDesc d = descs[blockIdx.x];
  ProbState* st = d.st;
  const int n = d.n;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int W = words_for(n), Wb = (n + 31) >> 5;
  const long long clk0 = clock64();
  auto stamp = [&](int k) { if (tid == 0) st->tail_clk[k] = (unsigned int)(clock64() - clk0); };
  TailMarks tm;
  tm.clk = st->tail_clk; tm.t0 = clk0;
  if (tid < 24) st->tail_clk[8 + tid] = 0u;
  if (st->flags & (FLAG_SHARD_FALLBACK | FLAG_SHARD_TIMEOUT)) return;   // sharded solve: the matrix is not valid, the host redoes the problem

  if (mode == 0) {
    if (!st->use32 || st->tc_overflow) {   // uniform: the state was written by earlier kernels
      if (tid == 0) st->tail_why = 1;
      return;
    }
    // ---- 1. exact decision of the listed pairs (graph_fix_kernel's first half) ----
    const unsigned cnt = min(st->amb_count, (unsigned)d.amb_cap);
    stamp(7);
    // two pairs of a thread in flight together (registers): the list entries first, then their 24 coordinate loads, then the arithmetic
    constexpr int FQ = 2;
    for (unsigned i0 = tid; i0 < cnt; i0 += FQ * kTailThreads) {
      uint32_t e[FQ];
      FixPair fp[FQ];
#pragma unroll
      for (int q = 0; q < FQ; ++q) { const unsigned idx = i0 + q * kTailThreads; e[q] = idx < cnt ? d.amb_list[idx] : 0xffffffffu; }
#pragma unroll
      for (int q = 0; q < FQ; ++q) if (e[q] != 0xffffffffu) fp[q] = load_listed_pair(d.src, d.dst, e[q]);
#pragma unroll
      for (int q = 0; q < FQ; ++q) if (e[q] != 0xffffffffu) decide_listed_pair(fp[q], d.adj, d.pitch, d.deg, e[q], kp.beta);
    }
    if (tid < 8) ts.key[tid] = 0ull;
    if (tid == 8) ts.deg_sum = 0ull;
    __threadfence();
    __syncthreads();
    stamp(0);
    // ---- 2. degree maxima, 2 * edges, seed of every id class (degree_kernel's fast path) ----
    {
      unsigned long long key = 0ull, sum = 0ull;   // kTailThreads % 8 == 0: a thread only ever sees vertices of class tid & 7
      for (int v = tid; v < n; v += kTailThreads) {
        const int c = __ldcg(d.deg + v);
        sum += (unsigned long long)c;
        const unsigned long long k = ((unsigned long long)c << 32) | (unsigned long long)(0xffffffffu - (unsigned)v);
        key = k > key ? k : key;
      }
#pragma unroll
      for (int o = 8; o <= 16; o <<= 1) {
        const unsigned long long other = __shfl_xor_sync(FULL, key, o);
        key = other > key ? other : key;
      }
      sum = (unsigned long long)__reduce_add_sync(FULL, (unsigned)sum);
      if (lane < 8 && key) atomicMax(&ts.key[lane], key);
      if (lane == 0) atomicAdd(&ts.deg_sum, sum);
    }
    __syncthreads();
    if (tid == 0) {
      unsigned long long m = 0;
      for (int k = 0; k < 8; ++k) { st->seed_key[k] = ts.key[k]; m = ts.key[k] > m ? ts.key[k] : m; }
      st->max_deg = (int)(m >> 32);
      st->deg_sum = ts.deg_sum;
      st->slow_pairs = cnt;
      st->deg_ready = 1;
    }
  } else {
    if (tid < 8) ts.key[tid] = st->seed_key[tid];
    __syncthreads();
  }

  stamp(1);
  // ---- 3 + 4. greedy clique bound from the seed of an id class (heuristic_block_kernel), then the (lb-1)-core of the graph
  // (prune_kernel: alive bitset + member list in shared memory, one warp per eight members).  When the core has exactly lb
  // vertices the bound is proven and the seeds of the other classes are not needed; otherwise the next seed may raise the bound
  // (the multi-kernel path always runs all seeds: the bound is the same whenever it matters).
  uint32_t* sh = reinterpret_cast<uint32_t*>(tsm);
  uint32_t* alive = sh;                                               // W words
  unsigned short* list = reinterpret_cast<unsigned short*>(sh + W);   // n entries
  if (tid == 0) {
    ts.lb = 0;
    // row cache: whatever shared memory the launch got beyond the heuristic's own arrays (the translation scratch reuses all of it later)
    size_t off = (heuristic_smem_bytes(n) + 15) & ~(size_t)15;
    rcache.P1 = reinterpret_cast<uint32_t*>(tsm + off); off += (size_t)W * 4;
    rcache.pre = reinterpret_cast<unsigned short*>(tsm + off); off += ((size_t)W * 2 + 15) & ~(size_t)15;
    rcache.rows = reinterpret_cast<uint32_t*>(tsm + off);
    rcache.cap = (size_t)dyn_smem_bytes > off ? (int)(((size_t)dyn_smem_bytes - off) / ((size_t)W * 4)) : 0;
    rcache.n_cached = 0; rcache.W = W; rcache.mbar = &rc_bar;
    rc_mbar_init(&rc_bar);
  }
  __syncthreads();
  unsigned long long heu_key = 0ull;
  int lb = 0, T = 0, n_alive = 0;
  bool core_is_clique = false;
  int n_cl = 0;
  for (int k = 0; k < B2R_HEU_SEEDS; ++k) {
    const unsigned long long skey = ts.key[k];
    if (skey == 0ull) continue;
    const int v = (int)(0xffffffffu - (unsigned)(skey & 0xffffffffu));
    const int size = heuristic_seed_cta<kTailThreads, true, true>(d, v, sh, &ts.lb, &rcache, k == 0 ? tm : TailMarks());
    __syncthreads();
    if (size > 0) {
      const unsigned long long hk = ((unsigned long long)size << 32) | (unsigned long long)(0xffffffffu - (unsigned)v);
      heu_key = hk > heu_key ? hk : heu_key;
      if (tid == 0 && size > ts.lb) ts.lb = size;
    }
    __syncthreads();
    if (k > 0 && ts.lb == lb) continue;   // this seed did not raise the bound: the core is the one already computed
    lb = ts.lb;
    if (k == 0) stamp(2);
    T = max(lb - 1, 0);
    for (int w = tid; w < W; w += kTailThreads) alive[w] = 0u;
    __syncthreads();
    for (int v0 = 0; v0 < n; v0 += kTailThreads) {
      const int vv = v0 + tid;
      const bool a = vv < n && __ldcg(d.deg + vv) >= T;
      const uint32_t bal = __ballot_sync(FULL, a);
      if (lane == 0 && (vv >> 5) < W) alive[vv >> 5] = bal;
    }
    for (;;) {
      n_alive = tail_list_bits(alive, Wb, list, &ts);
      if (tid == 0) { ts.changed = 0; ts.cmin = 0x7fffffff; }
      __syncthreads();
      // |alive| == lb: every member of the greedy clique has at least lb - 1 = T neighbours, so it is inside `alive`, and a set of
      // lb vertices that contains a clique of size lb IS that clique -- all of it survives any further pass: nothing to peel
      if (T > 0 && n_alive != lb && (W >> 2) <= kHeuThreadPerMemberW4) {
        // short rows: one thread per member (see heuristic_seed_cta)
        const int W4 = W >> 2;
        const uint4* A4 = reinterpret_cast<const uint4*>(alive);
        int cmin_t = 0x7fffffff;
        for (int m0 = 0; m0 < n_alive; m0 += kTailThreads) {
          const int m = m0 + tid;
          if (m < n_alive) {
            const int id = (int)list[m];
            const uint4* rp = cached_row(rcache, id);
            if (!rp) rp = reinterpret_cast<const uint4*>(d.adj + (size_t)id * d.pitch);
            int ww = lane % W4, c = 0;
            for (int k = 0; k < W4; ++k) {
              const uint4 x = rp[ww], mk = A4[ww];
              c += __popc(x.x & mk.x) + __popc(x.y & mk.y) + __popc(x.z & mk.z) + __popc(x.w & mk.w);
              ww = ww + 1 == W4 ? 0 : ww + 1;
            }
            cmin_t = min(cmin_t, c);
            if (c < T) { atomicAnd(&alive[id >> 5], ~(1u << (id & 31))); ts.changed = 1; }
          }
        }
        cmin_t = __reduce_min_sync(FULL, cmin_t);
        if (lane == 0 && cmin_t < ts.cmin) atomicMin(&ts.cmin, cmin_t);
      } else if (T > 0 && n_alive != lb) {
        constexpr int R = 8;
        for (int m0 = warp; m0 < n_alive; m0 += R * kTailWarps) {
          int c[R], ids[R];
          const uint4* rowp[R];
#pragma unroll
          for (int q = 0; q < R; ++q) {
            const int m = m0 + q * kTailWarps;
            ids[q] = m < n_alive ? (int)list[m] : -1;
            rowp[q] = nullptr;
            if (ids[q] >= 0) {
              rowp[q] = cached_row(rcache, ids[q]);
              if (!rowp[q]) rowp[q] = reinterpret_cast<const uint4*>(d.adj + (size_t)ids[q] * d.pitch);
            }
          }
          rows_and_count<R>(rowp, alive, W >> 2, lane, c);
#pragma unroll
          for (int q = 0; q < R; ++q) {
            if (ids[q] < 0) continue;
            if (lane == 0 && c[q] < ts.cmin) atomicMin(&ts.cmin, c[q]);
            if (c[q] < T && lane == 0) { atomicAnd(&alive[ids[q] >> 5], ~(1u << (ids[q] & 31))); ts.changed = 1; }
          }
        }
      }
      __syncthreads();
      if (!ts.changed) break;
    }
    // Note: a vertex counted against neighbours that die in the same pass is caught by the next pass (the loop ends on a pass
    // without removals), so the fixed point is the T-core whatever the order -- the same set prune_kernel leaves.
    // The last pass removed nothing, so ts.cmin is the smallest degree inside the core.  Every clique of at least lb vertices lies
    // inside the (lb-1)-core; when every core vertex is adjacent to all the others the core itself is a clique, hence THE maximum
    // clique, although the greedy bound missed it by a vertex or two (lb < omega = |core|).
    core_is_clique = T > 0 && n_alive > lb && ts.cmin == n_alive - 1;
    n_cl = n_alive;
    if (n_alive == lb || core_is_clique) break;   // the core IS the clique: proven
    if (!(n_alive == lb || core_is_clique) && lb > 0 && n_alive - lb <= kTailMaxExtra) {
      // ---- 4b. near-clique core: list its non-adjacent pairs (u < v), drop a minimum vertex cover of them (tail_small_cover) ----
      if (tid == 0) { ts.n_pairs = 0; ts.n_cover = -1; }
      __syncthreads();
      for (int m = warp; m < n_alive; m += kTailWarps) {
        const int id = (int)list[m];
        const uint32_t* row = reinterpret_cast<const uint32_t*>(cached_row(rcache, id));
        if (!row) row = d.adj + (size_t)id * d.pitch;
        for (int w = lane; w < Wb; w += 32) {
          uint32_t nb = alive[w] & ~row[w];
          if (w == (id >> 5)) nb &= ~(1u << (id & 31));
          while (nb) {
            const int j = 32 * w + (__ffs(nb) - 1);
            nb &= nb - 1;
            if (j > id) {
              const int slot = atomicAdd(&ts.n_pairs, 1);
              if (slot < kTailMaxPairs) { ts.pu[slot] = (unsigned short)id; ts.pv[slot] = (unsigned short)j; }
            }
          }
        }
      }
      __syncthreads();
      if (ts.n_pairs <= kTailMaxPairs) {
        {
          // the search order must not depend on which warp listed a pair first: order the pairs by (u, v) -- rank sort, one thread per
          // pair (the pairs are distinct)
          const int np = ts.n_pairs;
          unsigned key = 0u;
          int rank = 0;
          if (tid < np) {
            key = ((unsigned)ts.pu[tid] << 16) | ts.pv[tid];
            for (int j = 0; j < np; ++j) rank += ((((unsigned)ts.pu[j] << 16) | ts.pv[j]) < key) ? 1 : 0;
          }
          __syncthreads();
          if (tid < np) { ts.pu[rank] = (unsigned short)(key >> 16); ts.pv[rank] = (unsigned short)(key & 0xffffu); }
        }
        __syncthreads();
        tail_cover_masks(&ts);
        __syncthreads();
        if (tid == 0) {
          tail_small_cover(&ts, n_alive - lb);
          for (int q = 0; q < ts.n_cover; ++q) alive[ts.cover[q] >> 5] &= ~(1u << (ts.cover[q] & 31));
        }
        __syncthreads();
        if (ts.n_cover >= 0) {
          n_cl = tail_list_bits(alive, Wb, list, &ts);
          core_is_clique = true;     // what is left of the core is a maximum clique, the canonical one
        }
      }
    }
    if (core_is_clique) break;   // decided exactly: the other seeds are not needed
    // (the next seed's heuristic reuses this shared memory: `alive` and `list` are only valid when the loop is left by a break)
  }
  stamp(3);
  const int Wc = (((n_alive + 31) >> 5) + 3) & ~3;
  for (int m = tid; m < n_alive; m += kTailThreads) d.ids[m] = (int)list[m];
  if (tid == 0) {
    st->lb = lb; st->heu_key = heu_key;
    st->T = T; st->n_alive = n_alive; st->Wc = Wc; st->omega = core_is_clique ? n_cl : lb;
    if (Wc > kMaxCompactWords) st->flags |= FLAG_COMPACT_TOO_LARGE;
  }
  if (!(n_alive == lb || core_is_clique) || lb <= 0) {   // the core is neither the greedy clique nor close to a clique: the exact search decides (multi-kernel path)
    if (tid == 0) st->tail_why = 2;
    return;
  }
  // ---- 5. the clique: the (lb-1)-core itself (lb vertices each adjacent to the lb-1 others, or proven above), or the core minus the
  // cover of its non-adjacent pairs ----
  for (int m = tid; m < n_cl; m += kTailThreads) d.clique[m] = (int)list[m];
  if (tid == 0) st->clique_size = n_cl;
  __threadfence_block();
  __syncthreads();
  stamp(4);

  // ---- 6. rotation on the first thread group, then the three translation axes on all three ----
  if (tid < kRtThreads) rotation_body(d, kp, 4, tm, reinterpret_cast<double*>(tsm), (size_t)dyn_smem_bytes / sizeof(double));
  __syncthreads();
  stamp(5);
  translation_body(d, kp, reinterpret_cast<double*>(tsm), tm);
  __syncthreads();
  stamp(6);
  if (tid == 0) st->tail_done = 1;
}

size_t tail_smem_bytes(int n_max) {
  const size_t tr = 3 * kTrGroupDoubles * sizeof(double);
  const size_t he = heuristic_smem_bytes(n_max);
  return (tr > he ? tr : he) + 16;
}

// The kernel always gets all the shared memory an SM can give one CTA (it runs one CTA per SM by its thread count anyway): what
// the stages do not need themselves holds staged adjacency rows (RowCache).
int launch_tail(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, int mode, cudaStream_t st) {
  static PerDeviceOnce attr;
  static int dyn_max[64];
  int dev = 0;
  cudaGetDevice(&dev);
  attr.run([dev] {
    int optin = 0;
    cudaError_t e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaFuncAttributes fa;
    if (e == cudaSuccess) e = cudaFuncGetAttributes(&fa, tail_kernel);
    if (e != cudaSuccess) return e;
    dyn_max[dev & 63] = optin - (int)fa.sharedSizeBytes;
    return cudaFuncSetAttribute(tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn_max[dev & 63]);
  });
  int dyn = dyn_max[dev & 63];
  // not the whole SM: with the maximum carve-out (227 KB) the first stage of the kernel ran 60 k cycles longer on B200 (0.183 vs
  // 0.133 ms per batch) -- 160 KB holds 235 rows at N = 5000.  B2R_TAIL_SMEM_KB overrides (A/B runs).
  static const int cap_kb = [] { const char* e = getenv("B2R_TAIL_SMEM_KB"); return e ? atoi(e) : 160; }();
  if (cap_kb > 0 && dyn > cap_kb * 1024) dyn = cap_kb * 1024;
  if (dyn < (int)tail_smem_bytes(n_max)) dyn = (int)tail_smem_bytes(n_max);   // (attribute query failed: the launch will report it)
  tail_kernel<<<batch, kTailThreads, dyn, st>>>(d_desc, kp, mode, dyn);
  return 1;
}

}  // namespace b2r
