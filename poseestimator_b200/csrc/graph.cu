// TIM pairwise-consistency graph.
//
// Replaces, for `solver.solve(src, dst)` (HelpersRealtime.py:166, registration_utils.py:261), upstream's
// computeTIMs x2 + ScaleInliersSelector::solveForScale + the addEdge loop: for every
// pair i<j an edge exists iff | ||src_j-src_i|| - ||dst_j-dst_i|| | <= 2*noise_bound*sqrt(cbar2).
//
// Nothing O(N^2) is materialised except the answer: a symmetric bit matrix u32[N][pitch].
// One CTA (128 threads) owns a 128x128 tile of the upper triangle: thread t holds column point j in registers,
// row points stream from shared memory as broadcast 16-byte loads, every 32 rows a warp ballot yields the 32-bit
// word of row i, and the per-lane OR of the verdicts is the word of the transposed tile -- so the lower triangle
// costs no arithmetic.  Rows leave the CTA as 16-byte stores.
//
// Bit-exactness: the verdict must equal the reference's fp64 evaluation
//     a = sqrt((dx*dx + dy*dy) + dz*dz), b likewise, |a - b| <= beta        (no FMA contraction, A.4-3)
// A and B are formed with exactly those roundings (__dmul_rn/__dadd_rn).  The two DSQRTs are avoided by a
// sqrt-free sign test that is provably equivalent outside a tiny ambiguity band (DESIGN.md "graph filter"):
//     D = A + B - beta^2,  F = D^2 - 4AB,  in  <=>  D <= 0  or  F <= 0,
// trusted only when |F| is at least 2^-28 of (A + B + beta^2)^2 (seven orders above the rounding error of F and of
// the reference's own sqrt/sub); a 32-row chunk containing any other pair (about 1e-4 of the chunks) is redone with
// the exact two-sqrt path.  The hot loop is branch-free: 20 FP64-pipe instructions per pair, the sign and band tests
// run on the integer pipe.
#include "common.cuh"

namespace b2r {

__device__ __forceinline__ double sq3_ref(double dx, double dy, double dz) {
  // (dx*dx + dy*dy) + dz*dz, each operation rounded separately (reference order, no FMA)
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

__device__ __forceinline__ bool verdict_exact(double A, double B, double beta) {
  const double a = __dsqrt_rn(A), b = __dsqrt_rn(B);
  return fabs(__dsub_rn(a, b)) <= beta;
}

// estimate_scaling=True (A.4-5 + A.7-3): the edge set is the scale-inlier mask |b/a - s| <= beta * (1/a),
// s = the adaptive-voting estimate (scale.cu), every operation rounded as the reference's doubles are.
__device__ __forceinline__ bool verdict_scale(double A, double B, double beta, double s) {
  const double a = __dsqrt_rn(A), b = __dsqrt_rn(B);
  return fabs(__dsub_rn(__ddiv_rn(b, a), s)) <= __dmul_rn(beta, __ddiv_rn(1.0, a));
}

// Fast verdict, FP64 pipe: 6 sub + 6 mul + 4 add (A, B in reference rounding) + S1, H, AB, F = 20 instructions.
// Everything else runs on the integer pipe:  in = sign(H) | sign(F);  ambiguous <=> |F| < 2^-24 * Sigma^2 judged on the
// biased exponents (eF < 2*max(e(S1), e(beta^2)) - 1047 + shift), or F not finite / subnormal.
// Rigor (DESIGN.md "graph filter"): not ambiguous => |F| > 2^-28 * Sigma^2, seven orders of magnitude above both the
// rounding error of F and the reference's own sqrt/sub rounding, so sign(F) decides exactly like the reference.
__device__ __forceinline__ void verdict_fast(double A, double B, const KernelParams& kp, bool& in, bool& amb) {
  const double S1 = A + B;
  const double H = fma(0.5, S1, -kp.half_beta2);  // D / 2,  D = A + B - beta^2
  const double AB = A * B;
  const double F = fma(H, H, -AB);                // (D^2 - 4AB) / 4
  const int hH = __double2hiint(H), hF = __double2hiint(F), hS = __double2hiint(S1);
  in = (hH | hF) < 0;
  const int eF = (hF >> 20) & 0x7ff;
  const int eS = (hS >> 20) & 0x7ff;
  const int thr = max(2 * max(eS, kp.e_beta2) - 1047 + kp.filter_shift, 1);
  amb = (eF < thr) | (eF == 0x7ff);
}

#ifndef B2R_ROW_UNROLL
#define B2R_ROW_UNROLL 8
#endif
constexpr int kRowUnroll = B2R_ROW_UNROLL;

struct TilePts {
  double2 a[kTile];  // (sx, sy)
  double2 b[kTile];  // (sz, dx)
  double2 c[kTile];  // (dy, dz)
};

// Computes one 128x128 tile: rows i0.., columns j0...  Thread t owns column j0+t.
// dir[t] (shared) receives the 4 words of row i0+t for the column block; tw[0..3] (registers) the 4 words of
// row j0+t for the row block (transposed tile).
// kInterior: every row and column of the tile is < n and the tile is off the diagonal -> no validity tests at all.
// kMode: 0 sqrt-free filter + exact fallback, 1 exact only, 2 scale-inlier verdict (estimate_scaling=True).
template <int kMode, bool kInterior>
__device__ __forceinline__ void compute_tile(const double* __restrict__ src, const double* __restrict__ dst, int n,
                                             int i0, int j0, const KernelParams& kp, TilePts& sp,
                                             uint32_t (*dir)[4], uint32_t tw[4], unsigned& slow, double s_hat = 1.0) {
  constexpr bool kExactOnly = kMode != 0;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  {  // stage the row points (coalesced 24-byte AoS reads), zero-fill past n
    const int i = i0 + t;
    double v[6] = {0, 0, 0, 0, 0, 0};
    if (kInterior || i < n) {
      v[0] = src[3 * (size_t)i]; v[1] = src[3 * (size_t)i + 1]; v[2] = src[3 * (size_t)i + 2];
      v[3] = dst[3 * (size_t)i]; v[4] = dst[3 * (size_t)i + 1]; v[5] = dst[3 * (size_t)i + 2];
    }
    sp.a[t] = make_double2(v[0], v[1]);
    sp.b[t] = make_double2(v[2], v[3]);
    sp.c[t] = make_double2(v[4], v[5]);
  }
  const int j = j0 + t;
  double xj = 0, yj = 0, zj = 0, uj = 0, vj = 0, wj = 0;
  const bool jok = kInterior || j < n;
  if (jok) {
    xj = src[3 * (size_t)j]; yj = src[3 * (size_t)j + 1]; zj = src[3 * (size_t)j + 2];
    uj = dst[3 * (size_t)j]; vj = dst[3 * (size_t)j + 1]; wj = dst[3 * (size_t)j + 2];
  }
  __syncthreads();
  // The hot loop must stay inside the instruction cache: the chunk loop is NOT unrolled and the row loop only
  // kRowUnroll times (ncu on the fully unrolled version: 'no_instruction' was the top stall reason).
#pragma unroll 1
  for (int c = 0; c < 4; ++c) {
    uint32_t tword = 0;
    bool any_amb = false;
    if (!kExactOnly) {
#pragma unroll kRowUnroll
      for (int k = 0; k < 32; ++k) {
        const int il = 32 * c + k;
        const double2 pa = sp.a[il], pb = sp.b[il], pc = sp.c[il];
        const double A = sq3_ref(__dsub_rn(xj, pa.x), __dsub_rn(yj, pa.y), __dsub_rn(zj, pb.x));
        const double B = sq3_ref(__dsub_rn(uj, pb.y), __dsub_rn(vj, pc.x), __dsub_rn(wj, pc.y));
        bool in, amb;
        verdict_fast(A, B, kp, in, amb);
        if (!kInterior) {
          const int i = i0 + il;
          const bool ok = jok && (i < n) && (i != j);
          in = in && ok; amb = amb && ok;
        }
        const uint32_t bal = __ballot_sync(0xffffffffu, in);
        if (lane == 0) dir[il][warp] = bal;
        tword |= (in ? 1u : 0u) << k;
        any_amb |= amb;
      }
    }
    if (kExactOnly || __any_sync(0xffffffffu, any_amb)) {
      // rare (about 1e-4 of the chunks): redo the 32 rows with the reference's two square roots
      tword = 0;
      if (!kExactOnly) ++slow;
#pragma unroll 1
      for (int k = 0; k < 32; ++k) {
        const int il = 32 * c + k;
        const double2 pa = sp.a[il], pb = sp.b[il], pc = sp.c[il];
        const double A = sq3_ref(__dsub_rn(xj, pa.x), __dsub_rn(yj, pa.y), __dsub_rn(zj, pb.x));
        const double B = sq3_ref(__dsub_rn(uj, pb.y), __dsub_rn(vj, pc.x), __dsub_rn(wj, pc.y));
        const int i = i0 + il;
        bool in = kMode == 2 ? verdict_scale(A, B, kp.beta, s_hat) : verdict_exact(A, B, kp.beta);
        in = in && jok && (i < n) && (i != j);
        const uint32_t bal = __ballot_sync(0xffffffffu, in);
        if (lane == 0) dir[il][warp] = bal;
        tword |= (in ? 1u : 0u) << k;
      }
    }
    // tw[c] = tword without dynamic register indexing
    tw[0] = c == 0 ? tword : tw[0]; tw[1] = c == 1 ? tword : tw[1];
    tw[2] = c == 2 ? tword : tw[2]; tw[3] = c == 3 ? tword : tw[3];
  }
  __syncthreads();
}


// ------------------------------------------------------------------------------------------------------------------
// v2 of the filter kernel (mode 0 only): one CTA owns 128 rows x TWO adjacent 128-column blocks, every thread two
// columns, so each broadcast LDS of a row point feeds two pairs.  Inside the filter A and B are formed with FMA
// (3 FP64 instructions each instead of 5): |A' - A_ref| <= 6u*A moves the real verdict function by at most
// 3.3e-16 (sqrt(A)+sqrt(B)), which the ambiguity band absorbs with six orders of magnitude to spare (DESIGN.md "graph
// filter"); ambiguous chunks are redone with the reference's exact roundings, as before.  The band test runs on the
// high words only: |F| < 2^-22 * max(H^2, beta^4) judged on biased exponents, 5 integer instructions per pair.
// Per pair: 16 FP64-pipe instructions (was 20) and ~11 others (was ~25): the FP64 pipe is the only bound left.
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void verdict_fast2(double dx, double dy, double dz, double du, double dv, double dw, const KernelParams& kp,
                                              bool& in, bool& amb) {
  const double A = fma(dz, dz, fma(dy, dy, dx * dx));
  const double B = fma(dw, dw, fma(dv, dv, du * du));
  const double S1 = A + B;
  const double H = fma(0.5, S1, -kp.half_beta2);
  const double AB = A * B;
  const double F = fma(H, H, -AB);
  const unsigned hH = (unsigned)__double2hiint(H), hF = (unsigned)__double2hiint(F);
  in = (int)(hH | hF) < 0;
  const unsigned thr = max(hH << 1, kp.band_k2) - kp.band_k1;   // ((max(2 e(H), e(beta^4) + 1023) - 1045 + shift) << 20)
  amb = (hF & 0x7fffffffu) < thr;
}

template <bool kInterior>
__device__ __forceinline__ void compute_tile2(const double* __restrict__ src, const double* __restrict__ dst, int n, int i0, int j0,
                                              const KernelParams& kp, TilePts& sp, uint32_t (*dir)[8], uint32_t tw[8], unsigned& slow) {
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  {
    const int i = i0 + t;
    double v[6] = {0, 0, 0, 0, 0, 0};
    if (kInterior || i < n) {
      v[0] = src[3 * (size_t)i]; v[1] = src[3 * (size_t)i + 1]; v[2] = src[3 * (size_t)i + 2];
      v[3] = dst[3 * (size_t)i]; v[4] = dst[3 * (size_t)i + 1]; v[5] = dst[3 * (size_t)i + 2];
    }
    sp.a[t] = make_double2(v[0], v[1]);
    sp.b[t] = make_double2(v[2], v[3]);
    sp.c[t] = make_double2(v[4], v[5]);
  }
  const int ja = j0 + t, jb = j0 + kTile + t;
  const bool oka = kInterior || ja < n, okb = kInterior || jb < n;
  double xa = 0, ya = 0, za = 0, ua = 0, va = 0, wa = 0, xb = 0, yb = 0, zb = 0, ub = 0, vb = 0, wb = 0;
  if (oka) {
    xa = src[3 * (size_t)ja]; ya = src[3 * (size_t)ja + 1]; za = src[3 * (size_t)ja + 2];
    ua = dst[3 * (size_t)ja]; va = dst[3 * (size_t)ja + 1]; wa = dst[3 * (size_t)ja + 2];
  }
  if (okb) {
    xb = src[3 * (size_t)jb]; yb = src[3 * (size_t)jb + 1]; zb = src[3 * (size_t)jb + 2];
    ub = dst[3 * (size_t)jb]; vb = dst[3 * (size_t)jb + 1]; wb = dst[3 * (size_t)jb + 2];
  }
  __syncthreads();
#pragma unroll 1
  for (int c = 0; c < 4; ++c) {
    uint32_t worda = 0, wordb = 0;
    bool any_amb = false;
#pragma unroll 4
    for (int k = 0; k < 32; ++k) {
      const int il = 32 * c + k;
      const double2 pa = sp.a[il], pb = sp.b[il], pc = sp.c[il];
      bool ina, amba, inb, ambb;
      verdict_fast2(xa - pa.x, ya - pa.y, za - pb.x, ua - pb.y, va - pc.x, wa - pc.y, kp, ina, amba);
      verdict_fast2(xb - pa.x, yb - pa.y, zb - pb.x, ub - pb.y, vb - pc.x, wb - pc.y, kp, inb, ambb);
      if (!kInterior) {
        const int i = i0 + il;
        const bool ra = oka && (i < n) && (i != ja), rb = okb && (i < n) && (i != jb);
        ina = ina && ra; amba = amba && ra; inb = inb && rb; ambb = ambb && rb;
      }
      const uint32_t bala = __ballot_sync(0xffffffffu, ina), balb = __ballot_sync(0xffffffffu, inb);
      if (lane == 0) { dir[il][warp] = bala; dir[il][4 + warp] = balb; }
      worda |= (ina ? 1u : 0u) << k;
      wordb |= (inb ? 1u : 0u) << k;
      any_amb |= amba | ambb;
    }
    if (__any_sync(0xffffffffu, any_amb)) {  // rare: redo the 32 rows with the reference's roundings and square roots
      worda = wordb = 0;
      ++slow;
#pragma unroll 1
      for (int k = 0; k < 32; ++k) {
        const int il = 32 * c + k;
        const double2 pa = sp.a[il], pb = sp.b[il], pc = sp.c[il];
        const int i = i0 + il;
        const double Aa = sq3_ref(__dsub_rn(xa, pa.x), __dsub_rn(ya, pa.y), __dsub_rn(za, pb.x));
        const double Ba = sq3_ref(__dsub_rn(ua, pb.y), __dsub_rn(va, pc.x), __dsub_rn(wa, pc.y));
        const double Ab = sq3_ref(__dsub_rn(xb, pa.x), __dsub_rn(yb, pa.y), __dsub_rn(zb, pb.x));
        const double Bb = sq3_ref(__dsub_rn(ub, pb.y), __dsub_rn(vb, pc.x), __dsub_rn(wb, pc.y));
        const bool ina = verdict_exact(Aa, Ba, kp.beta) && oka && (i < n) && (i != ja);
        const bool inb = verdict_exact(Ab, Bb, kp.beta) && okb && (i < n) && (i != jb);
        const uint32_t bala = __ballot_sync(0xffffffffu, ina), balb = __ballot_sync(0xffffffffu, inb);
        if (lane == 0) { dir[il][warp] = bala; dir[il][4 + warp] = balb; }
        worda |= (ina ? 1u : 0u) << k;
        wordb |= (inb ? 1u : 0u) << k;
      }
    }
    tw[0] = c == 0 ? worda : tw[0]; tw[1] = c == 1 ? worda : tw[1]; tw[2] = c == 2 ? worda : tw[2]; tw[3] = c == 3 ? worda : tw[3];
    tw[4] = c == 0 ? wordb : tw[4]; tw[5] = c == 1 ? wordb : tw[5]; tw[6] = c == 2 ? wordb : tw[6]; tw[7] = c == 3 ? wordb : tw[7];
  }
  __syncthreads();
}

// work items of one problem: row block bi with the column-block pairs (bi + 2k, bi + 2k + 1), k = 0 .. ceil((T - bi)/2) - 1
__host__ __device__ inline int graph2_items(int T) { return (T / 2) * (T / 2 + 1) + ((T & 1) ? (T + 1) / 2 : 0); }

__global__ void __launch_bounds__(kTile) tim_graph2_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  const ProbDesc d = descs[blockIdx.y];
  const int n = d.n;
  const int T = (n + kTile - 1) / kTile;
  int p = blockIdx.x;
  int bi = 0, rowlen = (T + 1) / 2;
  while (bi < T && p >= rowlen) { p -= rowlen; ++bi; rowlen = (T - bi + 1) / 2; }
  if (bi >= T) return;
  const int bj = bi + 2 * p;
  __shared__ TilePts sp;
  __shared__ __align__(16) uint32_t dir[kTile][8];
  uint32_t tw[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
  unsigned slow = 0;
  if (bj != bi && (bj + 2) * kTile <= n) compute_tile2<true>(d.src, d.dst, n, bi * kTile, bj * kTile, kp, sp, dir, tw, slow);
  else compute_tile2<false>(d.src, d.dst, n, bi * kTile, bj * kTile, kp, sp, dir, tw, slow);
  const int t = threadIdx.x;
  const int i = bi * kTile + t;
  if (i < n) {
    uint32_t* row = d.adj + (size_t)i * d.pitch + 4 * bj;
    *reinterpret_cast<uint4*>(row) = *reinterpret_cast<const uint4*>(&dir[t][0]);
    if (bj + 1 < T) *reinterpret_cast<uint4*>(row + 4) = *reinterpret_cast<const uint4*>(&dir[t][4]);
  }
  if (bj != bi) {
    const int j = bj * kTile + t;
    if (j < n) *reinterpret_cast<uint4*>(d.adj + (size_t)j * d.pitch + 4 * bi) = make_uint4(tw[0], tw[1], tw[2], tw[3]);
  }
  if (bj + 1 < T) {
    const int j = (bj + 1) * kTile + t;
    if (j < n) *reinterpret_cast<uint4*>(d.adj + (size_t)j * d.pitch + 4 * bi) = make_uint4(tw[4], tw[5], tw[6], tw[7]);
  }
  if (slow && (threadIdx.x & 31) == 0) atomicAdd(&d.st->slow_pairs, 2048ull * slow);
}

// Batched, symmetric: grid = (T(T+1)/2 for the largest problem, batch).
template <int kMode>
__global__ void __launch_bounds__(kTile) tim_graph_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  constexpr bool kExactOnly = kMode != 0;
  const ProbDesc d = descs[blockIdx.y];
  const int n = d.n;
  const int T = (n + kTile - 1) / kTile;
  int p = blockIdx.x;
  if (p >= T * (T + 1) / 2) return;
  int ti = 0, rowlen = T;
  while (p >= rowlen) { p -= rowlen; ++ti; --rowlen; }
  const int tj = ti + p;
  __shared__ TilePts sp;
  __shared__ uint32_t dir[kTile][4];
  uint32_t tw[4] = {0u, 0u, 0u, 0u};
  unsigned slow = 0;
  const double s_hat = kMode == 2 ? d.st->scale : 1.0;
  if (ti != tj && (tj + 1) * kTile <= n) compute_tile<kMode, true>(d.src, d.dst, n, ti * kTile, tj * kTile, kp, sp, dir, tw, slow, s_hat);
  else compute_tile<kMode, false>(d.src, d.dst, n, ti * kTile, tj * kTile, kp, sp, dir, tw, slow, s_hat);
  const int t = threadIdx.x;
  const int i = ti * kTile + t;
  if (i < n) *reinterpret_cast<uint4*>(d.adj + (size_t)i * d.pitch + 4 * tj) = *reinterpret_cast<const uint4*>(dir[t]);
  if (ti != tj) {
    const int j = tj * kTile + t;
    if (j < n) *reinterpret_cast<uint4*>(d.adj + (size_t)j * d.pitch + 4 * ti) = make_uint4(tw[0], tw[1], tw[2], tw[3]);
  }
  if (!kExactOnly && slow && (threadIdx.x & 31) == 0) atomicAdd(&d.st->slow_pairs, 1024ull * slow);
}

// Row-sharded variant (config 5): full rows [row_begin,row_end), every column tile, no symmetry trick, so a rank
// writes only rows it owns.  grid = (column tiles, row tiles).
template <bool kExactOnly>
__global__ void __launch_bounds__(kTile) tim_graph_rows_kernel(const double* __restrict__ src, const double* __restrict__ dst,
                                                               int n, int row_begin, int row_end, uint32_t* __restrict__ adj,
                                                               int pitch, KernelParams kp, unsigned long long* slow_pairs) {
  const int i0 = row_begin + blockIdx.y * kTile;
  const int tj = blockIdx.x;
  __shared__ TilePts sp;
  __shared__ uint32_t dir[kTile][4];
  uint32_t tw[4] = {0u, 0u, 0u, 0u};
  unsigned slow = 0;
  // rows beyond row_end are computed but not stored (n_eff trick keeps the ballot uniform)
  compute_tile<kExactOnly ? 1 : 0, false>(src, dst, n, i0, tj * kTile, kp, sp, dir, tw, slow);
  const int i = i0 + threadIdx.x;
  if (i < row_end && i < n)
    *reinterpret_cast<uint4*>(adj + (size_t)(i - row_begin) * pitch + 4 * tj) = *reinterpret_cast<const uint4*>(dir[threadIdx.x]);
  if (!kExactOnly && slow && slow_pairs && (threadIdx.x & 31) == 0) atomicAdd(slow_pairs, 1024ull * slow);
}

// Degrees: one warp per row (8 rows per warp, 64 rows per CTA), popcount over the row's words.  Per problem it also
// leaves max degree, 2*edges and -- for the clique heuristic -- the highest-degree vertex of each id class (v & 7).
// Same-address atomics are reduced per CTA first (they serialise in L2 otherwise).
constexpr int kDegRowsPerBlock = 64;
__global__ void __launch_bounds__(256) degree_kernel(const ProbDesc* __restrict__ descs) {
  const ProbDesc d = descs[blockIdx.y];
  __shared__ unsigned long long s_key[8];
  __shared__ unsigned long long s_sum;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int v0 = blockIdx.x * kDegRowsPerBlock;
  if (v0 >= d.n) return;
  if (threadIdx.x < 8) s_key[threadIdx.x] = 0ull;
  if (threadIdx.x == 8) s_sum = 0ull;
  __syncthreads();
  const int W = 4 * ((d.n + kTile - 1) / kTile);
  unsigned long long wsum = 0;
#pragma unroll 2
  for (int k = 0; k < 8; ++k) {
    const int v = v0 + warp * 8 + k;
    if (v >= d.n) break;
    const uint32_t* row = d.adj + (size_t)v * d.pitch;
    int c = 0;
    for (int w = lane; w < W; w += 32) c += __popc(row[w]);
    c = __reduce_add_sync(0xffffffffu, c);
    if (lane == 0) {
      d.deg[v] = c;
      wsum += (unsigned long long)c;
      atomicMax(&s_key[v & 7], ((unsigned long long)c << 32) | (unsigned long long)(0xffffffffu - (unsigned)v));
    }
  }
  if (lane == 0) atomicAdd(&s_sum, wsum);
  __syncthreads();
  if (threadIdx.x < 8) {
    if (s_key[threadIdx.x]) atomicMax(&d.st->seed_key[threadIdx.x], s_key[threadIdx.x]);
  } else if (threadIdx.x == 8) {
    atomicAdd(&d.st->deg_sum, s_sum);
    unsigned long long m = 0;
    for (int k = 0; k < 8; ++k) m = s_key[k] > m ? s_key[k] : m;
    atomicMax(&d.st->max_deg, (int)(m >> 32));
  }
}

int launch_tim_graph(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, cudaStream_t st) {
  const int T = (n_max + kTile - 1) / kTile;
  dim3 grid(T * (T + 1) / 2, batch);
  if (kp.estimate_scaling) tim_graph_kernel<2><<<grid, kTile, 0, st>>>(d_desc, kp);
  else if (kp.exact_only) tim_graph_kernel<1><<<grid, kTile, 0, st>>>(d_desc, kp);
  else if (kp.graph_v1) tim_graph_kernel<0><<<grid, kTile, 0, st>>>(d_desc, kp);
  else tim_graph2_kernel<<<dim3(graph2_items(T), batch), kTile, 0, st>>>(d_desc, kp);
  return 1;
}

int launch_tim_graph_rows(const double* src, const double* dst, int n, int row_begin, int row_end, uint32_t* adj,
                          int pitch, const KernelParams& kp, unsigned long long* slow_pairs, cudaStream_t st) {
  const int T = (n + kTile - 1) / kTile;
  const int rows = row_end - row_begin;
  if (rows <= 0) return 0;
  dim3 grid(T, (rows + kTile - 1) / kTile);
  if (kp.exact_only) tim_graph_rows_kernel<true><<<grid, kTile, 0, st>>>(src, dst, n, row_begin, row_end, adj, pitch, kp, slow_pairs);
  else tim_graph_rows_kernel<false><<<grid, kTile, 0, st>>>(src, dst, n, row_begin, row_end, adj, pitch, kp, slow_pairs);
  return 1;
}

int launch_degree(const ProbDesc* d_desc, int batch, int n_max, cudaStream_t st) {
  dim3 grid((n_max + kDegRowsPerBlock - 1) / kDegRowsPerBlock, batch);
  degree_kernel<<<grid, 256, 0, st>>>(d_desc);
  return 1;
}

}  // namespace b2r
