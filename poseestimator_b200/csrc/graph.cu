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
#include <math_constants.h>

#include "common.cuh"
#include "exact.cuh"

namespace b2r {

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


// ------------------------------------------------------------------------------------------------------------------
// v3 of the filter kernel (mode 0 only): the same 128 x 256 tile and sign test, evaluated in PACKED FP32.
//
// v2 is bound by the FP64 pipe (64 lanes per SM per clock, 16 instructions per pair).  The verdict of almost every pair
// is decided by a wide margin, so v3 evaluates the sign test in fp32 on the FMA pipe with `*.f32x2` instructions (FADD2 /
// FMUL2 / FFMA2: the two columns of a thread travel in one 64-bit register pair, the row value is a scalar broadcast
// operand) and keeps fp64 only for the ambiguous 32-row chunks.  Per TWO pairs: 6 FADD2 (differences), 6 FMUL2/FFMA2 (A, B),
//     Dm = A - B,  S1 = A + B,  L = fma(-2 beta^2, S1, beta^4),  Dm2 = Dm * Dm,  G = Dm2 + L     (G = D^2 - 4AB, D = S1 - beta^2)
//     thr = fma(Dm2, a3, fma(S1, fma(S1, a2, a1), a0))                                            (error band, below)
// = 20 packed instructions, then per pair  in = (S1 < beta^2) | (G < 0),  ambiguous = !(|G| >= thr).
//
// Rigor (DESIGN.md "graph filter v3").  The fp32 copies are taken relative to the centre of each cloud's bounding box
// (graph_prep_kernel); M = largest |centred coordinate|, u = 2^-24, e = 2.001 u M bounds the absolute error of a computed
// coordinate difference (two conversions + one subtraction; its relative part is u).  Then
//   |A~ - A| <= 3.4642 e a + 6.00001 e^2 + 5.001 u A <= 8.47 u A + 3.47 u M^2      (a = sqrt(A); e a <= (e/2M) A + e M/2)
// so the errors of A~ and B~ together are below  E = alpha S1~ + gamma,  alpha = 8.5 u, gamma = 7 u M^2.  Propagating E and
// every fp32 rounding through Dm, Dm2, L and G gives
//   |G~ - G| <= 4.02 u Dm~^2 + 2 E |Dm~| + 2 E^2 + 2.95 beta^2 E + 3.02 u beta^4
//            <= (4.02 u + 1/k) Dm~^2 + (2 + k) E^2 + 2.95 beta^2 E + 3.02 u beta^4            (2 E |Dm| <= k E^2 + Dm^2 / k, any k > 0)
// and thr is twice that bound with (4 + k) in place of (2 + k), a polynomial in S1~ and Dm2~ with positive coefficients
// (rounded up).  A pair outside the band has |G| >= 0.49 thr, which (i) fixes the sign of G, (ii) when G > 0 puts
// |S1 - beta^2| >= 0.7 sqrt(thr) > 1.12 E + u beta^2, more than the error of the computed S1~ < beta^2 test, and
// (iii) exceeds 1e-12 (S1 + beta^2)^2, three hundred times the 3.7e-15 (S1 + beta^2)^2 within which the reference's own
// fp64 roundings (two square roots, one subtraction, the roundings of A and B) could flip the verdict.  k is chosen per
// problem to balance the two AM-GM terms at a typical pair on the decision boundary; any k keeps the bound valid.
// Problems with non-finite coordinates, M >= 1e8, beta outside [1e-6, 1e8] or a test-widened band >= 2^-6 run the fp64
// filter (v2) instead; NaN in G or thr reads as ambiguous.
// ------------------------------------------------------------------------------------------------------------------
#define B2R_F2_OP2(name, ptx)                                                                                          \
  __device__ __forceinline__ float2 name(float2 a, float2 b) {                                                        \
    float2 r;                                                                                                          \
    asm("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; " ptx " rc, ra, rb; mov.b64 {%0,%1}, rc;}"  \
        : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));                                             \
    return r;                                                                                                          \
  }
B2R_F2_OP2(f2add, "add.rn.f32x2")
B2R_F2_OP2(f2sub, "sub.rn.f32x2")
B2R_F2_OP2(f2mul, "mul.rn.f32x2")
__device__ __forceinline__ float2 f2fma(float2 a, float2 b, float2 c) {
  float2 r;
  asm("{.reg .b64 ra, rb, rc, rd; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; mov.b64 rc, {%6,%7}; fma.rn.f32x2 rd, ra, rb, rc; "
      "mov.b64 {%0,%1}, rd;}"
      : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
  return r;
}
__device__ __forceinline__ float2 f2bc(float v) { return make_float2(v, v); }
// w |= bit when p, as one predicated LOP3 (the compiler's own choice was SEL + SHF + LOP3)
__device__ __forceinline__ void set_bit_if(uint32_t& w, bool p, uint32_t bit) {
  asm("{.reg .pred q; setp.ne.u32 q, %1, 0; @q or.b32 %0, %0, %2;}" : "+r"(w) : "r"((uint32_t)p), "r"(bit));
}

#ifndef B2R_G3_UNROLL
#define B2R_G3_UNROLL 16
#endif
#ifndef B2R_G3_MINB
#define B2R_G3_MINB 4
#endif

constexpr int kG3Unroll = B2R_G3_UNROLL;

struct TilePts32 {
  float4 a[kTile];  // (sx, sy, sz, dx) relative to the cloud centres
  float2 b[kTile];  // (dy, dz)
};

// One CTA per problem: bounding boxes of src and dst -> centres, M, and the constants of the fp32 filter (ProbState).
__global__ void __launch_bounds__(256) graph_prep_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  const ProbDesc d = descs[blockIdx.x];
  const int n = d.n;
  double mn[6], mx[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) { mn[k] = CUDART_INF; mx[k] = -CUDART_INF; }
  bool bad = false;
  for (int i = threadIdx.x; i < n; i += 256) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double a = d.src[3 * (size_t)i + k], b = d.dst[3 * (size_t)i + k];
      bad |= !(fabs(a) <= 1.79e308) | !(fabs(b) <= 1.79e308);
      mn[k] = fmin(mn[k], a); mx[k] = fmax(mx[k], a);
      mn[3 + k] = fmin(mn[3 + k], b); mx[3 + k] = fmax(mx[3 + k], b);
    }
  }
  __shared__ double s_mn[8][6], s_mx[8][6];
  __shared__ int s_bad;
  if (threadIdx.x == 0) s_bad = 0;
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 6; ++k) {
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      mn[k] = fmin(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
      mx[k] = fmax(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
    }
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int k = 0; k < 6; ++k) { s_mn[threadIdx.x >> 5][k] = mn[k]; s_mx[threadIdx.x >> 5][k] = mx[k]; }
  }
  if (bad) atomicOr(&s_bad, 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    ProbState* st = d.st;
    double M = 0.0;
    double Mc[2] = {0.0, 0.0}, Pc[2] = {0.0, 0.0};   // per cloud: largest |centred coordinate|, bound of the largest squared norm
    for (int k = 0; k < 6; ++k) {
      double lo = s_mn[0][k], hi = s_mx[0][k];
      for (int w = 1; w < 8; ++w) { lo = fmin(lo, s_mn[w][k]); hi = fmax(hi, s_mx[w][k]); }
      const double c = 0.5 * lo + 0.5 * hi;
      st->ctr[k] = c;
      const double h = fmax(hi - c, c - lo) * (1.0 + 1e-6);   // covers the fp32 rounding of the centred copy
      M = fmax(M, h);
      Mc[k / 3] = fmax(Mc[k / 3], h);
      Pc[k / 3] += h * h;
    }
    const double u = 5.9604644775390625e-8;  // 2^-24
    const double beta = fabs(kp.beta), beta2 = beta * beta, beta4 = beta2 * beta2;
    const double tmin = kp.filter_tau_min;
    const bool ok = !s_bad && n > 0 && M < 1e8 && beta >= 1e-6 && beta <= 1e8 && tmin < 0.015625;
    st->use32 = ok ? 1 : 0;
    if (ok) {
      // v3 (coordinate differences): alpha = 8.5 u, gamma = 7 u M^2.  v4 (norm expansion, see compute_tile4): the error of
      // A~ and B~ together is below 4.47 u (A + B) + 3.47 u (Ms^2 + Md^2) + 12.004 u (Ps + Pd), P = largest squared norm of a
      // centred point (bounded by the squared half-diagonal of the bounding box).
      const double alpha = kp.graph_v3 ? 8.5 * u : 4.5 * u;
      const double gamma = kp.graph_v3 ? 7.0 * u * M * M : 1.0001 * u * (3.47 * (Mc[0] * Mc[0] + Mc[1] * Mc[1]) + 12.004 * (Pc[0] + Pc[1]));
      const double s_typ = 2.0 * M * M;                         // a typical S1; only steers k
      double k = beta * sqrt(2.0 * s_typ) / (alpha * s_typ + gamma);
      if (!(k >= 8.0)) k = 8.0;                                 // a3 < 1/2
      if (k > 1e12) k = 1e12;
      const double pad = 1.0 + 1e-5;                            // covers the fp32 evaluation of the band polynomial
      // band in (Dm^2, S1):  thr = a3 Dm^2 + a2 S1^2 + a1 S1 + a0;  with Dm^2 <= |G| (1 + 2u) - L it becomes a polynomial
      // in S1 alone that |G| is compared with:  |G| >= (q(S1) - a3 L) / (1 - a3 (1 + 2u))
      const double a3 = 2.0 * (4.03 * u + 1.0001 / k);
      const double a2 = 2.0 * (4.0 + k) * alpha * alpha + tmin;
      const double a1 = 2.0 * (2.0 * (4.0 + k) * alpha * gamma + 2.95 * beta2 * alpha) + 2.0 * tmin * beta2;
      const double a0 = 2.0 * ((4.0 + k) * gamma * gamma + 2.95 * beta2 * gamma + 3.02 * u * beta4) + tmin * beta4;
      const double den = 1.0 - a3 * (1.0 + 2.0 * u);
      st->f32c[0] = __double2float_rn(-2.0 * beta2);
      st->f32c[1] = __double2float_rn(beta4);
      st->f32c[2] = __double2float_rn(beta2);
      st->f32c[3] = __double2float_ru(a2 / den * pad);
      st->f32c[4] = __double2float_ru((a1 + a3 * 2.0 * beta2 * (1.0 + 3.0 * u)) / den * pad);
      st->f32c[5] = __double2float_ru((a0 - a3 * beta4 * (1.0 - 3.0 * u)) / den + 1.01 * u * beta4);
      st->f32c[6] = (float)a3;
      st->f32c[7] = (float)k;
    }
  }
}

// 32 x 32 bit transpose across a warp: lane r returns the word whose bit c is bit r of lane c's x (5 shuffle steps).
__device__ __forceinline__ uint32_t warp_transpose32(uint32_t x, int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const uint32_t m = s == 16 ? 0x0000ffffu : s == 8 ? 0x00ff00ffu : s == 4 ? 0x0f0f0f0fu : s == 2 ? 0x33333333u : 0x55555555u;
    const uint32_t y = __shfl_xor_sync(0xffffffffu, x, s);
    x = (lane & s) ? ((x & ~m) | ((y & ~m) >> s)) : ((x & m) | ((y & m) << s));
  }
  return x;
}

// dirT[w][r]: word w (column block a: 0..3, b: 4..7 by warp) of tile row r.  The hot loop has no warp-synchronous
// instruction (a ballot per row split it into one basic block per row and serialised the rows): every thread collects the
// verdict bits of its two columns over 32 rows, then the warp transposes the 32 x 32 bit block with shuffles.
template <bool kInterior>
__device__ __forceinline__ void compute_tile3(const double* __restrict__ src, const double* __restrict__ dst, int n, int i0, int j0,
                                              const ProbState* __restrict__ st, const KernelParams& kp, TilePts32& sp,
                                              uint32_t (*dirT)[kTile], uint32_t tw[8], unsigned& slow) {
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const double c0 = st->ctr[0], c1 = st->ctr[1], c2 = st->ctr[2], c3 = st->ctr[3], c4 = st->ctr[4], c5 = st->ctr[5];
  {
    const int i = i0 + t;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    float2 b = make_float2(0.f, 0.f);
    if (kInterior || i < n) {
      a.x = __double2float_rn(src[3 * (size_t)i] - c0); a.y = __double2float_rn(src[3 * (size_t)i + 1] - c1);
      a.z = __double2float_rn(src[3 * (size_t)i + 2] - c2); a.w = __double2float_rn(dst[3 * (size_t)i] - c3);
      b.x = __double2float_rn(dst[3 * (size_t)i + 1] - c4); b.y = __double2float_rn(dst[3 * (size_t)i + 2] - c5);
    }
    sp.a[t] = a;
    sp.b[t] = b;
  }
  const int ja = j0 + t, jb = j0 + kTile + t;
  const bool oka = kInterior || ja < n, okb = kInterior || jb < n;
  float2 X = f2bc(0.f), Y = X, Z = X, U = X, V = X, W = X;
  if (oka) {
    X.x = __double2float_rn(src[3 * (size_t)ja] - c0); Y.x = __double2float_rn(src[3 * (size_t)ja + 1] - c1);
    Z.x = __double2float_rn(src[3 * (size_t)ja + 2] - c2); U.x = __double2float_rn(dst[3 * (size_t)ja] - c3);
    V.x = __double2float_rn(dst[3 * (size_t)ja + 1] - c4); W.x = __double2float_rn(dst[3 * (size_t)ja + 2] - c5);
  }
  if (okb) {
    X.y = __double2float_rn(src[3 * (size_t)jb] - c0); Y.y = __double2float_rn(src[3 * (size_t)jb + 1] - c1);
    Z.y = __double2float_rn(src[3 * (size_t)jb + 2] - c2); U.y = __double2float_rn(dst[3 * (size_t)jb] - c3);
    V.y = __double2float_rn(dst[3 * (size_t)jb + 1] - c4); W.y = __double2float_rn(dst[3 * (size_t)jb + 2] - c5);
  }
  const float k1 = st->f32c[0], k0 = st->f32c[1], b2 = st->f32c[2], q2 = st->f32c[3], q1 = st->f32c[4], q0 = st->f32c[5];
  __syncthreads();
#pragma unroll 1
  for (int c = 0; c < 4; ++c) {
    uint32_t worda = 0, wordb = 0, ambwa = 0, ambwb = 0;
    bool any_amb = false;
#pragma unroll kG3Unroll
    for (int k = 0; k < 32; ++k) {
      const int il = 32 * c + k;
      const float4 pa = sp.a[il];
      const float2 pb = sp.b[il];
      const float2 d0 = f2sub(X, f2bc(pa.x)), d1 = f2sub(Y, f2bc(pa.y)), d2 = f2sub(Z, f2bc(pa.z));
      const float2 d3 = f2sub(U, f2bc(pa.w)), d4 = f2sub(V, f2bc(pb.x)), d5 = f2sub(W, f2bc(pb.y));
      const float2 A = f2fma(d2, d2, f2fma(d1, d1, f2mul(d0, d0)));
      const float2 B = f2fma(d5, d5, f2fma(d4, d4, f2mul(d3, d3)));
      const float2 Dm = f2sub(A, B), S1 = f2add(A, B);
      const float2 G = f2fma(Dm, Dm, f2fma(f2bc(k1), S1, f2bc(k0)));
      const float2 T = f2fma(S1, f2fma(S1, f2bc(q2), f2bc(q1)), f2bc(q0));
      const bool ina = (S1.x < b2) | (G.x < 0.f), inb = (S1.y < b2) | (G.y < 0.f);
      const bool amba = !(fabsf(G.x) >= T.x), ambb = !(fabsf(G.y) >= T.y);
      set_bit_if(worda, ina, 1u << k);
      set_bit_if(wordb, inb, 1u << k);
      if (kInterior) any_amb |= amba | ambb;
      else { set_bit_if(ambwa, amba, 1u << k); set_bit_if(ambwb, ambb, 1u << k); }
    }
    uint32_t maska = 0xffffffffu, maskb = 0xffffffffu;
    if (!kInterior) {  // rows past n, the diagonal and columns past n carry no pair
      const int r0 = i0 + 32 * c;
      const uint32_t rows = n - r0 >= 32 ? 0xffffffffu : (n - r0 <= 0 ? 0u : ((1u << (n - r0)) - 1u));
      maska = oka ? rows : 0u;
      maskb = okb ? rows : 0u;
      if (ja >= r0 && ja < r0 + 32) maska &= ~(1u << (ja - r0));
      if (jb >= r0 && jb < r0 + 32) maskb &= ~(1u << (jb - r0));
      worda &= maska; wordb &= maskb;
      any_amb = ((ambwa & maska) | (ambwb & maskb)) != 0u;
    }
    if (__any_sync(0xffffffffu, any_amb)) {  // rare: redo the 32 rows in fp64 with the reference's roundings and square roots
      worda = wordb = 0;
      ++slow;
      double xa = 0, ya = 0, za = 0, ua = 0, va = 0, wa = 0, xb = 0, yb = 0, zb = 0, ub = 0, vb = 0, wb = 0;
      if (oka) {
        xa = src[3 * (size_t)ja]; ya = src[3 * (size_t)ja + 1]; za = src[3 * (size_t)ja + 2];
        ua = dst[3 * (size_t)ja]; va = dst[3 * (size_t)ja + 1]; wa = dst[3 * (size_t)ja + 2];
      }
      if (okb) {
        xb = src[3 * (size_t)jb]; yb = src[3 * (size_t)jb + 1]; zb = src[3 * (size_t)jb + 2];
        ub = dst[3 * (size_t)jb]; vb = dst[3 * (size_t)jb + 1]; wb = dst[3 * (size_t)jb + 2];
      }
#pragma unroll 1
      for (int k = 0; k < 32; ++k) {
        const int i = i0 + 32 * c + k;
        const size_t ii = (kInterior || i < n) ? (size_t)i : 0;
        const double px = src[3 * ii], py = src[3 * ii + 1], pz = src[3 * ii + 2];
        const double qx = dst[3 * ii], qy = dst[3 * ii + 1], qz = dst[3 * ii + 2];
        const double Aa = sq3_ref(__dsub_rn(xa, px), __dsub_rn(ya, py), __dsub_rn(za, pz));
        const double Ba = sq3_ref(__dsub_rn(ua, qx), __dsub_rn(va, qy), __dsub_rn(wa, qz));
        const double Ab = sq3_ref(__dsub_rn(xb, px), __dsub_rn(yb, py), __dsub_rn(zb, pz));
        const double Bb = sq3_ref(__dsub_rn(ub, qx), __dsub_rn(vb, qy), __dsub_rn(wb, qz));
        worda |= (verdict_exact(Aa, Ba, kp.beta) ? 1u : 0u) << k;
        wordb |= (verdict_exact(Ab, Bb, kp.beta) ? 1u : 0u) << k;
      }
      worda &= maska; wordb &= maskb;
    }
    dirT[warp][32 * c + lane] = warp_transpose32(worda, lane);
    dirT[4 + warp][32 * c + lane] = warp_transpose32(wordb, lane);
    tw[0] = c == 0 ? worda : tw[0]; tw[1] = c == 1 ? worda : tw[1]; tw[2] = c == 2 ? worda : tw[2]; tw[3] = c == 3 ? worda : tw[3];
    tw[4] = c == 0 ? wordb : tw[4]; tw[5] = c == 1 ? wordb : tw[5]; tw[6] = c == 2 ? wordb : tw[6]; tw[7] = c == 3 ? wordb : tw[7];
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kTile, B2R_G3_MINB) tim_graph3_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  const ProbDesc d = descs[blockIdx.y];
  const int n = d.n;
  const int T = (n + kTile - 1) / kTile;
  int p = blockIdx.x;
  int bi = 0, rowlen = (T + 1) / 2;
  while (bi < T && p >= rowlen) { p -= rowlen; ++bi; rowlen = (T - bi + 1) / 2; }
  if (bi >= T) return;
  const int bj = bi + 2 * p;
  __shared__ union { TilePts p64; TilePts32 p32; } sp;
  __shared__ __align__(16) union { uint32_t rows[kTile][8]; uint32_t cols[8][kTile]; } dir;
  uint32_t tw[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
  unsigned slow = 0;
  const int t = threadIdx.x;
  uint4 lo, hi;
  if (d.st->use32) {
    if (bj != bi && (bj + 2) * kTile <= n) compute_tile3<true>(d.src, d.dst, n, bi * kTile, bj * kTile, d.st, kp, sp.p32, dir.cols, tw, slow);
    else compute_tile3<false>(d.src, d.dst, n, bi * kTile, bj * kTile, d.st, kp, sp.p32, dir.cols, tw, slow);
    lo = make_uint4(dir.cols[0][t], dir.cols[1][t], dir.cols[2][t], dir.cols[3][t]);
    hi = make_uint4(dir.cols[4][t], dir.cols[5][t], dir.cols[6][t], dir.cols[7][t]);
  } else {
    compute_tile2<false>(d.src, d.dst, n, bi * kTile, bj * kTile, kp, sp.p64, dir.rows, tw, slow);
    lo = *reinterpret_cast<const uint4*>(&dir.rows[t][0]);
    hi = *reinterpret_cast<const uint4*>(&dir.rows[t][4]);
  }
  const int i = bi * kTile + t;
  if (i < n) {
    uint32_t* row = d.adj + (size_t)i * d.pitch + 4 * bj;
    *reinterpret_cast<uint4*>(row) = lo;
    if (bj + 1 < T) *reinterpret_cast<uint4*>(row + 4) = hi;
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

// ------------------------------------------------------------------------------------------------------------------
// v4 of the filter kernel (default): v3's packed-fp32 sign test with the squared lengths formed by NORM EXPANSION,
//     A~ = fma(Z, -2 pz, fma(Y, -2 py, fma(X, -2 px, Hs_j + hs_i))),   Hs = fl32(|x^|^2) of the centred fp32 copy,
// 4 packed instructions instead of 6 (B likewise), then Dm, S1, L, G, thr(S1) as in v3 and Q = S1 - beta^2 whose sign
// replaces the S1 < beta^2 compare: 15 packed instructions per two pairs (v3: 18).  Per pair the integer pipe does
// LOP3 (G | Q), one funnel shift that pushes the sign bit into the column's 32-row word, and one FSETP that ORs
// !(|G| >= thr) into a per-chunk predicate (v3: ~10 integer-pipe instructions per two pairs more).
//
// Rigor.  x^ = fl32(x - c) per coordinate, |x^ - (x - c)| <= 1.0005 u M.  With A^ = |x^_j - x^_i|^2 (real arithmetic):
//   |A^ - A| <= 3.4642 e a + 3 e^2,  e = 2.001 u M  (as v3);
//   the computed chain returns A^ + (h_j - r_j) + (h_i - r_i) + eps0 + eps1 + eps2 + eps3  (r = |x^|^2, h = fl32(r) taken from
//   an fp64 evaluation, eps_m = the rounding of step m); every exact partial sum is  sum_{k<=m} d^_k^2 + sum_{k>m} (x^_jk^2 + x^_ik^2)
//   <= 2 (r_i + r_j), the last one is A^, so |A~ - A^| <= 6.002 u (r_i + r_j) + 1.0001 u A^ <= 12.004 u P + 1.0001 u A^.
// Hence |A~ - A| + |B~ - B| <= 4.47 u (A + B) + 3.47 u (Ms^2 + Md^2) + 12.004 u (Ps + Pd) = alpha S1 + gamma with alpha = 4.5 u
// (the slack pays for S1 -> S1~), and steps 3-5 of the v3 analysis (Dm, S1, L, G and the band polynomial) apply verbatim.
// sign(fl(S1~ - beta^2~)) is exact (one rounded subtraction never changes the sign), so Q < 0 <=> S1~ < beta^2~ as in v3.
//
// Ambiguous pairs (about 30 per 12.5 M at config 3) are fixed ONE BY ONE: a 32-row chunk with any ambiguous pair replays
// its rows (same instruction sequence, hence the same fp32 values), and only the flagged pairs are re-evaluated with the
// reference's fp64 roundings and square roots (v3 redid all 2048 pairs of the chunk in fp64).
// ------------------------------------------------------------------------------------------------------------------
#ifndef B2R_G4_UNROLL
#define B2R_G4_UNROLL 16
#endif
#ifndef B2R_G4_MINB
#define B2R_G4_MINB 5
#endif

constexpr int kG4Unroll = B2R_G4_UNROLL;

struct TilePts4 {
  float4 a[kTile];  // (-2 sx, -2 sy, -2 sz, |s|^2) of the centred fp32 copy of the row point
  float4 b[kTile];  // (-2 dx, -2 dy, -2 dz, |d|^2)
};

struct Cols4 { float2 X, Y, Z, Hs, U, V, W, Hd; };
struct Consts4 { float2 k1, k0, q2, q1, q0, nb2; };

// centred fp32 copy of point i of both clouds and the fp32 roundings of its squared norms (evaluated in fp64: exact products,
// sums rounded at 2^-53)
__device__ __forceinline__ void load_point4(const double* __restrict__ src, const double* __restrict__ dst, size_t i, const double c[6],
                                            float v[6], float& hs, float& hd) {
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    v[k] = __double2float_rn(src[3 * i + k] - c[k]);
    v[3 + k] = __double2float_rn(dst[3 * i + k] - c[3 + k]);
  }
  const double x = v[0], y = v[1], z = v[2], p = v[3], q = v[4], r = v[5];
  hs = __double2float_rn(fma(z, z, fma(y, y, x * x)));
  hd = __double2float_rn(fma(r, r, fma(q, q, p * p)));
}

__device__ __forceinline__ void eval4(const Cols4& cl, const Consts4& cs, const float4 ra, const float4 rb, float2& G, float2& S1, float2& T) {
  const float2 A = f2fma(cl.Z, f2bc(ra.z), f2fma(cl.Y, f2bc(ra.y), f2fma(cl.X, f2bc(ra.x), f2add(cl.Hs, f2bc(ra.w)))));
  const float2 B = f2fma(cl.W, f2bc(rb.z), f2fma(cl.V, f2bc(rb.y), f2fma(cl.U, f2bc(rb.x), f2add(cl.Hd, f2bc(rb.w)))));
  const float2 Dm = f2sub(A, B);
  S1 = f2add(A, B);
  G = f2fma(Dm, Dm, f2fma(cs.k1, S1, cs.k0));
  T = f2fma(S1, f2fma(S1, cs.q2, cs.q1), cs.q0);
}

template <bool kInterior>
__device__ __forceinline__ void compute_tile4(const double* __restrict__ src, const double* __restrict__ dst, int n, int i0, int j0,
                                              const ProbState* __restrict__ st, const KernelParams& kp, TilePts4& sp,
                                              uint32_t (*dirT)[kTile], uint32_t tw[8], unsigned& slow) {
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  double ctr[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) ctr[k] = st->ctr[k];
  {
    const int i = i0 + t;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
    if (kInterior || i < n) {
      float v[6], hs, hd;
      load_point4(src, dst, (size_t)i, ctr, v, hs, hd);
      a = make_float4(-2.f * v[0], -2.f * v[1], -2.f * v[2], hs);
      b = make_float4(-2.f * v[3], -2.f * v[4], -2.f * v[5], hd);
    }
    sp.a[t] = a;
    sp.b[t] = b;
  }
  const int ja = j0 + t, jb = j0 + kTile + t;
  const bool oka = kInterior || ja < n, okb = kInterior || jb < n;
  Cols4 cl;
  cl.X = cl.Y = cl.Z = cl.Hs = cl.U = cl.V = cl.W = cl.Hd = f2bc(0.f);
  if (oka) {
    float v[6], hs, hd;
    load_point4(src, dst, (size_t)ja, ctr, v, hs, hd);
    cl.X.x = v[0]; cl.Y.x = v[1]; cl.Z.x = v[2]; cl.U.x = v[3]; cl.V.x = v[4]; cl.W.x = v[5]; cl.Hs.x = hs; cl.Hd.x = hd;
  }
  if (okb) {
    float v[6], hs, hd;
    load_point4(src, dst, (size_t)jb, ctr, v, hs, hd);
    cl.X.y = v[0]; cl.Y.y = v[1]; cl.Z.y = v[2]; cl.U.y = v[3]; cl.V.y = v[4]; cl.W.y = v[5]; cl.Hs.y = hs; cl.Hd.y = hd;
  }
  Consts4 cs;
  cs.k1 = f2bc(st->f32c[0]); cs.k0 = f2bc(st->f32c[1]); cs.nb2 = f2bc(-st->f32c[2]);
  cs.q2 = f2bc(st->f32c[3]); cs.q1 = f2bc(st->f32c[4]); cs.q0 = f2bc(st->f32c[5]);
  __syncthreads();
#pragma unroll 1
  for (int c = 0; c < 4; ++c) {
    uint32_t worda = 0, wordb = 0;
    bool any_amb = false;
#pragma unroll kG4Unroll
    for (int k = 0; k < 32; ++k) {
      const int il = 32 * c + k;
      float2 G, S1, T;
      eval4(cl, cs, sp.a[il], sp.b[il], G, S1, T);
      const float2 Q = f2add(S1, cs.nb2);
      // row k lands in bit 31 - k: word = (word << 1) | sign(G | Q)
      worda = __funnelshift_l(__float_as_uint(G.x) | __float_as_uint(Q.x), worda, 1);
      wordb = __funnelshift_l(__float_as_uint(G.y) | __float_as_uint(Q.y), wordb, 1);
      any_amb |= !(fabsf(G.x) >= T.x) | !(fabsf(G.y) >= T.y);
    }
    worda = __brev(worda);
    wordb = __brev(wordb);
    if (__any_sync(0xffffffffu, any_amb)) {
      // rare: replay the chunk and re-evaluate only the flagged pairs with the reference's fp64 roundings and square roots
#pragma unroll 1
      for (int k = 0; k < 32; ++k) {
        const int il = 32 * c + k;
        float2 G, S1, T;
        eval4(cl, cs, sp.a[il], sp.b[il], G, S1, T);
        const bool amba = !(fabsf(G.x) >= T.x), ambb = !(fabsf(G.y) >= T.y);
        if (amba | ambb) {
          const int i = i0 + il;
          const size_t ii = (kInterior || i < n) ? (size_t)i : 0;
          const double px = src[3 * ii], py = src[3 * ii + 1], pz = src[3 * ii + 2];
          const double qx = dst[3 * ii], qy = dst[3 * ii + 1], qz = dst[3 * ii + 2];
          if (amba) {
            const size_t jj = oka ? (size_t)ja : 0;
            const double A = sq3_ref(__dsub_rn(src[3 * jj], px), __dsub_rn(src[3 * jj + 1], py), __dsub_rn(src[3 * jj + 2], pz));
            const double B = sq3_ref(__dsub_rn(dst[3 * jj], qx), __dsub_rn(dst[3 * jj + 1], qy), __dsub_rn(dst[3 * jj + 2], qz));
            worda = (worda & ~(1u << k)) | ((verdict_exact(A, B, kp.beta) ? 1u : 0u) << k);
            ++slow;
          }
          if (ambb) {
            const size_t jj = okb ? (size_t)jb : 0;
            const double A = sq3_ref(__dsub_rn(src[3 * jj], px), __dsub_rn(src[3 * jj + 1], py), __dsub_rn(src[3 * jj + 2], pz));
            const double B = sq3_ref(__dsub_rn(dst[3 * jj], qx), __dsub_rn(dst[3 * jj + 1], qy), __dsub_rn(dst[3 * jj + 2], qz));
            wordb = (wordb & ~(1u << k)) | ((verdict_exact(A, B, kp.beta) ? 1u : 0u) << k);
            ++slow;
          }
        }
      }
    }
    if (!kInterior) {  // rows past n, the diagonal and columns past n carry no pair
      const int r0 = i0 + 32 * c;
      const uint32_t rows = n - r0 >= 32 ? 0xffffffffu : (n - r0 <= 0 ? 0u : ((1u << (n - r0)) - 1u));
      uint32_t maska = oka ? rows : 0u, maskb = okb ? rows : 0u;
      if (ja >= r0 && ja < r0 + 32) maska &= ~(1u << (ja - r0));
      if (jb >= r0 && jb < r0 + 32) maskb &= ~(1u << (jb - r0));
      worda &= maska; wordb &= maskb;
    }
    dirT[warp][32 * c + lane] = warp_transpose32(worda, lane);
    dirT[4 + warp][32 * c + lane] = warp_transpose32(wordb, lane);
    tw[0] = c == 0 ? worda : tw[0]; tw[1] = c == 1 ? worda : tw[1]; tw[2] = c == 2 ? worda : tw[2]; tw[3] = c == 3 ? worda : tw[3];
    tw[4] = c == 0 ? wordb : tw[4]; tw[5] = c == 1 ? wordb : tw[5]; tw[6] = c == 2 ? wordb : tw[6]; tw[7] = c == 3 ? wordb : tw[7];
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kTile, B2R_G4_MINB) tim_graph4_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  const ProbDesc d = descs[blockIdx.y];
  const int n = d.n;
  const int T = (n + kTile - 1) / kTile;
  int p = blockIdx.x;
  int bi = 0, rowlen = (T + 1) / 2;
  while (bi < T && p >= rowlen) { p -= rowlen; ++bi; rowlen = (T - bi + 1) / 2; }
  if (bi >= T) return;
  const int bj = bi + 2 * p;
  __shared__ union { TilePts p64; TilePts4 p32; } sp;
  __shared__ __align__(16) union { uint32_t rows[kTile][8]; uint32_t cols[8][kTile]; } dir;
  uint32_t tw[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
  unsigned slow = 0;
  const int t = threadIdx.x;
  uint4 lo, hi;
  const bool use32 = d.st->use32 != 0;
  if (use32) {
    if (bj != bi && (bj + 2) * kTile <= n) compute_tile4<true>(d.src, d.dst, n, bi * kTile, bj * kTile, d.st, kp, sp.p32, dir.cols, tw, slow);
    else compute_tile4<false>(d.src, d.dst, n, bi * kTile, bj * kTile, d.st, kp, sp.p32, dir.cols, tw, slow);
    lo = make_uint4(dir.cols[0][t], dir.cols[1][t], dir.cols[2][t], dir.cols[3][t]);
    hi = make_uint4(dir.cols[4][t], dir.cols[5][t], dir.cols[6][t], dir.cols[7][t]);
  } else {
    compute_tile2<false>(d.src, d.dst, n, bi * kTile, bj * kTile, kp, sp.p64, dir.rows, tw, slow);
    lo = *reinterpret_cast<const uint4*>(&dir.rows[t][0]);
    hi = *reinterpret_cast<const uint4*>(&dir.rows[t][4]);
  }
  const int i = bi * kTile + t;
  if (i < n) {
    uint32_t* row = d.adj + (size_t)i * d.pitch + 4 * bj;
    *reinterpret_cast<uint4*>(row) = lo;
    if (bj + 1 < T) *reinterpret_cast<uint4*>(row + 4) = hi;
  }
  if (bj != bi) {
    const int j = bj * kTile + t;
    if (j < n) *reinterpret_cast<uint4*>(d.adj + (size_t)j * d.pitch + 4 * bi) = make_uint4(tw[0], tw[1], tw[2], tw[3]);
  }
  if (bj + 1 < T) {
    const int j = (bj + 1) * kTile + t;
    if (j < n) *reinterpret_cast<uint4*>(d.adj + (size_t)j * d.pitch + 4 * bi) = make_uint4(tw[4], tw[5], tw[6], tw[7]);
  }
  // v4 counts re-evaluated pairs one by one; the fp64 filter (use32 == 0) counts 2048-pair chunks per warp
  if (slow) {
    if (use32) atomicAdd(&d.st->slow_pairs, (unsigned long long)slow);
    else if ((threadIdx.x & 31) == 0) atomicAdd(&d.st->slow_pairs, 2048ull * slow);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Companion of the tensor-core filter (graph_tc.cu): (a) the pairs it listed as ambiguous are decided here, one thread per
// pair, with the reference's fp64 roundings and square roots, and both symmetric bits are set or cleared atomically;
// (b) problems the fp32 path may not decide (use32 == 0: non-finite input, extreme extents) or whose list overflowed run the
// fp64 filter (v2 tiles) over all their work items.  grid = (x, batch), 128 threads.
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTile) graph_fix_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  const ProbDesc d = descs[blockIdx.y];
  ProbState* st = d.st;
  if (st->tail_done || st->tail_why == 2) return;   // the fused tail (tail.cu) already fixed this problem's listed pairs
  const int n = d.n;
  const int t = threadIdx.x;
  if (st->use32 && !st->tc_overflow) {
    const unsigned cnt = min(st->amb_count, (unsigned)d.amb_cap);
    for (unsigned idx = blockIdx.x * kTile + t; idx < cnt; idx += gridDim.x * kTile) {
      fix_listed_pair(d.src, d.dst, d.adj, d.pitch, d.deg, d.amb_list[idx], kp.beta);
    }
    if (blockIdx.x == 0 && t == 0) { st->slow_pairs = cnt; st->deg_ready = 1; }
    return;
  }
  const int T = (n + kTile - 1) / kTile;
  __shared__ TilePts sp;
  __shared__ __align__(16) uint32_t dir[kTile][8];
  const int items = graph2_items(T);
  unsigned slow = 0;
  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    int p = item, bi = 0, rowlen = (T + 1) / 2;
    while (bi < T && p >= rowlen) { p -= rowlen; ++bi; rowlen = (T - bi + 1) / 2; }
    const int bj = bi + 2 * p;
    uint32_t tw[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    compute_tile2<false>(d.src, d.dst, n, bi * kTile, bj * kTile, kp, sp, dir, tw, slow);
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
    __syncthreads();
  }
  if (slow && (threadIdx.x & 31) == 0) atomicAdd(&st->slow_pairs, 2048ull * slow);
}

int launch_graph_fix(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, cudaStream_t st) {
  const int T = (n_max + kTile - 1) / kTile;
  int x = 1184 / (batch > 0 ? batch : 1);
  if (x < 8) x = 8;
  if (x > graph2_items(T)) x = graph2_items(T);
  if (x < 1) x = 1;
  graph_fix_kernel<<<dim3(x, batch), kTile, 0, st>>>(d_desc, kp);
  return 1;
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
  if (d.st->tail_done || d.st->tail_why == 2) return;   // the fused tail (tail.cu) already reduced this problem's degrees
  if (threadIdx.x < 8) s_key[threadIdx.x] = 0ull;
  if (threadIdx.x == 8) s_sum = 0ull;
  __syncthreads();
  const int W4 = (d.n + kTile - 1) / kTile;   // 16-byte groups per row
  unsigned long long wsum = 0;
  if (d.st->deg_ready) {
    // the tensor-core filter already left the degrees in deg[] (popcounts of the words it wrote, corrected by the fix-up):
    // only the per-problem maxima and the seed keys remain
    if (threadIdx.x < kDegRowsPerBlock) {
      const int v = v0 + threadIdx.x;
      if (v < d.n) {
        const int cr = d.deg[v];
        wsum = (unsigned long long)cr;
        atomicMax(&s_key[v & 7], ((unsigned long long)cr << 32) | (unsigned long long)(0xffffffffu - (unsigned)v));
      }
    }
    wsum = __reduce_add_sync(0xffffffffu, (unsigned)wsum);
  } else {
  // four rows in flight per warp (16-byte loads, all issued before the reductions): the kernel is a pure HBM / L2 stream
#pragma unroll 1
  for (int k0 = 0; k0 < 8; k0 += 4) {
    int c[4] = {0, 0, 0, 0};
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int v = v0 + warp * 8 + k0 + r;
      if (v < d.n) {
        const uint4* row = reinterpret_cast<const uint4*>(d.adj + (size_t)v * d.pitch);
        for (int w = lane; w < W4; w += 32) {
          const uint4 q = row[w];
          c[r] += __popc(q.x) + __popc(q.y) + __popc(q.z) + __popc(q.w);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int v = v0 + warp * 8 + k0 + r;
      const int cr = __reduce_add_sync(0xffffffffu, c[r]);
      if (lane == 0 && v < d.n) {
        d.deg[v] = cr;
        wsum += (unsigned long long)cr;
        atomicMax(&s_key[v & 7], ((unsigned long long)cr << 32) | (unsigned long long)(0xffffffffu - (unsigned)v));
      }
    }
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
  else if (kp.graph_v2) tim_graph2_kernel<<<dim3(graph2_items(T), batch), kTile, 0, st>>>(d_desc, kp);
  else if (!kp.graph_v3 && !kp.graph_v4) return launch_tim_graph_tc_and_fix(d_desc, batch, n_max, kp, st);
  else {
    graph_prep_kernel<<<batch, 256, 0, st>>>(d_desc, kp);
    if (kp.graph_v3) tim_graph3_kernel<<<dim3(graph2_items(T), batch), kTile, 0, st>>>(d_desc, kp);
    else tim_graph4_kernel<<<dim3(graph2_items(T), batch), kTile, 0, st>>>(d_desc, kp);
    return 2;
  }
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
