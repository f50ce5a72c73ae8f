// TIM pairwise-consistency graph, filter v5: the pair test on the 5th-generation tensor cores.
//
// Same contract as graph.cu (the bit of pair (i, j) equals the reference's fp64 verdict | ||src_j - src_i|| - ||dst_j - dst_i|| | <= beta,
// upstream computeTIMs x2 + the addEdge loop behind `solver.solve`, HelpersRealtime.py:166 / registration_utils.py:261).
//
// Work in units of beta: x' = (x - centre) / beta, A' = |s'_j - s'_i|^2, B' = |d'_j - d'_i|^2.  With Dm = A' - B', L = 1 - 2 (A' + B'),
//     G = Dm^2 + L = ((a' - b')^2 - 1) ((a' + b')^2 - 1),           edge  <=>  G <= 0   whenever A' + B' > 1.
// Dm and L are BILINEAR in per-point quantities (r = squared norm):
//     L_ij  = p_i + p_j + 4 s_i.s_j + 4 d_i.d_j,       p  = 1/2 - 2 (r_s + r_d)
//     Dm_ij = m_i + m_j - 2 s_i.s_j + 2 d_i.d_j,       m  = r_s - r_d
// so one tcgen05.mma (kind::tf32, M = 128 rows, N = 64 columns, K = 32) per output produces a 128 x 64 block of each in TMEM.
// Every operand entry is an exact TF32 value: coordinates are rounded ONCE to 22 significant bits (hi + lo, both TF32) and the
// products hi.hi + hi.lo + lo.hi are laid along K (3xTF32, the lo.lo term is bounded); the fp32 scalars p and m are split into
// three TF32 pieces (exact).  The epilogue (4 warps, thread = row, tcgen05.ld 32 columns at a time) is per pair:
//     G = fma(Dm, Dm, L),  T = fma(L, -t1, t0)  (packed f32x2),  sign(G) funnel-shifted into the adjacency word of the row,
//     one FSETP each for  !(|G| >= T)  and  L > Lcut  ORed into a per-chunk "ambiguous" predicate.
// Ambiguous pairs (|G| inside the error band, or both lengths possibly below beta) are appended to a per-problem list and
// decided afterwards, one thread per pair, with the reference's fp64 roundings and square roots (graph_fix_kernel in graph.cu).
//
// Error budget (w = 2^-22, u = 2^-24, Pi = Ps + Pd = bounds of the largest squared norms, all in beta units), DESIGN.md
// "graph filter v5":  geometry (22-bit copies) 17.34 w Pi on A' + B';  scalars u (1 + 4 Pi) resp. 2 u Pi;  dropped lo.lo 4 w Pi;
// tensor-core accumulation ASSUMED within 2^-20 of the sum of |products| (measured on B200: below 2^-23, tests/test_gpu_teaser.py);
//     E_L = w (71.8 Pi + 4.3),   E_D = 45.9 w Pi,
//     |G~ - G| <= u |G~| + 2 E_D |Dm~| + E_D^2 + E_L + E_scale,     2 E_D |Dm~| <= Dm~^2 / k + k E_D^2,  Dm~^2 <= (|G~| + |L~|)(1 + 2u)
// and T is twice that bound plus the 4e-15 (A' + B' + 1)^2 inside which the reference's own roundings could flip the verdict,
// written as a linear function of L~ (|L~| <= 2.001 + 2 E_L - L~).
#include <math_constants.h>
#include <stdlib.h>

#include "common.cuh"
#include "exact.cuh"

namespace b2r {

namespace {

constexpr int kTcRowTile = 128;     // rows per CTA (TMEM lanes)
constexpr int kTcUnit = 64;         // columns per MMA / TMEM stage
#ifndef B2R_TC_UPI
#define B2R_TC_UPI 32
#endif
constexpr int kTcUnitsPerItem = B2R_TC_UPI; // column units one CTA walks (32 = 16 column tiles; 16: 0.490 ms, 32: 0.475 ms per 64 x 5000 stage)
constexpr int kTcEpiWarps = 8;      // epilogue warps: lane quarter w & 3, 32-column half w >> 2 of every unit
constexpr int kTcThreads = 32 * (kTcEpiWarps + 1);   // + one producer / MMA-issuer warp
constexpr uint32_t kTcTmemCols = 256;  // 2 stages x (Dm 64 + L 64)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar_addr) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_addr) : "memory");
}
// (a suspend-time hint on this try_wait changes nothing measurable: 0.4755 vs 0.4751 ms per 64 x 5000 stage)
__device__ __forceinline__ void mbar_wait_a(uint32_t bar_addr, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(bar_addr), "r"(parity) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tc_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand tile, 128-byte swizzle: row r at r * 128 B, its 16-byte chunk c at position c ^ (r & 7); 8-row groups 1024 B apart.
// Same encodings as nn_gemm.cu (verified on B200): start >> 4, LBO = 1, SBO = 1024 >> 4, version 1, layout 2 = SWIZZLE_128B.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ constexpr uint32_t umma_idesc_tf32(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ float to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ float2 f2fma(float2 a, float2 b, float2 c) {
  float2 r;
  asm("{.reg .b64 ra, rb, rc, rd; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; mov.b64 rc, {%6,%7}; fma.rn.f32x2 rd, ra, rb, rc; "
      "mov.b64 {%0,%1}, rd;}"
      : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
  return r;
}

// The same 32 x 32 bit transpose in 13 instructions instead of ~50: the two byte-granular steps are one PRMT each (per-lane
// selector), the three sub-byte steps one rotate (per-lane amount) + one LOP3 bit-select (per-lane mask) each.
struct TposeConst { uint32_t sel16, sel8, km4, km2, km1; int sh4, sh2, sh1; };
__device__ __forceinline__ TposeConst tpose_const(int lane) {
  TposeConst k;
  k.sel16 = (lane & 16) ? 0x3276u : 0x5410u;
  k.sel8 = (lane & 8) ? 0x3715u : 0x6240u;
  k.km4 = (lane & 4) ? 0xf0f0f0f0u : 0x0f0f0f0fu;  k.sh4 = (lane & 4) ? 28 : 4;
  k.km2 = (lane & 2) ? 0xccccccccu : 0x33333333u;  k.sh2 = (lane & 2) ? 30 : 2;
  k.km1 = (lane & 1) ? 0xaaaaaaaau : 0x55555555u;  k.sh1 = (lane & 1) ? 31 : 1;
  // opaque to the compiler: otherwise it re-derives all eight from the lane id (R2P + SEL + MOV, ~16 instructions) at every use
  asm volatile("" : "+r"(k.sel16), "+r"(k.sel8), "+r"(k.km4), "+r"(k.km2), "+r"(k.km1), "+r"(k.sh4), "+r"(k.sh2), "+r"(k.sh1));
  return k;
}
__device__ __forceinline__ uint32_t warp_transpose32_fast(uint32_t x, const TposeConst& k) {
  uint32_t y = __shfl_xor_sync(0xffffffffu, x, 16);
  x = __byte_perm(x, y, k.sel16);
  y = __shfl_xor_sync(0xffffffffu, x, 8);
  x = __byte_perm(x, y, k.sel8);
  y = __shfl_xor_sync(0xffffffffu, x, 4);
  y = __funnelshift_l(y, y, k.sh4);
  x = (x & k.km4) | (y & ~k.km4);
  y = __shfl_xor_sync(0xffffffffu, x, 2);
  y = __funnelshift_l(y, y, k.sh2);
  x = (x & k.km2) | (y & ~k.km2);
  y = __shfl_xor_sync(0xffffffffu, x, 1);
  y = __funnelshift_l(y, y, k.sh1);
  x = (x & k.km1) | (y & ~k.km1);
  return x;
}
__device__ __forceinline__ float2 f2mul(float2 a, float2 b) {
  float2 r;
  asm("{.reg .b64 ra, rb, rd; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; mul.rn.f32x2 rd, ra, rb; mov.b64 {%0,%1}, rd;}"
      : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return r;
}

}  // namespace

// ---------------------------------------------------------------------------------------------------------------------
// prep: one CTA per problem: bounding boxes -> centres, extents, band constants (ProbState).  graph_tc_ops_kernel then writes
// the operand rows of every point (zero-filled up to the tile edge).
// ---------------------------------------------------------------------------------------------------------------------
// band constants of one problem from the bounding boxes of its two clouds (one thread); st == nullptr: only the decision
// "the fp32 path may decide this problem" is returned (the other slices of graph_tc_setup_kernel)
__device__ bool tc_write_constants(ProbState* st, const double (&lo6)[6], const double (&hi6)[6], bool bad, int n, const KernelParams& kp) {
  const double beta = fabs(kp.beta);
  const double kappa = 1.0 / beta;
  double Pi = 0.0, Mmax = 0.0;
  for (int k = 0; k < 6; ++k) {
    const double lo = lo6[k], hi = hi6[k];
    const double c = 0.5 * lo + 0.5 * hi;
    if (st) st->ctr[k] = c;
    const double h = fmax(hi - c, c - lo) * kappa * (1.0 + 1e-6);   // covers the roundings of the 22-bit copy
    Mmax = fmax(Mmax, h);
    Pi += h * h;
  }
  const double w22 = 2.384185791015625e-7, u = 5.9604644775390625e-8;   // 2^-22, 2^-24
  const double E_L = w22 * (71.8 * Pi + 4.3), E_D = 45.9 * w22 * Pi;
  const double tmin = kp.filter_tau_min;                                   // tests widen the band with it
  // the fp32 path decides only well-conditioned problems; everything else goes to the fp64 filter (graph_fix_kernel)
  const bool ok = !bad && n > 0 && beta >= 1e-300 && kappa <= 1e300 && Mmax < 1e12 && Pi > 1e-20 && E_L < 1e3 && tmin < 0.015625;
  if (!st) return ok;
  st->use32 = ok ? 1 : 0;
  st->kappa = kappa;
  if (ok) {
    double k = sqrt(2.0 * Pi) / E_D;
    if (!(k >= 8.0)) k = 8.0;
    if (k > 1e12) k = 1e12;
    const double E_scale = 2e-15 * (1.0 + 8.0 * Pi), refc = 4e-15 * (1.0 + 8.0 * Pi);
    const double den = 1.0 - 2.0 * (1.0 + 2.0 * u) / k;
    const double pad = 1.0 + 1e-5;                                         // fp32 evaluation of T
    // T(L) = t0 - t1 L;  tmin (tests) widens the band relative to (A' + B' + 1)^2 <= (1 + 8 Pi)(3.001 - L) / 2
    const double t1 = (2.0 / k + refc + tmin * (1.0 + 8.0 * Pi)) / den * pad;
    const double labs = 2.001 + 2.0 * E_L;                                 // |L~| <= labs - L~  (L~ <= 1 + E_L)
    const double t0 = (labs * 2.0 / k + 2.0 * (k + 1.0) * E_D * E_D + 2.0 * E_L + 2.0 * E_scale + (labs + 1.0) * (refc + tmin * (1.0 + 8.0 * Pi))) / den * pad;
    st->tcc[0] = -__double2float_ru(t1);
    st->tcc[1] = __double2float_ru(t0);
    st->tcc[2] = -__double2float_ru(1.001 + 1.01 * E_L);                   // L~ <= this  =>  A' + B' > 1
    st->tcc[3] = (float)k;
    st->tc_pi = __double2float_ru(Pi);
  }
  return ok;
}

// ---------------------------------------------------------------------------------------------------------------------
// operand rows (K = 32 TF32 values = 128 bytes = 8 chunks of 16 bytes, chunk c of row r stored at c ^ (r & 7))
//   k:      0-2     3-5     6-8     9-11    12-14   15-17   18-20    21-23    24-26    27-29   30-31
//   column  s_hi    s_lo    s_hi    d_hi    d_lo    d_hi    p (3)    1 1 1    m (3)    1 1 1   0
//   row L   4 s_hi  4 s_hi  4 s_lo  4 d_hi  4 d_hi  4 d_lo  1 1 1    p (3)    0        0       0
//   row D  -2 s_hi -2 s_hi -2 s_lo  2 d_hi  2 d_hi  2 d_lo  0        0        1 1 1    m (3)   0
// ---------------------------------------------------------------------------------------------------------------------
namespace {

struct Split { float hi[6], lo[6], p[3], m[3]; };

__device__ __forceinline__ Split split_record(const float4 a, const float4 b) {
  Split s;
  const float v[6] = {a.x, a.y, a.z, b.x, b.y, b.z};
#pragma unroll
  for (int k = 0; k < 6; ++k) { s.hi[k] = to_tf32(v[k]); s.lo[k] = v[k] - s.hi[k]; }   // exact: v has 22 significant bits
  s.p[0] = to_tf32(a.w); { const float r = a.w - s.p[0]; s.p[1] = to_tf32(r); s.p[2] = r - s.p[1]; }
  s.m[0] = to_tf32(b.w); { const float r = b.w - s.m[0]; s.m[1] = to_tf32(r); s.m[2] = r - s.m[1]; }
  return s;
}

__device__ __forceinline__ void store_row(unsigned char* tile, int r, const float (&v)[32]) {
#pragma unroll
  for (int c = 0; c < 8; ++c)
    *reinterpret_cast<float4*>(tile + r * 128 + ((c ^ (r & 7)) << 4)) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
}

__device__ __forceinline__ void store_col_row(unsigned char* tile, int r, const Split& s) {
  float v[32];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    v[k] = s.hi[k]; v[3 + k] = s.lo[k]; v[6 + k] = s.hi[k];
    v[9 + k] = s.hi[3 + k]; v[12 + k] = s.lo[3 + k]; v[15 + k] = s.hi[3 + k];
    v[18 + k] = s.p[k]; v[21 + k] = 1.f; v[24 + k] = s.m[k]; v[27 + k] = 1.f;
  }
  v[30] = 0.f; v[31] = 0.f;
  store_row(tile, r, v);
}

// row-L and row-D operand rows of point r from its column row (all scalings are powers of two: exact)
__device__ __forceinline__ void rows_from_col_row(const unsigned char* stage, unsigned char* tile_l, unsigned char* tile_d, int r) {
  float c[32];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const float4 q = *reinterpret_cast<const float4*>(stage + r * 128 + ((k ^ (r & 7)) << 4));
    c[4 * k] = q.x; c[4 * k + 1] = q.y; c[4 * k + 2] = q.z; c[4 * k + 3] = q.w;
  }
  float v[32];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    v[k] = 4.f * c[k]; v[3 + k] = 4.f * c[k]; v[6 + k] = 4.f * c[3 + k];
    v[9 + k] = 4.f * c[9 + k]; v[12 + k] = 4.f * c[9 + k]; v[15 + k] = 4.f * c[12 + k];
    v[18 + k] = 1.f; v[21 + k] = c[18 + k]; v[24 + k] = 0.f; v[27 + k] = 0.f;
  }
  v[30] = 0.f; v[31] = 0.f;
  store_row(tile_l, r, v);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    v[k] = -2.f * c[k]; v[3 + k] = -2.f * c[k]; v[6 + k] = -2.f * c[3 + k];
    v[9 + k] = 2.f * c[9 + k]; v[12 + k] = 2.f * c[9 + k]; v[15 + k] = 2.f * c[12 + k];
    v[18 + k] = 0.f; v[21 + k] = 0.f; v[24 + k] = 1.f; v[27 + k] = c[24 + k];
  }
  store_row(tile_d, r, v);
}

constexpr int kTcAmbBuf = 64;   // ambiguous pairs a warp collects in shared memory before it touches the global list
struct TcSmemCtl {
  uint64_t rowfull, colfull[4], full[2], empty[2];
  uint32_t tmem_base;
  int amb_cnt[8];
  uint32_t amb_buf[8][kTcAmbBuf];
};

constexpr int kTcRowBytes = kTcRowTile * 128;   // 16 KB
constexpr int kTcColBytes = kTcUnit * 128;      // 8 KB
constexpr int kTcCtlBytes = 3 * 1024;           // TcSmemCtl (barriers + the warps' pair buffers)
static_assert(sizeof(TcSmemCtl) <= kTcCtlBytes, "control block overlaps the staging tile");
constexpr int kTcStageBytes = kTcRowTile * 2 * kTcUnitsPerItem * 4;   // the item's adjacency words: 128 rows x 2 words per unit (>= the 16 KB of operand rows staged there during set-up)
static_assert(kTcStageBytes >= kTcRowBytes, "staging tile smaller than the operand rows it holds during set-up");
constexpr int kTcSmemBytes = 2 * kTcRowBytes + 4 * kTcColBytes + kTcCtlBytes + kTcStageBytes + 1024;   // + alignment slack = 84 KB; two CTAs per SM by registers (2 x 288 x 96) and TMEM (2 x 256 columns)

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// 1-D bulk copy global -> shared through the async proxy (the operand rows are stored pre-swizzled, so no tensor map is needed)
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
               "r"(smem_u32(bar))
               : "memory");
}

}  // namespace

// Column-operand row of every point (128 bytes), already in the 128-byte-swizzled order the MMA reads (row j of a 64- or
// 128-row tile uses j & 7, and tiles start at multiples of 64).  The row-L / row-D operands of a CTA's 128 rows are derived
// from these rows in shared memory (rows_from_col_row).  grid = (n_pad / 128, batch).
__device__ __forceinline__ void tc_store_operand_row(const ProbDesc& d, const double (&ctr)[6], double kappa, int i, int n) {
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
  if (i < n) {
    d.deg[i] = 0;   // the filter kernel accumulates the vertex degrees (popcounts of the words it writes)
    float v[6];
    double r[2] = {0.0, 0.0};
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const double x = ((k < 3 ? d.src[3 * (size_t)i + k] : d.dst[3 * (size_t)i + k - 3]) - ctr[k]) * kappa;
      const float x32 = __double2float_rn(x);
      const float hi = to_tf32(x32);
      const float lo = to_tf32(x32 - hi);
      v[k] = hi + lo;                       // exact: 22 significant bits
      r[k / 3] = fma((double)v[k], (double)v[k], r[k / 3]);
    }
    a = make_float4(v[0], v[1], v[2], __double2float_rn(0.5 - 2.0 * (r[0] + r[1])));
    b = make_float4(v[3], v[4], v[5], __double2float_rn(r[0] - r[1]));
  }
  const Split sp = split_record(a, b);
  store_col_row(reinterpret_cast<unsigned char*>(d.pts32), i, sp);
}

// prep + operand rows in ONE launch: grid = (slices, batch).  Every CTA of a problem reduces the bounding boxes of the whole
// problem itself (48 n bytes out of L2; min / max are exact, so all slices agree bit for bit), slice 0 writes the constants, and
// each slice writes the operand rows of its share of the points.  No grid-wide synchronisation is needed.
__global__ void __launch_bounds__(1024) graph_tc_setup_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  const ProbDesc d = descs[blockIdx.y];
  const int n = d.n;
  const int t = threadIdx.x;
  double mn[6], mx[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) { mn[k] = CUDART_INF; mx[k] = -CUDART_INF; }
  bool bad = false;
  for (int i = t; i < n; i += 1024) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double a = d.src[3 * (size_t)i + k], b = d.dst[3 * (size_t)i + k];
      bad |= !(fabs(a) <= 1.79e308) | !(fabs(b) <= 1.79e308);
      mn[k] = fmin(mn[k], a); mx[k] = fmax(mx[k], a);
      mn[3 + k] = fmin(mn[3 + k], b); mx[3 + k] = fmax(mx[3 + k], b);
    }
  }
  __shared__ double s_mn[32][6], s_mx[32][6];
  __shared__ double s_ctr[6], s_kappa;
  __shared__ int s_bad, s_use32;
  if (t == 0) s_bad = 0;
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 6; ++k) {
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      mn[k] = fmin(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
      mx[k] = fmax(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
    }
  }
  if ((t & 31) == 0) {
#pragma unroll
    for (int k = 0; k < 6; ++k) { s_mn[t >> 5][k] = mn[k]; s_mx[t >> 5][k] = mx[k]; }
  }
  if (bad) atomicOr(&s_bad, 1);
  __syncthreads();
  if (t == 0) {
    double lo6[6], hi6[6];
    for (int k = 0; k < 6; ++k) {
      double lo = s_mn[0][k], hi = s_mx[0][k];
      for (int w = 1; w < 32; ++w) { lo = fmin(lo, s_mn[w][k]); hi = fmax(hi, s_mx[w][k]); }
      lo6[k] = lo; hi6[k] = hi;
      s_ctr[k] = 0.5 * lo + 0.5 * hi;    // the same expression tc_write_constants stores
    }
    // slice 0 writes the state; the others take the same decision without touching it
    s_use32 = tc_write_constants(blockIdx.x == 0 ? d.st : nullptr, lo6, hi6, s_bad != 0, n, kp) ? 1 : 0;
    s_kappa = 1.0 / fabs(kp.beta);
  }
  __syncthreads();
  if (!s_use32) return;
  double ctr[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) ctr[k] = s_ctr[k];
  const double kappa = s_kappa;
  const int n_pad = (n + kTile - 1) / kTile * kTile;
  const int per = ((n_pad / kTile + gridDim.x - 1) / gridDim.x) * kTile;   // whole 128-row blocks per slice
  const int i_begin = blockIdx.x * per, i_end = min(i_begin + per, n_pad);
  for (int i = i_begin + t; i < i_end; i += 1024) tc_store_operand_row(d, ctr, kappa, i, n);
}

// work items of one problem: row block bi walks its column units v = 0 .. 2 (T - bi) - 1 in chunks of kTcUnitsPerItem
__host__ __device__ inline int tc_items(int T) {
  int s = 0;
  for (int bi = 0; bi < T; ++bi) s += (2 * (T - bi) + kTcUnitsPerItem - 1) / kTcUnitsPerItem;
  return s;
}

// ---------------------------------------------------------------------------------------------------------------------
// The filter kernel.  grid = (work items of the largest problem, batch); an item is a 128-row block and up to 16 column units of
// 64 columns.  Warp 8 (one lane) feeds the tensor cores: operand rows by cp.async.bulk into a 4-slot ring, 2 x 4 tcgen05.mma per
// unit into one of two TMEM stages (Dm: 64 columns, L: 64 columns), tcgen05.commit on the stage's `full` barrier.  Warps 0-7 are
// the epilogue: warp w owns TMEM lanes 32 (w & 3) .. + 31 (rows) and columns 32 (w >> 2) .. + 31 of every unit.
//
// What the epilogue is built around (ncu of the earlier kernels, profiles/r01_graph_kernel_ncu_full.md and
// profiles/r02_graph6_ncu.md: latency bound -- four epilogue warps per scheduler, issue slot busy 48 % -- with the LSU and
// L1 -> L2 paths at ~50 % from 4-byte scattered adjacency stores, 5.3 sectors written per 32 useful bytes):
//   * accumulators fetched 16 columns at a time (32 live registers: the loop state fits the 96 registers two resident CTAs
//     leave a thread), four independent 4-column bit accumulators per fetch, the TMEM stage a compile-time constant;
//   * edge units (diagonal tile, last column tile, partial row block) take a masked path, interior units none of it;
//   * an ambiguous chunk replays only the flagged 4-column groups;
//   * STAGED: the unit's adjacency word goes to shared memory (the 16 KB that staged the CTA's own operand rows; word k of row
//     r at k ^ (r & 31), conflict free for every access pattern below) and NOTHING else happens per unit.  After the last unit
//     the eight warps flush the item: rows as 128-byte coalesced stores, and the mirrored blocks as sixteen INDEPENDENT
//     32 x 32 bit transposes per warp (13 instructions each, PRMT / rotate + bit-select with per-lane constants) whose four
//     words per column leave as one 16-byte store and ONE degree atomic.  Per item: 1.5 k sectors written instead of 8 k, a
//     quarter of the degree atomics, and the transposes' shuffle chains overlap each other instead of stalling the unit loop.
// Bits of ambiguous pairs are written as 0 and are NOT counted in the degrees: the exact fix-up (graph_fix_kernel / tail_kernel)
// only ever SETS bits and increments, without reading anything back.
// ---------------------------------------------------------------------------------------------------------------------
template <int S_> struct StageC { static constexpr int value = S_; };

template <bool LD16, bool STAGED, bool PEERS = false>
__global__ void __launch_bounds__(kTcThreads, 2) tim_graph6_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  const ProbDesc d = descs[blockIdx.y];
  const int n = d.n;
  if (!d.st->use32) return;
  const int T = (n + kTile - 1) / kTile;
  // item -> (row block, first unit) in closed form (every warp of every CTA needs it before anything else; the loop over row
  // blocks this replaces was a 14-iteration dependent chain on the uniform datapath).  Row block bi owns ceil(m / G) items,
  // m = T - bi; counted from the END, S(m) = G q (q + 1) / 2 + r (q + 1) items belong to blocks of fewer than m + 1 rows (m = G q + r).
  constexpr int G = kTcUnitsPerItem / 2;   // row blocks whose item count is the same: ceil(m / G)
  int bi, v0;
  {
    const int qT = T / G, rT = T % G;
    const int total = G * qT * (qT + 1) / 2 + rT * (qT + 1);
    // PEERS (b2r_solve_sharded): rank r of w takes the items r, r + w, r + 2 w, ... -- neighbouring items have the same size
    const int item = PEERS ? (int)blockIdx.x * kp.shard_world + kp.shard_rank : (int)blockIdx.x;
    if (item >= total) return;
    const int pe = total - 1 - item;
    int q = (int)((sqrtf(1.0f + (8.0f / G) * (float)pe) - 1.0f) * 0.5f);
    q += (G * (q + 1) * (q + 2) / 2 <= pe) ? 1 : 0;
    q -= (G * q * (q + 1) / 2 > pe) ? 1 : 0;
    const int rem = pe - G * q * (q + 1) / 2;
    const int r = rem / (q + 1);
    const int m = G * q + r + 1;
    bi = T - m;
    v0 = (q - (rem - r * (q + 1))) * kTcUnitsPerItem;
  }
  const int nu = min(kTcUnitsPerItem, 2 * (T - bi) - v0);
  const int i0 = bi * kTile;

  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* sm = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  unsigned char* sRowL = sm;
  unsigned char* sRowD = sm + kTcRowBytes;
  unsigned char* sCol = sm + 2 * kTcRowBytes;   // ring of 4 slots
  TcSmemCtl* g = reinterpret_cast<TcSmemCtl*>(sm + 2 * kTcRowBytes + 4 * kTcColBytes);
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const unsigned char* ops_col = reinterpret_cast<const unsigned char*>(d.pts32);
  // column rows of the CTA's own 128 points (16 KB) during set-up; STAGED: the item's adjacency words (128 rows x 32) afterwards
  unsigned char* sStage = sm + 2 * kTcRowBytes + 4 * kTcColBytes + kTcCtlBytes;

  if (t == 32 * kTcEpiWarps) {
    mbar_init(&g->rowfull, 1);
    for (int s = 0; s < 4; ++s) mbar_init(&g->colfull[s], 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&g->full[s], 1); mbar_init(&g->empty[s], kTcEpiWarps); }
    for (int w = 0; w < kTcEpiWarps; ++w) g->amb_cnt[w] = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_expect_tx(&g->rowfull, kTcRowBytes);
    bulk_g2s(sStage, ops_col + (size_t)i0 * 128, kTcRowBytes, &g->rowfull);
    for (int u = 0; u < 3 && u < nu; ++u) {  // the first column units start travelling before the set-up barrier
      const int j0 = i0 + (v0 + u) * kTcUnit;
      mbar_expect_tx(&g->colfull[u & 3], kTcColBytes);
      bulk_g2s(sCol + (u & 3) * kTcColBytes, ops_col + (size_t)j0 * 128, kTcColBytes, &g->colfull[u & 3]);
    }
  }
  __syncthreads();
  if (warp == kTcEpiWarps) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&g->tmem_base)), "r"(kTcTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  } else if (t < kTcRowTile) {
    mbar_wait(&g->rowfull, 0);
    rows_from_col_row(sStage, sRowL, sRowD, t);
    fence_proxy_async();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = g->tmem_base;

  if (warp == kTcEpiWarps) {
    // ===== producer (one lane) =====
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_tf32(kTcRowTile, kTcUnit);
      const uint32_t rowl = smem_u32(sRowL), rowd = smem_u32(sRowD);
      auto copy_unit = [&](int u) {
        const int j0 = i0 + (v0 + u) * kTcUnit;
        mbar_expect_tx(&g->colfull[u & 3], kTcColBytes);
        bulk_g2s(sCol + (u & 3) * kTcColBytes, ops_col + (size_t)j0 * 128, kTcColBytes, &g->colfull[u & 3]);
      };
      for (int u = 0; u < nu; ++u) {
        const int s = u & 1;
        mbar_wait(&g->colfull[u & 3], (u >> 2) & 1);
        mbar_wait(&g->empty[s], ((u >> 1) & 1) ^ 1);   // epilogue done with the TMEM stage
        tc_fence_after();
        const uint32_t cb = smem_u32(sCol + (u & 3) * kTcColBytes);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) tc_mma_tf32(tmem + s * 128, umma_desc_sw128(rowd) + 2 * ks, umma_desc_sw128(cb) + 2 * ks, idesc, ks != 0 ? 1u : 0u);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) tc_mma_tf32(tmem + s * 128 + 64, umma_desc_sw128(rowl) + 2 * ks, umma_desc_sw128(cb) + 2 * ks, idesc, ks != 0 ? 1u : 0u);
        tc_commit(&g->full[s]);
        if (u + 3 < nu) {  // slot (u + 3) & 3 held unit u - 1: its MMAs must have retired before the slot is overwritten
          if (u >= 1) mbar_wait(&g->full[(u - 1) & 1], ((u - 1) >> 1) & 1);
          copy_unit(u + 3);
        }
      }
    }
    __syncwarp();
  } else {
    // ===== epilogue =====
    const int q = warp & 3, ch = warp >> 2;
    const int r = 32 * q + lane;
    const int i = i0 + r;
    const ProbState* stp = d.st;
    const float nt1s = stp->tcc[0], t0s = stp->tcc[1];
    const float2 nt1 = make_float2(nt1s, nt1s), t0 = make_float2(t0s, t0s);
    const uint32_t lcut_bits = __float_as_uint(stp->tcc[2]);   // lcut < 0:  L > lcut  <=>  bits(L) < bits(lcut) as unsigned
    const bool rows_in = i0 + kTile <= n;
    // units [ub, ue) of this item are interior: off the diagonal tile, all 64 columns and all 128 rows inside the problem
    const int ub = v0 == 0 ? 2 : 0;
    const int ue = rows_in ? max(0, min(nu, (n - i0) / kTcUnit - v0)) : 0;
    int jc = i0 + v0 * kTcUnit + 32 * ch;                                     // first column of this warp's chunk in unit 0
    uint32_t tbase = tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(ch * 32);
    asm volatile("" : "+r"(tbase));   // keep it in a register (re-derived from the thread id per unit otherwise)
    uint32_t bar_full = smem_u32(&g->full[0]);          // full[S] at + 8 S, empty[S] at + 16 + 8 S
    asm volatile("" : "+r"(bar_full));
    int rowdeg = 0;
    uint32_t ph = 0;
    // STAGED: word k = 2 u + ch of row r lives at sStage + 4 (kSW r + (k ^ (r & 31)));  r & 31 == lane for the writer
    constexpr int kSW = 2 * kTcUnitsPerItem;   // staged words per row
    uint32_t* const stage_w = reinterpret_cast<uint32_t*>(sStage);
    uint32_t stage_row = smem_u32(sStage) + 4u * kSW * (uint32_t)r;
    asm volatile("" : "+r"(stage_row));
    // !STAGED: running pointers of the direct stores
    TposeConst tk;
    if (!STAGED) tk = tpose_const(lane);
    uint32_t* rowp = d.adj + (size_t)min(i, n - 1) * d.pitch + (jc >> 5);     // += 2 words per unit
    uint32_t* colp = d.adj + (size_t)(jc + lane) * d.pitch + (i0 >> 5) + q;   // += 64 rows per unit (dereferenced only inside the problem)
    int* degp = d.deg + jc + lane;
    const size_t colstep = (size_t)kTcUnit * d.pitch;

    // pair (i, jc + c) for every set bit c of aw goes to the warp's shared-memory buffer (flushed once per item: a global atomic
    // with its L2 round trip here would make this warp lag a whole unit behind the others, and they share the TMEM stages)
    auto list_pairs = [&](uint32_t aw) {
      while (aw) {
        const int c = __ffs(aw) - 1;
        aw &= aw - 1;
        const uint32_t e = ((uint32_t)i << 16) | (uint32_t)(jc + c);
        const int pos = atomicAdd(&g->amb_cnt[warp], 1);
        if (pos < kTcAmbBuf) g->amb_buf[warp][pos] = e;
        else {  // buffer full (wide test bands): straight to the global list
          const unsigned base = atomicAdd(&d.st->amb_count, 1u);
          if (base < (unsigned)d.amb_cap) d.amb_list[base] = e;
          else d.st->tc_overflow = 1;
        }
      }
    };
    // the unit's finished word: into the staging tile, or straight to the row and (transposed) to the 32 mirrored rows
    auto emit = [&](uint32_t word, const int u, const bool interior) {
      rowdeg += __popc(word);
      if (STAGED) {
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(stage_row + (uint32_t)(((2 * u + ch) ^ lane) << 2)), "r"(word) : "memory");
      } else {
        if (interior) {
          const uint32_t tw = warp_transpose32_fast(word, tk);
          *colp = tw;
          if (tw) atomicAdd(degp, __popc(tw));
          *rowp = word;
        } else {
          if (u >= ub) {
            const uint32_t tw = warp_transpose32_fast(word, tk);
            if (jc + lane < n) {
              *colp = tw;
              if (tw) atomicAdd(degp, __popc(tw));
            }
          }
          if (i < n) *rowp = word;
        }
        rowp += 2; colp += colstep; degp += kTcUnit;
      }
      jc += kTcUnit;
    };
    auto edge_mask = [&]() -> uint32_t {
      const int nc = n - jc;
      uint32_t mask = nc >= 32 ? 0xffffffffu : (nc <= 0 ? 0u : ((1u << nc) - 1u));
      if (i >= n) mask = 0u;
      if (i >= jc && i < jc + 32) mask &= ~(1u << (i - jc));
      return mask;
    };

    // 32 columns per fetch (64 live accumulator registers)
    auto unit32 = [&](auto stage_c, const int u) {
      constexpr int S = decltype(stage_c)::value;
      mbar_wait_a(bar_full + 8 * S, ph);
      tc_fence_after();
      uint32_t dm[32], lv[32];
      tc_ld_32x32(tbase + S * 128, dm);
      tc_ld_32x32(tbase + S * 128 + 64, lv);
      tc_wait_ld();
      tc_fence_before();   // the accumulators are in registers: release the TMEM stage right away
      __syncwarp();
      if (lane == 0) mbar_arrive_a(bar_full + 16 + 8 * S);
      // four independent 8-column accumulators; within one the higher column goes first, so column c ends in bit c & 7
      uint32_t wq[4] = {0u, 0u, 0u, 0u}, lmin[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
      bool aq[4] = {false, false, false, false};
#pragma unroll
      for (int c = 6; c >= 0; c -= 2) {
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          const int cc = 8 * h + c;
          const float2 D = make_float2(__uint_as_float(dm[cc]), __uint_as_float(dm[cc + 1]));
          const float2 L = make_float2(__uint_as_float(lv[cc]), __uint_as_float(lv[cc + 1]));
          const float2 G = f2fma(D, D, L);
          wq[h] = __funnelshift_l(__float_as_uint(G.y), wq[h], 1);
          wq[h] = __funnelshift_l(__float_as_uint(G.x), wq[h], 1);
          lmin[h] = __vimin3_u32(lmin[h], lv[cc], lv[cc + 1]);
          const float2 Tb = f2fma(L, nt1, t0);
          aq[h] |= !(fabsf(G.x) >= Tb.x) | !(fabsf(G.y) >= Tb.y);
        }
      }
      uint32_t word = wq[0] + (wq[1] << 8) + (wq[2] << 16) + (wq[3] << 24);
      bool fl[4];
#pragma unroll
      for (int h = 0; h < 4; ++h) fl[h] = aq[h] | (lmin[h] < lcut_bits);
      const bool interior = (u >= ub) & (u < ue);
      const uint32_t mask = interior ? 0xffffffffu : edge_mask();
      if (fl[0] | fl[1] | fl[2] | fl[3]) {
        uint32_t aw = 0;
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          if (fl[h]) {
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              const int cc = 8 * h + c;
              const float Dv = __uint_as_float(dm[cc]), Lv = __uint_as_float(lv[cc]);
              const float G = fmaf(Dv, Dv, Lv);
              const bool a = !(fabsf(G) >= fmaf(Lv, nt1s, t0s));
              aw |= (a | (lv[cc] < lcut_bits)) ? (1u << cc) : 0u;
            }
          }
        }
        aw &= mask;
        word &= ~aw;
        list_pairs(aw);
      }
      emit(word & mask, u, interior);
    };

    // 16 columns per fetch
    auto unit16 = [&](auto stage_c, const int u) {
      constexpr int S = decltype(stage_c)::value;
      mbar_wait_a(bar_full + 8 * S, ph);
      tc_fence_after();
      uint32_t word = 0u, aw = 0u;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        uint32_t dm[16], lv[16];
        tc_ld_32x16(tbase + S * 128 + 16 * half, dm);
        tc_ld_32x16(tbase + S * 128 + 64 + 16 * half, lv);
        tc_wait_ld();
        if (half == 1) {  // both halves are in registers: release the TMEM stage
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_a(bar_full + 16 + 8 * S);
        }
        uint32_t wq[4] = {0u, 0u, 0u, 0u}, lmin[4];
        bool aq[4] = {false, false, false, false};
#pragma unroll
        for (int c = 2; c >= 0; c -= 2) {
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            const int cc = 4 * h + c;
            const float2 D = make_float2(__uint_as_float(dm[cc]), __uint_as_float(dm[cc + 1]));
            const float2 L = make_float2(__uint_as_float(lv[cc]), __uint_as_float(lv[cc + 1]));
            const float2 G = f2fma(D, D, L);
            wq[h] = __funnelshift_l(__float_as_uint(G.y), wq[h], 1);
            wq[h] = __funnelshift_l(__float_as_uint(G.x), wq[h], 1);
            lmin[h] = __vimin3_u32(c == 2 ? 0xffffffffu : lmin[h], lv[cc], lv[cc + 1]);
            const float2 Tb = f2fma(L, nt1, t0);
            aq[h] |= !(fabsf(G.x) >= Tb.x) | !(fabsf(G.y) >= Tb.y);
          }
        }
        word |= (wq[0] + (wq[1] << 4) + (wq[2] << 8) + (wq[3] << 12)) << (16 * half);
        // first a combined test (one branch per half), per-group flags only on the rare path
        const uint32_t lm = min(__vimin3_u32(lmin[0], lmin[1], lmin[2]), lmin[3]);
        if ((lm < lcut_bits) | aq[0] | aq[1] | aq[2] | aq[3]) {
#pragma unroll
          for (int h = 0; h < 4; ++h) {
            if (aq[h] | (lmin[h] < lcut_bits)) {
#pragma unroll
              for (int c = 0; c < 4; ++c) {
                const int cc = 4 * h + c;
                const float Dv = __uint_as_float(dm[cc]), Lv = __uint_as_float(lv[cc]);
                const float G = fmaf(Dv, Dv, Lv);
                const bool a = !(fabsf(G) >= fmaf(Lv, nt1s, t0s));
                aw |= (a | (lv[cc] < lcut_bits)) ? (1u << (16 * half + cc)) : 0u;
              }
            }
          }
        }
      }
      const bool interior = (u >= ub) & (u < ue);
      const uint32_t mask = interior ? 0xffffffffu : edge_mask();
      if (aw) {
        aw &= mask;
        word &= ~aw;
        list_pairs(aw);
      }
      emit(word & mask, u, interior);
    };

#pragma unroll 1
    for (int u = 0; u < nu; u += 2) {
      if (LD16) {
        unit16(StageC<0>{}, u);
        if (u + 1 < nu) unit16(StageC<1>{}, u + 1);
      } else {
        unit32(StageC<0>{}, u);
        if (u + 1 < nu) unit32(StageC<1>{}, u + 1);
      }
      ph ^= 1u;
    }
    if (i < n && rowdeg) atomicAdd(&d.deg[i], rowdeg);
    __syncwarp();
    const int cnt = min(g->amb_cnt[warp], kTcAmbBuf);
    if (cnt > 0) {
      unsigned base = 0;
      if (lane == 0) base = atomicAdd(&d.st->amb_count, (unsigned)cnt);
      base = __shfl_sync(0xffffffffu, base, 0);
      for (int k = lane; k < cnt; k += 32) {
        if (base + k < (unsigned)d.amb_cap) d.amb_list[base + k] = g->amb_buf[warp][k];
        else d.st->tc_overflow = 1;
      }
    }
    if (STAGED) {
      asm volatile("bar.sync 1, %0;" ::"n"(32 * kTcEpiWarps) : "memory");   // the eight epilogue warps: every word of the item is staged
      tk = tpose_const(lane);
      const int kw = 2 * nu;                          // words per row of this item
      const int wbase = (i0 >> 5) + 2 * v0;           // its first word in the adjacency row
      // rows: warp w flushes rows 16 w .. 16 w + 15, 128-byte stores
#pragma unroll 4
      for (int rr = 0; rr < 16; ++rr) {
        const int r2 = 16 * warp + rr, i2 = i0 + r2;
#pragma unroll
        for (int h = 0; h < kSW / 32; ++h) {
          const uint32_t x = stage_w[kSW * r2 + 32 * h + (lane ^ (r2 & 31))];
          if (i2 < n && 32 * h + lane < kw) {
            const size_t off = (size_t)i2 * d.pitch + wbase + 32 * h + lane;
            d.adj[off] = x;
            if (PEERS) for (int pp = 0; pp < kp.n_peer; ++pp) kp.peer_adj[pp][off] = x;   // the same 128-byte row segment into every peer's matrix (NVLink)
          }
        }
      }
      // mirrored blocks: chunk k = 2 u + ch holds columns jk .. jk + 31; lane c of the transposes owns column jk + c
#pragma unroll
      for (int kk = 0; kk < kSW / 8; ++kk) {
        const int k = warp + 8 * kk;
        if (k >= kw || (k >> 1) < ub) continue;       // beyond the item / diagonal tile (its rows are their own mirror)
        uint32_t x[4];
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) x[qq] = stage_w[kSW * (32 * qq + lane) + (k ^ lane)];
#pragma unroll
        for (int qq = 0; qq < 4; ++qq) x[qq] = warp_transpose32_fast(x[qq], tk);
        const int j = i0 + v0 * kTcUnit + 32 * k + lane;
        if (j < n) {
          const size_t off = (size_t)j * d.pitch + (i0 >> 5);
          *reinterpret_cast<uint4*>(d.adj + off) = make_uint4(x[0], x[1], x[2], x[3]);
          if (PEERS) for (int pp = 0; pp < kp.n_peer; ++pp) *reinterpret_cast<uint4*>(kp.peer_adj[pp] + off) = make_uint4(x[0], x[1], x[2], x[3]);
          const int c = __popc(x[0]) + __popc(x[1]) + __popc(x[2]) + __popc(x[3]);
          if (c) atomicAdd(d.deg + j, c);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kTcEpiWarps) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTcTmemCols) : "memory");
  }
}

// B2R_G6 (A/B runs): 1 = direct stores, 16-column fetches; 2 = staged stores, 16-column fetches; 3 = staged stores, 32-column
// fetches (default).  64 x 5000 graph stage on B200: 0.539 / 0.521 / 0.500 ms
static int g6_variant() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B2R_G6");
    v = (e && e[0] >= '1' && e[0] <= '3') ? e[0] - '0' : 3;
  }
  return v;
}

int launch_tim_graph_tc(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, cudaStream_t st) {
  static PerDeviceOnce attr;
  attr.run([] {
    cudaError_t e = cudaFuncSetAttribute(tim_graph6_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(tim_graph6_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(tim_graph6_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes);
    return e;
  });
  const int T = (n_max + kTile - 1) / kTile;
  // prep + operand rows in one launch: about one CTA per SM (every slice re-reduces the problem's bounding boxes), never more
  // slices than 128-row blocks
  int slices = 148 / batch;
  if (slices > T) slices = T;
  if (slices < 1) slices = 1;
  graph_tc_setup_kernel<<<dim3(slices, batch), 1024, 0, st>>>(d_desc, kp);
  const dim3 grid(tc_items(T), batch);
  switch (g6_variant()) {
    case 1: tim_graph6_kernel<true, false><<<grid, kTcThreads, kTcSmemBytes, st>>>(d_desc, kp); break;
    case 2: tim_graph6_kernel<true, true><<<grid, kTcThreads, kTcSmemBytes, st>>>(d_desc, kp); break;
    default: tim_graph6_kernel<false, true><<<grid, kTcThreads, kTcSmemBytes, st>>>(d_desc, kp); break;
  }
  return 2;
}

// Exact decision of this rank's listed pairs on every copy of the matrix: bits only (atomicOr locally and over NVLink); the degrees
// of a sharded solve are recomputed from the complete matrix after the exchange.
__global__ void __launch_bounds__(256) graph_fix_peers_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  const ProbDesc d = descs[0];
  ProbState* st = d.st;
  if (!st->use32 || st->tc_overflow) return;
  const unsigned cnt = min(st->amb_count, (unsigned)d.amb_cap);
  for (unsigned idx = blockIdx.x * 256 + threadIdx.x; idx < cnt; idx += gridDim.x * 256) {
    const uint32_t e = d.amb_list[idx];
    const FixPair p = load_listed_pair(d.src, d.dst, e);
    const size_t i = e >> 16, j = e & 0xffffu;
    const double A = sq3_ref(__dsub_rn(p.s[3], p.s[0]), __dsub_rn(p.s[4], p.s[1]), __dsub_rn(p.s[5], p.s[2]));
    const double B = sq3_ref(__dsub_rn(p.t[3], p.t[0]), __dsub_rn(p.t[4], p.t[1]), __dsub_rn(p.t[5], p.t[2]));
    if (!verdict_exact(A, B, kp.beta)) continue;
    const bool both = (i >> 7) != (j >> 7);
    for (int pp = -1; pp < kp.n_peer; ++pp) {
      uint32_t* adj = pp < 0 ? d.adj : kp.peer_adj[pp];
      atomicOr(adj + i * d.pitch + (j >> 5), 1u << (j & 31));
      if (both) atomicOr(adj + j * d.pitch + (i >> 5), 1u << (i & 31));
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) st->slow_pairs = cnt;
}

int launch_tim_graph_tc_sharded(const ProbDesc* d_desc, int n, const KernelParams& kp, cudaStream_t st) {
  static PerDeviceOnce attr;
  attr.run([] { return cudaFuncSetAttribute(tim_graph6_kernel<false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kTcSmemBytes); });
  const int T = (n + kTile - 1) / kTile;
  int slices = 148;
  if (slices > T) slices = T;
  graph_tc_setup_kernel<<<dim3(slices, 1), 1024, 0, st>>>(d_desc, kp);
  const int w = kp.shard_world > 1 ? kp.shard_world : 1;
  const dim3 grid((tc_items(T) + w - 1) / w, 1);
  tim_graph6_kernel<false, true, true><<<grid, kTcThreads, kTcSmemBytes, st>>>(d_desc, kp);
  graph_fix_peers_kernel<<<64, 256, 0, st>>>(d_desc, kp);
  return 3;
}

int launch_tim_graph_tc_and_fix(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, cudaStream_t st) {
  return launch_tim_graph_tc(d_desc, batch, n_max, kp, st) + launch_graph_fix(d_desc, batch, n_max, kp, st);
}

}  // namespace b2r

// ---------------------------------------------------------------------------------------------------------------------
// Test hook (b2r_debug_tc_values): the raw Dm~ and L~ of one 128 x 64 unit, exactly as the filter kernel's tcgen05.mma
// leaves them in TMEM, so that tests can measure the tensor-core error against the bound the band assumes.
// ---------------------------------------------------------------------------------------------------------------------
namespace b2r {

__global__ void __launch_bounds__(128, 1) graph_tc_debug_kernel(const ProbDesc* __restrict__ descs, int prob, int bi, int v, float* __restrict__ out_dm,
                                                                  float* __restrict__ out_l) {
  const ProbDesc d = descs[prob];
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* sm = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  unsigned char* sRowL = sm;
  unsigned char* sRowD = sm + kTcRowBytes;
  unsigned char* sCol = sm + 2 * kTcRowBytes;
  unsigned char* sStage = sCol + kTcColBytes;
  TcSmemCtl* g = reinterpret_cast<TcSmemCtl*>(sStage + kTcRowBytes);
  const int t = threadIdx.x, warp = t >> 5;
  const unsigned char* ops = reinterpret_cast<const unsigned char*>(d.pts32);
  const int i0 = bi * kTile;
  const int j0 = (bi + (v >> 1)) * kTile + (v & 1) * kTcUnit;
  if (t == 0) {
    mbar_init(&g->rowfull, 1);
    mbar_init(&g->full[0], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_expect_tx(&g->rowfull, kTcRowBytes + kTcColBytes);
    bulk_g2s(sStage, ops + (size_t)i0 * 128, kTcRowBytes, &g->rowfull);
    bulk_g2s(sCol, ops + (size_t)j0 * 128, kTcColBytes, &g->rowfull);
  }
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&g->tmem_base)), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  mbar_wait(&g->rowfull, 0);
  rows_from_col_row(sStage, sRowL, sRowD, t);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = g->tmem_base;
  if (t == 0) {
    constexpr uint32_t idesc = umma_idesc_tf32(kTcRowTile, kTcUnit);
    const uint32_t rowl = smem_u32(sRowL), rowd = smem_u32(sRowD), cb = smem_u32(sCol);
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) tc_mma_tf32(tmem, umma_desc_sw128(rowd) + 2 * ks, umma_desc_sw128(cb) + 2 * ks, idesc, ks != 0 ? 1u : 0u);
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) tc_mma_tf32(tmem + 64, umma_desc_sw128(rowl) + 2 * ks, umma_desc_sw128(cb) + 2 * ks, idesc, ks != 0 ? 1u : 0u);
    tc_commit(&g->full[0]);
  }
  mbar_wait(&g->full[0], 0);
  tc_fence_after();
#pragma unroll 1
  for (int ch = 0; ch < 2; ++ch) {
    uint32_t dm[32], lv[32];
    const uint32_t ta = tmem + ((uint32_t)(32 * warp) << 16) + (uint32_t)(ch * 32);
    tc_ld_32x32(ta, dm);
    tc_ld_32x32(ta + 64, lv);
    tc_wait_ld();
#pragma unroll
    for (int c = 0; c < 32; ++c) {
      out_dm[t * kTcUnit + ch * 32 + c] = __uint_as_float(dm[c]);
      out_l[t * kTcUnit + ch * 32 + c] = __uint_as_float(lv[c]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u) : "memory");
  }
}

int launch_graph_tc_debug(const ProbDesc* d_desc, int prob, int bi, int v, float* out_dm, float* out_l, cudaStream_t st) {
  static PerDeviceOnce attr;
  attr.run([] { return cudaFuncSetAttribute(graph_tc_debug_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024); });
  graph_tc_debug_kernel<<<1, 128, 64 * 1024, st>>>(d_desc, prob, bi, v, out_dm, out_l);
  return 1;
}

}  // namespace b2r

// ---------------------------------------------------------------------------------------------------------------------
// Test hook (b2r_debug_tc_sweep): EVERY unit of one problem is recomputed on the tensor cores exactly as the filter kernel does and
// compared on the device with fp64 -- (a) the exact sum of the same operand products (the accumulation error the band ASSUMES to be
// below 2^-20 of the sum of |products|), (b) the values from the unrounded coordinates (the total error the band bounds by E_D, E_L).
// grid = T (T + 1) units, 128 threads (thread = row).  out: [0] max accumulation ratio, [1] max |Dm~-Dm| / E_D, [2] max |L~-L| / E_L,
// [3] pairs checked, [4] Pi, [5] E_D, [6] E_L.
// ---------------------------------------------------------------------------------------------------------------------
namespace b2r {

namespace {
__device__ __forceinline__ float sw_elem(const unsigned char* tile, int r, int k) {
  return *reinterpret_cast<const float*>(tile + r * 128 + ((((k >> 2) ^ (r & 7))) << 4) + (k & 3) * 4);
}
__device__ __forceinline__ void atomic_max_nonneg(double* addr, double v) {
  atomicMax(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)__double_as_longlong(v));
}
}  // namespace

__global__ void __launch_bounds__(128, 1) graph_tc_sweep_kernel(const ProbDesc* __restrict__ descs, int prob, double* __restrict__ out) {
  const ProbDesc d = descs[prob];
  const ProbState* st = d.st;
  const int n = d.n;
  const int T = (n + kTile - 1) / kTile;
  int p = blockIdx.x, bi = 0;
  while (bi < T && p >= 2 * (T - bi)) { p -= 2 * (T - bi); ++bi; }
  if (bi >= T) return;
  const int v = p;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* sm = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  unsigned char* sRowL = sm;
  unsigned char* sRowD = sm + kTcRowBytes;
  unsigned char* sCol = sm + 2 * kTcRowBytes;
  unsigned char* sStage = sCol + kTcColBytes;
  TcSmemCtl* g = reinterpret_cast<TcSmemCtl*>(sStage + kTcRowBytes);
  const int t = threadIdx.x, warp = t >> 5;
  const unsigned char* ops = reinterpret_cast<const unsigned char*>(d.pts32);
  const int i0 = bi * kTile;
  const int j0 = i0 + v * kTcUnit;
  if (t == 0) {
    mbar_init(&g->rowfull, 1);
    mbar_init(&g->full[0], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_expect_tx(&g->rowfull, kTcRowBytes + kTcColBytes);
    bulk_g2s(sStage, ops + (size_t)i0 * 128, kTcRowBytes, &g->rowfull);
    bulk_g2s(sCol, ops + (size_t)j0 * 128, kTcColBytes, &g->rowfull);
  }
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&g->tmem_base)), "r"(128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  mbar_wait(&g->rowfull, 0);
  rows_from_col_row(sStage, sRowL, sRowD, t);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = g->tmem_base;
  if (t == 0) {
    constexpr uint32_t idesc = umma_idesc_tf32(kTcRowTile, kTcUnit);
    const uint32_t rowl = smem_u32(sRowL), rowd = smem_u32(sRowD), cb = smem_u32(sCol);
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) tc_mma_tf32(tmem, umma_desc_sw128(rowd) + 2 * ks, umma_desc_sw128(cb) + 2 * ks, idesc, ks != 0 ? 1u : 0u);
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) tc_mma_tf32(tmem + 64, umma_desc_sw128(rowl) + 2 * ks, umma_desc_sw128(cb) + 2 * ks, idesc, ks != 0 ? 1u : 0u);
    tc_commit(&g->full[0]);
  }
  mbar_wait(&g->full[0], 0);
  tc_fence_after();
  const int i = i0 + t;
  const double kappa = st->kappa;
  const double Pi = (double)st->tc_pi;
  const double w22 = 2.384185791015625e-7;
  const double E_L = w22 * (71.8 * Pi + 4.3), E_D = 45.9 * w22 * Pi;
  double si[3] = {0, 0, 0}, di[3] = {0, 0, 0};
  if (i < n) {
#pragma unroll
    for (int k = 0; k < 3; ++k) { si[k] = (d.src[3 * (size_t)i + k] - st->ctr[k]) * kappa; di[k] = (d.dst[3 * (size_t)i + k] - st->ctr[3 + k]) * kappa; }
  }
  double worst_tc = 0.0, worst_d = 0.0, worst_l = 0.0;
  unsigned checked = 0;
#pragma unroll 1
  for (int ch = 0; ch < 2; ++ch) {
    uint32_t dm[32], lv[32];
    const uint32_t ta = tmem + ((uint32_t)(32 * warp) << 16) + (uint32_t)(ch * 32);
    tc_ld_32x32(ta, dm);
    tc_ld_32x32(ta + 64, lv);
    tc_wait_ld();
#pragma unroll 1
    for (int c = 0; c < 32; ++c) {
      const int cj = ch * 32 + c, j = j0 + cj;
      if (i >= n || j >= n) continue;
      double sd = 0.0, ad = 0.0, sl = 0.0, al = 0.0;
      for (int k = 0; k < 32; ++k) {
        const double col = (double)sw_elem(sCol, cj, k);
        const double pd = (double)sw_elem(sRowD, t, k) * col, pl = (double)sw_elem(sRowL, t, k) * col;
        sd += pd; ad += fabs(pd); sl += pl; al += fabs(pl);
      }
      const double tdm = (double)__uint_as_float(dm[c]), tl = (double)__uint_as_float(lv[c]);
      if (ad > 0.0) worst_tc = fmax(worst_tc, fabs(tdm - sd) / ad);
      if (al > 0.0) worst_tc = fmax(worst_tc, fabs(tl - sl) / al);
      double A = 0.0, B = 0.0;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const double a = (d.src[3 * (size_t)j + k] - st->ctr[k]) * kappa - si[k];
        const double b = (d.dst[3 * (size_t)j + k] - st->ctr[3 + k]) * kappa - di[k];
        A += a * a; B += b * b;
      }
      worst_d = fmax(worst_d, fabs(tdm - (A - B)) / E_D);
      worst_l = fmax(worst_l, fabs(tl - (1.0 - 2.0 * (A + B))) / E_L);
      ++checked;
    }
  }
  atomic_max_nonneg(out + 0, worst_tc);
  atomic_max_nonneg(out + 1, worst_d);
  atomic_max_nonneg(out + 2, worst_l);
  if (checked) atomicAdd(out + 3, (double)checked);
  if (blockIdx.x == 0 && t == 0) { out[4] = Pi; out[5] = E_D; out[6] = E_L; }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128u) : "memory");
  }
}

int launch_graph_tc_sweep(const ProbDesc* d_desc, int prob, int n, const KernelParams& kp, double* out, cudaStream_t st) {
  static PerDeviceOnce attr;
  attr.run([] { return cudaFuncSetAttribute(graph_tc_sweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024); });
  (void)kp;
  const int T = (n + kTile - 1) / kTile;
  graph_tc_sweep_kernel<<<T * (T + 1), 128, 64 * 1024, st>>>(d_desc, prob, out);
  return 1;
}

}  // namespace b2r
