// Nearest neighbour in feature space on the tensor cores (north-star item (b)).
//
// Replaces the feature-matching step of get_correspondences (HelpersRealtime.py:194-213 /
// registration_utils.py:177-196) in its north-star form: nn[i] = argmin_j ||fq_i - fr_j||^2, both directions, then the
// mutual filter (presteps.cu).  The one dense contraction of the path, -2 * Fq * Fr^T, runs as a tcgen05 GEMM:
//
//   prep      f64 descriptors -> f32 -> 3xTF32 split, K-concatenated so that ONE GEMM with K = 3*40 (padded to 128)
//             yields hi*hi + hi*lo + lo*hi (fp32-class accuracy from TF32 tensor cores), plus the squared norms
//   gemm      one CTA per (128-query tile, range of 128-reference tiles): TMA (cp.async.bulk.tensor, 128B swizzle)
//             stages the operands, one thread issues tcgen05.mma.kind::tf32 (M=128, N=128, K=8) into double-buffered
//             TMEM accumulators, four epilogue warps read them back with tcgen05.ld and keep a per-row top-4 of
//             ||fr_j||^2 - 2 fq_i.fr_j in registers -- the distance matrix never exists in memory
//   finish    one warp per query: candidates of all column ranges, shortlist within the proven error margin of the
//             approximate minimum, EXACT fp64 re-rank in the oracle's order (sequential sum of (a-b)^2, lowest index
//             wins ties); rows whose shortlist cannot be proven complete are redone by exact brute force.
// Result: indices identical to the exact fp64 search (tests compare with the oracle bit for bit).
#include <cuda.h>
#include <math_constants.h>

#include "solver.cuh"
#include "../../include/b2r_debug.h"

namespace b2r {

namespace {

constexpr int kBM = 128;          // queries per CTA
constexpr int kBN = 128;          // references per MMA tile
constexpr int kKSeg = 40;         // one TF32 segment: feature dim padded to a multiple of 8
constexpr int kKTot = 128;        // 3 segments (120) padded to four 128-byte swizzle atoms
constexpr int kKBlocks = 4;       // 32 floats (128 B) each
constexpr int kTopK = 4;
constexpr int kMaxSplits = 16;
constexpr int kEpiWarps = 8;      // lane quarter x column half of the 128 x 128 accumulator
constexpr int kGemmThreads = 64 + 32 * kEpiWarps; // warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2..9 epilogue
constexpr uint32_t kTileBytes = kBM * 128;                 // one k-block of one operand tile: 128 rows x 128 B
constexpr uint32_t kOperandBytes = kKBlocks * kTileBytes;  // 64 KB
// |D~_ij - D_ij| <= kEpsRel * (|fq_i - c|^2 + |fr_j - c|^2) for descriptors centred on c (nn_centre_kernel).  With u = 2^-24 and
// x.y = sum_k x_k y_k of the centred fp64 rows (sum|x_k y_k| <= (nq + nr) / 2):
//   fp32 copy of a centred coordinate x' = x (1 + d), |d| <= u                          -> 2u              sum|xy|
//   3xTF32 split x' = hi + lo + r, |r| <= 2^-22 |x'|; dropped lo.lo and r terms          -> 3 * 2^-22       sum|xy|
//   products of two TF32 values are exact in fp32; fp32 accumulation on the tensor core over K = 128 (16 chained K = 8
//   instructions): ASSUMED within 2^-19 of sum|products| -- measured 2^-22.2 for the 4-instruction chain of the graph filter
//   (tests/test_gpu_tc_sweep.py), i.e. 2.3x margin over a linear growth with the chain length; the total is MEASURED on every
//   candidate pair of a run by b2r_debug_nn_gemm_error (tests/test_gpu_presteps.py)       -> 2^-19 (1 + 2^-10) sum|xy|
//   => |acc - x.y| <= 2.75e-6 sum|xy|, times 2 for -2 x.y, with sum|xy| <= (nq + nr) / 2  -> 2.75e-6 (nq + nr)
//   norms rounded up to fp32 (2^-23 each), final fma rounding u |v| <= 2u (nq + nr)       -> 0.25e-6 (nq + nr)
//   = 3.0e-6 (nq + nr)  ->  3.2e-6.
// The epilogue ranks references by the LOWER BOUND v_lo = fma(-2, acc, nr_lo), nr_lo = rd(nr (1 - kEpsRel)): the reference's own
// share of the error is folded into the value it is ranked by, so the completeness test of a row needs no global norm maximum
// (one far-away descriptor no longer widens every row's margin; nn_fallback_rows on the config-2 frame 31 % -> ~1 %).
constexpr double kEpsRel = 3.2e-6;
constexpr int kCentreRows = 4096;   // rows of the reference cloud whose mean is the common centre c

// ---- PTX wrappers ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
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
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint64_t* bar, void* dst, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)),
               "l"((uint64_t)map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
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
// 32 lanes x 32 columns of 32-bit accumulators -> 32 registers per thread (thread = lane/row, register = column)
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
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand tile with 128-byte swizzle (the layout TMA writes): rows 128 B apart, 8-row groups 1024 B apart.
// bits [0,14) start address >> 4, [16,30) LBO >> 4 (= 1, ignored for swizzled K-major), [32,46) SBO >> 4 (= 64),
// [46,48) descriptor version 1 (Blackwell), [61,64) layout type 2 = SWIZZLE_128B.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// kind::tf32, fp32 accumulate, A and B K-major: c_format [4,6) = 1, a_format [7,10) = 2, b_format [10,13) = 2,
// n >> 3 at [17,23), m >> 4 at [24,29)
__device__ __forceinline__ constexpr uint32_t umma_idesc_tf32(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ float to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

// ---- prep -----------------------------------------------------------------------------------------------------------
// Common centre of both descriptor sets: the mean of the first kCentreRows reference rows (one CTA, fixed summation order, so the
// centre -- and with it every rounded operand -- is the same on every run).  Distances do not change under a common shift;
// the error bound is relative to the CENTRED norms, which for FPFH histograms (every row sums to 300, norms ~ 1e4) are one to two
// orders smaller than the raw ones.
__global__ void __launch_bounds__(1024) nn_centre_kernel(const double* __restrict__ f, int n, int dim, double* __restrict__ centre) {
  __shared__ double part[32][kKSeg];
  const int rows = n < kCentreRows ? n : kCentreRows;
  const int t = threadIdx.x % kKSeg, r0 = threadIdx.x / kKSeg;      // 1024 / 40 = 25 row lanes (threads 1000..1023 idle)
  double s = 0.0;
  if (r0 < 25 && t < dim)
    for (int r = r0; r < rows; r += 25) s += f[(size_t)r * dim + t];
  if (r0 < 25) part[r0][t] = s;
  __syncthreads();
  if (threadIdx.x < kKSeg) {
    double a = 0.0;
    for (int k = 0; k < 25; ++k) a += part[k][threadIdx.x];
    centre[threadIdx.x] = (threadIdx.x < dim && rows > 0) ? a / rows : 0.0;
  }
}

// out: [n_pad][128] floats.  role 0 (left operand): [hi | hi | lo | 0]; role 1 (right operand): [hi | lo | hi | 0].
__global__ void __launch_bounds__(256) nn_prep_kernel(const double* __restrict__ f, int n, int dim, int n_pad, const double* __restrict__ centre,
                                                      float* __restrict__ out_l, float* __restrict__ out_r, float* __restrict__ norm,
                                                      float* __restrict__ norm_lo) {
  const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
  if (e >= (long long)n_pad * kKTot) return;
  const int row = (int)(e / kKTot), k = (int)(e % kKTot);
  const int seg = k / kKSeg, t = k % kKSeg;
  float x = 0.f;
  if (row < n && t < dim && seg < 3) x = (float)(f[(size_t)row * dim + t] - centre[t]);
  const float hi = to_tf32(x);
  const float lo = to_tf32(x - hi);
  out_l[e] = seg == 2 ? lo : hi;
  out_r[e] = seg == 1 ? lo : hi;
  if (k == 0) {
    float nv = CUDART_INF_F;
    if (row < n) {
      double s = 0;
      for (int q = 0; q < dim; ++q) { const double v = f[(size_t)row * dim + q] - centre[q]; s += v * v; }
      nv = __double2float_ru(s);                 // rounded up: the norms only enter error bounds and the approximate distance
    }
    norm[row] = nv;
    norm_lo[row] = row < n ? __double2float_rd((double)nv * (1.0 - kEpsRel)) : CUDART_INF_F;
  }
}

// ---- gemm + top-k epilogue --------------------------------------------------------------------------------------------
struct GemmSmem {
  uint64_t a_full, b_full[2], b_empty[2], acc_full[2], acc_empty[2];
  uint32_t tmem_base;
  alignas(16) float nr[2][kBN];
};

__global__ void __launch_bounds__(kGemmThreads, 1)
nn_gemm_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_r, const float* __restrict__ norm_r_lo,
               int n_ref, int col_tiles, int splits, float* __restrict__ cand_v, int* __restrict__ cand_j) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space (LDS, not LD.E)
  unsigned char* sA = sm;
  unsigned char* sB = sm + kOperandBytes;                 // 2 stages
  GemmSmem* g = reinterpret_cast<GemmSmem*>(sm + 3 * kOperandBytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row_tile = blockIdx.x, split = blockIdx.y;
  const int per = (col_tiles + splits - 1) / splits;
  const int c0 = split * per;
  const int c1 = min(c0 + per, col_tiles);
  const int nt = max(c1 - c0, 0);

  if (threadIdx.x == 0) {
    mbar_init(&g->a_full, 1);
    for (int s = 0; s < 2; ++s) { mbar_init(&g->b_full[s], 1); mbar_init(&g->b_empty[s], 1); mbar_init(&g->acc_full[s], 1); mbar_init(&g->acc_empty[s], kEpiWarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // TMEM: 2 accumulator stages x 128 columns
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&g->tmem_base)), "r"(256u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = g->tmem_base;

  if (warp == 0) {
    if (lane == 0) {  // ===== TMA producer =====
      if (nt > 0) mbar_expect_tx(&g->a_full, kOperandBytes);
      for (int kb = 0; kb < kKBlocks && nt > 0; ++kb) tma_load_2d(&map_q, &g->a_full, sA + kb * kTileBytes, kb * 32, row_tile * kBM);
      for (int t = 0; t < nt; ++t) {
        const int s = t & 1;
        const uint32_t ph = (t >> 1) & 1;
        mbar_wait(&g->b_empty[s], ph ^ 1);
        mbar_expect_tx(&g->b_full[s], kOperandBytes);
        for (int kb = 0; kb < kKBlocks; ++kb)
          tma_load_2d(&map_r, &g->b_full[s], sB + s * kOperandBytes + kb * kTileBytes, kb * 32, (c0 + t) * kBN);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {  // ===== MMA issuer =====
      constexpr uint32_t idesc = umma_idesc_tf32(kBM, kBN);
      if (nt > 0) mbar_wait(&g->a_full, 0);
      for (int t = 0; t < nt; ++t) {
        const int s = t & 1;
        const uint32_t ph = (t >> 1) & 1;
        mbar_wait(&g->b_full[s], ph);
        mbar_wait(&g->acc_empty[s], ph ^ 1);
        tc_fence_after();
        const uint32_t a_base = smem_u32(sA), b_base = smem_u32(sB + s * kOperandBytes);
#pragma unroll
        for (int kb = 0; kb < kKBlocks; ++kb) {
          const uint64_t ad = umma_desc_sw128(a_base + kb * kTileBytes), bd = umma_desc_sw128(b_base + kb * kTileBytes);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)  // 8 tf32 = 32 bytes per step inside the 128-byte swizzle row: +2 in the encoded address
            tc_mma_tf32(tmem + s * kBN, ad + 2 * ks, bd + 2 * ks, idesc, (kb | ks) != 0 ? 1u : 0u);
        }
        tc_commit(&g->b_empty[s]);   // operands consumed -> the producer may refill the stage
        tc_commit(&g->acc_full[s]);  // accumulator complete -> epilogue
      }
    }
  } else {  // ===== epilogue: warps 2..9; warp & 3 = TMEM lane quarter it may read, (warp - 2) >> 2 = column half of every tile =====
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = quarter * 32 + lane;
    const int e = threadIdx.x - 64;
    float v0 = CUDART_INF_F, v1 = CUDART_INF_F, v2 = CUDART_INF_F, v3 = CUDART_INF_F;
    int j0 = -1, j1 = -1, j2 = -1, j3 = -1;
    for (int t = 0; t < nt; ++t) {
      const int s = t & 1;
      const uint32_t ph = (t >> 1) & 1;
      const int jbase = (c0 + t) * kBN;
      if (e < kBN) g->nr[s][e] = (jbase + e < n_ref) ? norm_r_lo[jbase + e] : CUDART_INF_F;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      mbar_wait(&g->acc_full[s], ph);
      tc_fence_after();
#pragma unroll 1
      for (int ch = 0; ch < kBN / 64; ++ch) {
        uint32_t acc[32];
        const int cb = half * 64 + ch * 32;
        tc_ld_32x32(tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(s * kBN + cb), acc);
        float nrv[32];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const float4 q4 = *reinterpret_cast<const float4*>(&g->nr[s][cb + 4 * c]);
          nrv[4 * c] = q4.x; nrv[4 * c + 1] = q4.y; nrv[4 * c + 2] = q4.z; nrv[4 * c + 3] = q4.w;
        }
        tc_wait_ld();
        // v = lower bound of D_ij - nq_i (up to eps nq_i).  Fast path: one min over the 32 values against the row's 4th best;
        // the insertion code runs only for chunks in which some row of the warp improves its list.
        float m = CUDART_INF_F;
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          nrv[c] = fmaf(-2.f, __uint_as_float(acc[c]), nrv[c]);
          m = fminf(m, nrv[c]);
        }
        if (m < v3) {
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            const float v = nrv[c];
            if (v < v3) {
              v3 = v; j3 = jbase + cb + c;
              if (v3 < v2) { const float tv = v2; v2 = v3; v3 = tv; const int tj = j2; j2 = j3; j3 = tj; }
              if (v2 < v1) { const float tv = v1; v1 = v2; v2 = tv; const int tj = j1; j1 = j2; j2 = tj; }
              if (v1 < v0) { const float tv = v0; v0 = v1; v1 = tv; const int tj = j0; j0 = j1; j1 = tj; }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&g->acc_empty[s]);
    }
    const size_t o = ((size_t)(row_tile * kBM + row) * (2 * splits) + 2 * split + half) * kTopK;
    *reinterpret_cast<float4*>(cand_v + o) = make_float4(v0, v1, v2, v3);
    *reinterpret_cast<int4*>(cand_j + o) = make_int4(j0, j1, j2, j3);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(256u) : "memory");
  }
}

// ---- finish: shortlist, exact re-rank, fallback -------------------------------------------------------------------------
__device__ __forceinline__ double exact_d2(const double* __restrict__ a, const double* __restrict__ b, int dim) {
  double s = 0;
  for (int t = 0; t < dim; ++t) { const double d = __dsub_rn(a[t], b[t]); s = __dadd_rn(s, __dmul_rn(d, d)); }
  return s;
}

constexpr int kFinWarps = 4;

// groups = 2 * splits candidate groups of kTopK per row (column range x column half), each sorted ascending by v_lo.
// err_max (debug, may be NULL): max over ALL candidates of |(nq + v) - D| / (nq + nr_j) as float bits, v = v_lo + (nr - nr_lo).
__global__ void __launch_bounds__(kFinWarps * 32) nn_finish_kernel(const double* __restrict__ fq, int nq, const double* __restrict__ fr, int nr,
                                                                   int dim, const float* __restrict__ norm_q, const float* __restrict__ norm_r,
                                                                   const float* __restrict__ norm_r_lo, const float* __restrict__ cand_v,
                                                                   const int* __restrict__ cand_j, int groups, int* __restrict__ nn,
                                                                   double* __restrict__ dist, unsigned int* __restrict__ fb_count, int* __restrict__ fb_list,
                                                                   unsigned int* __restrict__ err_max) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = blockIdx.x * kFinWarps + warp;
  if (i >= nq) return;
  const int nc = groups * kTopK;
  const double nqi = (double)norm_q[i];
  const double eq = kEpsRel * nqi;
  // D_ij - nq_i lies in [v_lo - eq, v_lo + eq + 2.001 eps nr_j] for a kept reference and above kept - eq for every reference a
  // group did NOT keep (kept = the group's last value; +inf while the group still had room).
  double min_ub = CUDART_INF, kept = CUDART_INF;
  bool any = false;
  for (int c = lane; c < nc; c += 32) {
    const int j = cand_j[(size_t)i * nc + c];
    const double v = (double)cand_v[(size_t)i * nc + c];
    if (j >= 0) { any = true; min_ub = fmin(min_ub, v + eq + 2.001 * kEpsRel * (double)norm_r[j]); }
    if ((c & (kTopK - 1)) == kTopK - 1 && j >= 0) kept = fmin(kept, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    min_ub = fmin(min_ub, __shfl_xor_sync(0xffffffffu, min_ub, o));
    kept = fmin(kept, __shfl_xor_sync(0xffffffffu, kept, o));
  }
  any = __any_sync(0xffffffffu, any);
  double best = CUDART_INF; int bj = -1;
  const bool complete = any && (kept - eq > min_ub);     // false for NaN
  if (complete || err_max) {
    float worst = 0.f;
    for (int c = lane; c < nc; c += 32) {
      const int j = cand_j[(size_t)i * nc + c];
      if (j < 0) continue;
      const double v = (double)cand_v[(size_t)i * nc + c];
      const bool listed = complete && (v - eq <= min_ub);
      if (listed || err_max) {
        const double d = exact_d2(fq + (size_t)i * dim, fr + (size_t)j * dim, dim);
        if (listed && (d < best || (d == best && (unsigned)j < (unsigned)bj))) { best = d; bj = j; }
        if (err_max) {
          const double nrj = (double)norm_r[j];
          const double approx = nqi + v + (nrj - (double)norm_r_lo[j]);
          worst = fmaxf(worst, (float)(fabs(approx - d) / (nqi + nrj + 1e-300)));
        }
      }
    }
    if (err_max) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) worst = fmaxf(worst, __shfl_xor_sync(0xffffffffu, worst, o));
      if (lane == 0) atomicMax(err_max, __float_as_uint(worst));
    }
  }
  if (!complete) {  // warp-uniform: the row goes to nn_fallback_kernel (exact search over every reference, reference tiles shared by 8 rows)
    if (lane == 0) fb_list[atomicAdd(fb_count, 1u)] = i;
    return;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double od = __shfl_xor_sync(0xffffffffu, best, o); const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
    if (od < best || (od == best && (unsigned)oj < (unsigned)bj)) { best = od; bj = oj; }
  }
  if (lane == 0) { nn[i] = bj; if (dist) dist[i] = best; }
}

// Rows whose shortlist could not be proven complete: exact fp64 search over EVERY reference.  Eight rows per CTA (one per warp) share
// each 128-reference tile through shared memory, so a reference row is read from L2 once per eight rows instead of once per row
// (the in-warp loop this replaces moved 264 B x references x rows through L2: 0.3 ms for the ~800 near-tie rows of a config-2 frame).
// A tile is a straight copy of 128 consecutive descriptor rows (stride = dim doubles: conflict-free for the odd FPFH dimension 33);
// the next tile is already in registers while the current one is evaluated, and a lane carries four references at once (four
// independent summation chains instead of one).
constexpr int kFbChunks = 4;
constexpr int kFbRows = 8, kFbTile = 128, kFbLoads = (kFbTile * kKSeg + kFbRows * 32 - 1) / (kFbRows * 32);

// gridDim.y = chunks of the reference range (whole tiles each): a frame with ~800 fallback rows has ~100 row groups, which leave a
// third of the SMs idle and run 29 tiles one after the other; with four chunks 400 CTAs run 7 - 8 tiles each.  A chunk leaves its
// (distance, index) per row in part_d / part_j; the CTA that arrives last at the group's counter (part_cnt, zeroed by the host)
// merges the chunks in ascending reference order -- the same lowest-index tie rule -- and writes the result.
__global__ void __launch_bounds__(kFbRows * 32) nn_fallback_kernel(const double* __restrict__ fq, const double* __restrict__ fr, int nr, int dim,
                                                                   const int* __restrict__ fb_list, const unsigned int* __restrict__ fb_count,
                                                                   int* __restrict__ nn, double* __restrict__ dist, unsigned int* __restrict__ n_fallback,
                                                                   double* part_d, int* part_j, unsigned int* part_cnt) {
  __shared__ double tile[kFbTile * kKSeg];
  __shared__ double qv[kFbRows][kKSeg];
  __shared__ unsigned int s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned int count = *fb_count;
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 && n_fallback) atomicAdd(n_fallback, count);
  const size_t total = (size_t)nr * dim;
  const int chunks = (int)gridDim.y, chunk = (int)blockIdx.y;
  const int tiles_total = (nr + kFbTile - 1) / kFbTile;
  const int tiles_per_chunk = (tiles_total + chunks - 1) / chunks;
  const int j_begin = min(nr, chunk * tiles_per_chunk * kFbTile), j_end = min(nr, (chunk + 1) * tiles_per_chunk * kFbTile);
  for (unsigned int g0 = blockIdx.x * kFbRows; g0 < count; g0 += gridDim.x * kFbRows) {
    const unsigned int slot = g0 + warp;
    const int i = slot < count ? fb_list[slot] : -1;
    __syncthreads();
    if (i >= 0) for (int t = lane; t < dim; t += 32) qv[warp][t] = fq[(size_t)i * dim + t];
    double best = CUDART_INF; int bj = -1;
    double pre[kFbLoads];
    const int tile_elems = kFbTile * dim;
    const size_t first = (size_t)j_begin * dim;
#pragma unroll
    for (int q = 0; q < kFbLoads; ++q) { const size_t e = (size_t)q * (kFbRows * 32) + threadIdx.x; pre[q] = ((int)e < tile_elems && first + e < total) ? fr[first + e] : 0.0; }
    for (int j0 = j_begin; j0 < j_end; j0 += kFbTile) {
      __syncthreads();
#pragma unroll
      for (int q = 0; q < kFbLoads; ++q) { const int e = q * (kFbRows * 32) + threadIdx.x; if (e < tile_elems) tile[e] = pre[q]; }
      const size_t nbase = (size_t)(j0 + kFbTile) * dim;
#pragma unroll
      for (int q = 0; q < kFbLoads; ++q) { const size_t e = (size_t)q * (kFbRows * 32) + threadIdx.x; pre[q] = ((int)e < tile_elems && nbase + e < total) ? fr[nbase + e] : 0.0; }
      __syncthreads();
      if (i >= 0) {
        const int rows = min(kFbTile, nr - j0);
        double sacc[4] = {0.0, 0.0, 0.0, 0.0};
        for (int t = 0; t < dim; ++t) {
          const double qt = qv[warp][t];
#pragma unroll
          for (int h = 0; h < 4; ++h) { const double dd = __dsub_rn(qt, tile[(lane + 32 * h) * dim + t]); sacc[h] = __dadd_rn(sacc[h], __dmul_rn(dd, dd)); }
        }
#pragma unroll
        for (int h = 0; h < 4; ++h) {       // ascending reference index per lane: h = 0 .. 3
          const int jl = lane + 32 * h;
          const int j = j0 + jl;
          if (jl < rows && (sacc[h] < best || (sacc[h] == best && (unsigned)j < (unsigned)bj))) { best = sacc[h]; bj = j; }
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const double od = __shfl_xor_sync(0xffffffffu, best, o); const int oj = __shfl_xor_sync(0xffffffffu, bj, o);
      if (od < best || (od == best && (unsigned)oj < (unsigned)bj)) { best = od; bj = oj; }
    }
    if (chunks == 1) {
      if (i >= 0 && lane == 0) { nn[i] = bj; if (dist) dist[i] = best; }
      continue;
    }
    if (i >= 0 && lane == 0) { part_d[(size_t)slot * chunks + chunk] = best; part_j[(size_t)slot * chunks + chunk] = bj; }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned int group = g0 / kFbRows;
      const unsigned int old = atomicAdd(&part_cnt[group], 1u);
      s_last = old == (unsigned int)chunks - 1u ? 1u : 0u;
      if (s_last) part_cnt[group] = 0u;
    }
    __syncthreads();
    if (s_last) {   // every chunk of this group has published its partial results
      __threadfence();
      if (i >= 0 && lane == 0) {
        double fb = CUDART_INF; int fj = -1;
        for (int c = 0; c < chunks; ++c) {
          const double od = __ldcg(&part_d[(size_t)slot * chunks + c]); const int oj = __ldcg(&part_j[(size_t)slot * chunks + c]);
          if (od < fb || (od == fb && (unsigned)oj < (unsigned)fj)) { fb = od; fj = oj; }
        }
        nn[i] = fj; if (dist) dist[i] = fb;
      }
    }
  }
}

}  // namespace

}  // namespace b2r

using namespace b2r;

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled() {
  // process-wide (a driver entry point, not per-device state); guarded against two threads resolving it at once
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  static std::mutex mtx;
  std::lock_guard<std::mutex> lock(mtx);
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
    else
      cudaGetLastError();
  }
  return fn;
}

// operand buffer [n_pad][128] f32, row-major; box = 32 floats (one 128-byte swizzle row) x 128 rows
bool make_operand_map(CUtensorMap* m, const float* base, int n_pad) {
  EncodeTiledFn fn = encode_tiled();
  if (!fn) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)kKTot, (cuuint64_t)n_pad};
  const cuuint64_t strides[1] = {(cuuint64_t)kKTot * sizeof(float)};
  const cuuint32_t box[2] = {32, (cuuint32_t)kBM};
  const cuuint32_t estr[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

struct Prepped {
  float* l; float* r; float* norm; float* norm_lo; int n, n_pad;
};

inline size_t pad256(size_t b) { return (b + 255) / 256 * 256; }

size_t prepped_bytes(int n) {
  const size_t n_pad = (size_t)(n + kBM - 1) / kBM * kBM;
  return 2 * pad256(n_pad * kKTot * 4) + 2 * pad256(n_pad * 4);
}

char* carve_prepped(char* p, int n, Prepped* out) {
  const size_t n_pad = (size_t)(n + kBM - 1) / kBM * kBM;
  out->n = n; out->n_pad = (int)n_pad;
  out->l = (float*)p; p += pad256(n_pad * kKTot * 4);
  out->r = (float*)p; p += pad256(n_pad * kKTot * 4);
  out->norm = (float*)p; p += pad256(n_pad * 4);
  out->norm_lo = (float*)p; p += pad256(n_pad * 4);
  return p;
}

int choose_splits(int row_tiles, int col_tiles) {
  int s = 296 / (row_tiles > 0 ? row_tiles : 1);
  static const int forced = [] { const char* e = getenv("B2R_NN_SPLITS"); return e ? atoi(e) : 0; }();   // A/B runs
  if (forced > 0) { s = forced < kMaxSplits ? forced : kMaxSplits; if (s > col_tiles) s = col_tiles; const int per = (col_tiles + s - 1) / s; return (col_tiles + per - 1) / per; }
  // Long column ranges per CTA beat many CTAs: a thread's top-4 list only stops changing after a few thousand values, and until then
  // nearly every 32-value chunk takes the insertion path (32768^2: 1.32 / 1.47 / 1.46 / 1.53 / 1.69 ms for 1 / 2 / 4 / 8 / 16 splits).
  if (s < 1) s = 1;
  if (s > kMaxSplits) s = kMaxSplits;
  if (s > col_tiles) s = col_tiles;
  const int per = (col_tiles + s - 1) / s;
  return (col_tiles + per - 1) / per;  // no empty split
}

size_t cand_bytes(int nq, int nr) {
  const int rt = (nq + kBM - 1) / kBM, ct = (nr + kBN - 1) / kBN;
  const int s = choose_splits(rt, ct);
  return 2 * pad256((size_t)rt * kBM * 2 * s * kTopK * 4) + pad256((size_t)rt * kBM * 4) + 512 +
         pad256((size_t)rt * kBM * kFbChunks * 8) + pad256((size_t)rt * kBM * kFbChunks * 4) + pad256(((size_t)rt * kBM / kFbRows + 1) * 4);   // fallback chunks
}

b2r_status launch_centre(b2r_solver* s, const double* f_ref, int n_ref, int dim, double* centre) {
  nn_centre_kernel<<<1, 1024, 0, s->stream>>>(f_ref, n_ref, dim, centre);
  ++s->launches;
  return B2R_OK;
}

b2r_status launch_prep(b2r_solver* s, const double* f, int dim, const double* centre, Prepped& p) {
  const long long total = (long long)p.n_pad * kKTot;
  nn_prep_kernel<<<(int)((total + 255) / 256), 256, 0, s->stream>>>(f, p.n, dim, p.n_pad, centre, p.l, p.r, p.norm, p.norm_lo);
  ++s->launches;
  return B2R_OK;
}

// nn[i] = exact argmin_j |fq_i - fr_j|^2 via the tensor-core shortlist.  scratch: cand_bytes(nq, nr) bytes.
b2r_status launch_direction(b2r_solver* s, const double* fq, const Prepped& q, const double* fr, const Prepped& r, int dim, char* scratch,
                            int* nn, double* dist, unsigned int* n_fallback, unsigned int* err_max = nullptr) {
  static PerDeviceOnce attr;
  const size_t smem = 3 * (size_t)kOperandBytes + sizeof(GemmSmem) + 1024;
  CK(attr.run([&] { return cudaFuncSetAttribute(nn_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); }));
  const int rt = q.n_pad / kBM, ct = r.n_pad / kBN;
  const int splits = choose_splits(rt, ct);
  float* cand_v = (float*)scratch;
  int* cand_j = (int*)(scratch + pad256((size_t)rt * kBM * 2 * splits * kTopK * 4));
  int* fb_list = (int*)(scratch + 2 * pad256((size_t)rt * kBM * 2 * splits * kTopK * 4));
  unsigned int* fb_count = (unsigned int*)((char*)fb_list + pad256((size_t)rt * kBM * 4));
  CK(cudaMemsetAsync(fb_count, 0, 4, s->stream));
  CUtensorMap mq, mr;
  if (!make_operand_map(&mq, q.l, q.n_pad) || !make_operand_map(&mr, r.r, r.n_pad))
    return fail(s, B2R_ERR_CUDA, "cuTensorMapEncodeTiled failed (driver entry point missing or bad operand buffer)");
  nn_gemm_kernel<<<dim3(rt, splits), kGemmThreads, smem, s->stream>>>(mq, mr, r.norm_lo, r.n, ct, splits, cand_v, cand_j);
  nn_finish_kernel<<<(q.n + kFinWarps - 1) / kFinWarps, kFinWarps * 32, 0, s->stream>>>(fq, q.n, fr, r.n, dim, q.norm, r.norm, r.norm_lo,
                                                                                        cand_v, cand_j, 2 * splits, nn, dist, fb_count, fb_list, err_max);
  int fb_ctas = (q.n + kFbRows - 1) / kFbRows;
  if (fb_ctas > 148 * 4) fb_ctas = 148 * 4;
  const int fb_tiles = (r.n + kFbTile - 1) / kFbTile;
  const int fb_chunks = fb_tiles >= 8 ? kFbChunks : 1;
  // partial results of the chunks and the arrival counters of the row groups, behind the fallback list
  char* fbp = (char*)fb_count + 256;
  double* part_d = (double*)fbp; fbp += pad256((size_t)rt * kBM * kFbChunks * 8);
  int* part_j = (int*)fbp; fbp += pad256((size_t)rt * kBM * kFbChunks * 4);
  unsigned int* part_cnt = (unsigned int*)fbp;
  if (fb_chunks > 1) CK(cudaMemsetAsync(part_cnt, 0, ((size_t)rt * kBM / kFbRows + 1) * 4, s->stream));
  nn_fallback_kernel<<<dim3(fb_ctas, fb_chunks), kFbRows * 32, 0, s->stream>>>(fq, fr, r.n, dim, fb_list, fb_count, nn, dist, n_fallback, part_d,
                                                                              part_j, part_cnt);
  s->launches += 3;
  CK(cudaGetLastError());
  return B2R_OK;
}

}  // namespace

extern "C" {

b2r_status b2r_nn_gemm(b2r_solver* s, const double* fq_dev, int nq, const double* fr_dev, int nr, int dim, int32_t* nn_dev,
                       double* dist_dev, int* n_fallback_out) {
  B2R_NVTX("b2r:nn_gemm");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!fq_dev || !fr_dev || !nn_dev || nq < 0 || nr < 0 || dim < 1) return fail(s, B2R_ERR_INVALID_ARG, "nn_gemm arguments");
  if (n_fallback_out) *n_fallback_out = 0;
  if (nq == 0) return B2R_OK;
  if (dim > kKSeg || nr == 0) return b2r_nn_exact(s, fq_dev, nq, fr_dev, nr, dim, nn_dev, dist_dev);  // K = 3 x 40 TF32 columns at most
  CK(cudaSetDevice(s->device));
  CK(s->d_nn.ensure(prepped_bytes(nq) + prepped_bytes(nr) + cand_bytes(nq, nr) + 2048));
  char* p = (char*)s->d_nn.p;
  Prepped q, r;
  p = carve_prepped(p, nq, &q);
  p = carve_prepped(p, nr, &r);
  unsigned int* n_fb = (unsigned int*)p; p += 256;
  double* centre = (double*)p; p += 512;
  CK(cudaMemsetAsync(n_fb, 0, 4, s->stream));
  b2r_status st = launch_centre(s, fr_dev, nr, dim, centre);
  if (st != B2R_OK) return st;
  st = launch_prep(s, fq_dev, dim, centre, q);
  if (st != B2R_OK) return st;
  st = launch_prep(s, fr_dev, dim, centre, r);
  if (st != B2R_OK) return st;
  st = launch_direction(s, fq_dev, q, fr_dev, r, dim, p, nn_dev, dist_dev, n_fb);
  if (st != B2R_OK) return st;
  if (n_fallback_out) {
    unsigned int h = 0;
    { const b2r_status rr = b2r_internal_read_words(s, n_fb, 1, &h); if (rr != B2R_OK) return rr; }
    *n_fallback_out = (int)h;
  }
  return B2R_OK;
}

b2r_status b2r_debug_nn_gemm_error(b2r_solver* s, const double* fq_dev, int nq, const double* fr_dev, int nr, int dim, int32_t* nn_dev,
                                   double* out3) {
  B2R_NVTX("b2r:debug_nn_gemm_error");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!fq_dev || !fr_dev || !nn_dev || !out3 || nq < 1 || nr < 1 || dim < 1 || dim > kKSeg) return fail(s, B2R_ERR_INVALID_ARG, "debug_nn_gemm_error arguments");
  CK(cudaSetDevice(s->device));
  CK(s->d_nn.ensure(prepped_bytes(nq) + prepped_bytes(nr) + cand_bytes(nq, nr) + 2048));
  char* p = (char*)s->d_nn.p;
  Prepped q, r;
  p = carve_prepped(p, nq, &q);
  p = carve_prepped(p, nr, &r);
  unsigned int* n_fb = (unsigned int*)p; p += 256;     // [0] fallback rows, [1] max error ratio (float bits)
  double* centre = (double*)p; p += 512;
  CK(cudaMemsetAsync(n_fb, 0, 8, s->stream));
  b2r_status st = launch_centre(s, fr_dev, nr, dim, centre);
  if (st == B2R_OK) st = launch_prep(s, fq_dev, dim, centre, q);
  if (st == B2R_OK) st = launch_prep(s, fr_dev, dim, centre, r);
  if (st == B2R_OK) st = launch_direction(s, fq_dev, q, fr_dev, r, dim, p, nn_dev, nullptr, n_fb, n_fb + 1);
  if (st != B2R_OK) return st;
  unsigned int h[2] = {0, 0};
  CK(cudaMemcpyAsync(h, n_fb, 8, cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  float ratio;
  memcpy(&ratio, &h[1], 4);
  out3[0] = ratio; out3[1] = kEpsRel; out3[2] = h[0];
  return B2R_OK;
}

b2r_status b2r_mutual_nn(b2r_solver* s, const double* fa_dev, int na, const double* fb_dev, int nb, int dim, int32_t* nn01_dev,
                         int32_t* nn10_dev, int32_t* pairs_dev, int* n_pairs_out, int* n_fallback_out) {
  B2R_NVTX("b2r:mutual_nn");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!fa_dev || !fb_dev || !nn01_dev || !nn10_dev || !pairs_dev || !n_pairs_out || na < 0 || nb < 0 || dim < 1)
    return fail(s, B2R_ERR_INVALID_ARG, "mutual_nn arguments");
  *n_pairs_out = 0;
  if (n_fallback_out) *n_fallback_out = 0;
  if (na == 0 || nb == 0) return B2R_OK;
  CK(cudaSetDevice(s->device));
  if (dim > kKSeg) {
    b2r_status st = b2r_nn_exact(s, fa_dev, na, fb_dev, nb, dim, nn01_dev, nullptr);
    if (st != B2R_OK) return st;
    st = b2r_nn_exact(s, fb_dev, nb, fa_dev, na, dim, nn10_dev, nullptr);
    if (st != B2R_OK) return st;
    return b2r_mutual_pairs(s, nn01_dev, na, nn10_dev, nb, pairs_dev, n_pairs_out);
  }
  const size_t cb = cand_bytes(na, nb) > cand_bytes(nb, na) ? cand_bytes(na, nb) : cand_bytes(nb, na);
  CK(s->d_nn.ensure(prepped_bytes(na) + prepped_bytes(nb) + cb + 2048));
  char* p = (char*)s->d_nn.p;
  Prepped a, b;
  p = carve_prepped(p, na, &a);
  p = carve_prepped(p, nb, &b);
  unsigned int* n_fb = (unsigned int*)p; p += 256;
  double* centre = (double*)p; p += 512;
  CK(cudaMemsetAsync(n_fb, 0, 4, s->stream));
  b2r_status st = launch_centre(s, fb_dev, nb, dim, centre);   // one centre for both directions
  if (st != B2R_OK) return st;
  st = launch_prep(s, fa_dev, dim, centre, a);
  if (st != B2R_OK) return st;
  st = launch_prep(s, fb_dev, dim, centre, b);
  if (st != B2R_OK) return st;
  st = launch_direction(s, fa_dev, a, fb_dev, b, dim, p, nn01_dev, nullptr, n_fb);   // same stream: scratch reuse is ordered
  if (st != B2R_OK) return st;
  st = launch_direction(s, fb_dev, b, fa_dev, a, dim, p, nn10_dev, nullptr, n_fb);
  if (st != B2R_OK) return st;
  st = b2r_mutual_pairs(s, nn01_dev, na, nn10_dev, nb, pairs_dev, n_pairs_out);  // synchronises
  if (st != B2R_OK) return st;
  if (n_fallback_out) {
    unsigned int h = 0;
    { const b2r_status rr = b2r_internal_read_words(s, n_fb, 1, &h); if (rr != B2R_OK) return rr; }
    *n_fallback_out = (int)h;
  }
  return B2R_OK;
}

}  // extern "C"
