// Scale estimation by adaptive voting over all TIM length ratios (SURVEY.md 8a-6, Appendix A.4-5 + A.7).
//
// Replaces upstream's TLSScaleSolver::solveForScale -> ScalarTLSEstimator::estimate, reached from `solver.solve`
// when Params.estimate_scaling is True (upstream default; the reference sets False at registration_utils.py:254 and
// HelpersRealtime.py:160, the north-star lists the scale vote).  For every pair k=(i<j), in upstream's TIM order:
//     a = ||src_j - src_i||, b = ||dst_j - dst_i||, X_k = b / a, r_k = beta * (1 / a),  beta = 2*noise_bound*sqrt(cbar2)
// 2P interval endpoints are sorted by (value, entry before exit, index); a sweep keeps the consensus set's running
// sums and the estimate is the weighted mean at the first minimum of the TLS cost.
//
// HBM-bound design: the endpoints are laid out entries-then-exits in index order, so ONE stable radix sort on the
// 64-bit value (radix.cu) realises the full three-level order; the sweep is a two-level scan (per-tile partial sums,
// scan of the tile sums, per-tile replay with the cost evaluated at every endpoint) and a (cost, position) argmin.
// The running sums are associated hierarchically instead of strictly left to right, so the estimate agrees with a
// sequential evaluation to a few ulp of the accumulated magnitude, not bit for bit (stated in DESIGN.md).
#include <math_constants.h>

#include "common.cuh"

namespace b2r {

namespace {

constexpr int kSwThreads = 256;
constexpr int kSwItems = 8;
constexpr int kSwTile = kSwThreads * kSwItems;  // endpoints per CTA
constexpr int kScaleMaxChunk = 4096;             // problems per batched launch (their endpoint counts live in problem 0's workspace)
constexpr int kNQ = 7;                           // count, w, w*x, r(entered), x, x*x, r(all entries)

__device__ __forceinline__ uint64_t ordered_key(double v) {
  const uint64_t b = (uint64_t)__double_as_longlong(v);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}

__device__ __forceinline__ double sq3_ref(double dx, double dy, double dz) {
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

// Workspace of ONE problem, laid out for the batch's largest N (n_max) so that every problem of a batch has the same offsets; the
// problems of a batch sit ws_stride bytes apart.
struct ScaleLayout {
  size_t X, R, keys, keys_alt, pay, pay_alt, table, tile_sums, best, e_count, total;   // byte offsets
  size_t seg_items;      // key / payload items between two problems (ws_stride / item size must match: see launch)
  int n_tiles_max;
};
__host__ __device__ inline size_t sl_pad(size_t b) { return (b + 255) / 256 * 256; }
__host__ inline ScaleLayout scale_layout(int n_max) {
  const size_t P = (size_t)n_max * (n_max - 1) / 2, E = 2 * P;
  ScaleLayout L;
  L.n_tiles_max = (int)((E + kSwTile - 1) / kSwTile);
  size_t o = 0;
  L.X = o; o += sl_pad(P * 8);
  L.R = o; o += sl_pad(P * 8);
  L.keys = o; o += sl_pad(E * 8);
  L.keys_alt = o; o += sl_pad(E * 8);
  L.pay = o; o += sl_pad(E * 4);
  L.pay_alt = o; o += sl_pad(E * 4);
  L.table = o; o += sl_pad(radix_table_words(E) * 4);
  L.tile_sums = o; o += sl_pad(((size_t)L.n_tiles_max + 1) * kNQ * 8);
  L.best = o; o += sl_pad((size_t)L.n_tiles_max * 3 * 8);
  L.e_count = o; o += 8 * kScaleMaxChunk;
  L.total = o;
  L.seg_items = 0;
  return L;
}

// grid = (ceil(n_max/256), n_max - 1, batch): thread (i = blockIdx.y, j) for j > i of problem blockIdx.z.
__global__ void __launch_bounds__(256) scale_tims_kernel(const ProbDesc* __restrict__ descs, double beta, char* __restrict__ ws, size_t ws_stride,
                                                         ScaleLayout L) {
  const ProbDesc d = descs[blockIdx.z];
  const double* __restrict__ src = d.src;
  const double* __restrict__ dst = d.dst;
  const int n = d.n;
  char* base = ws + (size_t)blockIdx.z * ws_stride;
  double* __restrict__ X = reinterpret_cast<double*>(base + L.X);
  double* __restrict__ R = reinterpret_cast<double*>(base + L.R);
  uint64_t* __restrict__ keys = reinterpret_cast<uint64_t*>(base + L.keys);
  uint32_t* __restrict__ pay = reinterpret_cast<uint32_t*>(base + L.pay);
  const long long P = (long long)n * (n - 1) / 2;
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0)
    reinterpret_cast<unsigned long long*>(ws + L.e_count)[blockIdx.z] = (unsigned long long)(2 * P);   // endpoints of every problem, contiguous at the head
  const int i = blockIdx.y;
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j <= i || j >= n) return;
  const long long k = (long long)i * n - (long long)i * (i + 1) / 2 + (j - i - 1);
  const double a2 = sq3_ref(__dsub_rn(src[3 * (size_t)j], src[3 * (size_t)i]), __dsub_rn(src[3 * (size_t)j + 1], src[3 * (size_t)i + 1]),
                            __dsub_rn(src[3 * (size_t)j + 2], src[3 * (size_t)i + 2]));
  const double b2 = sq3_ref(__dsub_rn(dst[3 * (size_t)j], dst[3 * (size_t)i]), __dsub_rn(dst[3 * (size_t)j + 1], dst[3 * (size_t)i + 1]),
                            __dsub_rn(dst[3 * (size_t)j + 2], dst[3 * (size_t)i + 2]));
  const double a = __dsqrt_rn(a2), b = __dsqrt_rn(b2);
  const double x = __ddiv_rn(b, a);
  const double r = __dmul_rn(beta, __ddiv_rn(1.0, a));
  X[k] = x; R[k] = r;
  keys[k] = ordered_key(__dsub_rn(x, r)); pay[k] = (uint32_t)k;
  keys[P + k] = ordered_key(__dadd_rn(x, r)); pay[P + k] = (uint32_t)(P + k);
}

struct Q7 { double q[kNQ]; };

__device__ __forceinline__ void endpoint_terms(uint32_t p, long long P, const double* __restrict__ X, const double* __restrict__ R,
                                               double (&t)[kNQ]) {
  const bool entry = (long long)p < P;
  const long long idx = entry ? (long long)p : (long long)p - P;
  const double eps = entry ? 1.0 : -1.0;
  const double x = X[idx], r = R[idx];
  const double w = 1.0 / (r * r);
  t[0] = eps; t[1] = eps * w; t[2] = eps * w * x; t[3] = eps * r; t[4] = eps * x; t[5] = eps * x * x;
  t[6] = entry ? r : 0.0;
}

// block-wide exclusive scan of kNQ doubles per thread (thread order); total of the block in tot[] (all threads)
__device__ void block_exclusive_scan(double (&v)[kNQ], double (&tot)[kNQ], double* sm /* kNQ * 8 */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double inc[kNQ];
#pragma unroll
  for (int q = 0; q < kNQ; ++q) {
    double x = v[q];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const double y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    inc[q] = x;
  }
  __syncthreads();
  if (lane == 31)
#pragma unroll
    for (int q = 0; q < kNQ; ++q) sm[q * 8 + warp] = inc[q];
  __syncthreads();
#pragma unroll
  for (int q = 0; q < kNQ; ++q) {
    double wb = 0, t = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) { const double s = sm[q * 8 + w]; if (w < warp) wb += s; t += s; }
    v[q] = wb + inc[q] - v[q];
    tot[q] = t;
  }
}

// pass 1 (kFinal=false): per-tile totals.  pass 2 (kFinal=true): replay with the scanned tile offsets, evaluate the
// cost at every endpoint, leave the tile's best (cost, position, estimate).
template <bool kFinal>
__global__ void __launch_bounds__(kSwThreads) tls_sweep_kernel(const ProbDesc* __restrict__ descs, char* __restrict__ ws, size_t ws_stride,
                                                               ScaleLayout L, int sorted_in_alt) {
  const int n = descs[blockIdx.y].n;
  const long long P = (long long)n * (n - 1) / 2, E = 2 * P;
  const int n_tiles = (int)((E + kSwTile - 1) / kSwTile);
  if ((int)blockIdx.x >= n_tiles) return;
  char* base = ws + (size_t)blockIdx.y * ws_stride;
  const uint32_t* __restrict__ pay = reinterpret_cast<const uint32_t*>(base + (sorted_in_alt ? L.pay_alt : L.pay));
  const double* __restrict__ X = reinterpret_cast<const double*>(base + L.X);
  const double* __restrict__ R = reinterpret_cast<const double*>(base + L.R);
  double* __restrict__ tile_sums = reinterpret_cast<double*>(base + L.tile_sums);   // n_tiles x kNQ (+ totals)
  double* __restrict__ best = reinterpret_cast<double*>(base + L.best);             // n_tiles x 3
  __shared__ double sm[kNQ * 8];
  __shared__ double sbest[3 * 8];
  const long long e0 = (long long)blockIdx.x * kSwTile + (long long)threadIdx.x * kSwItems;
  double t[kSwItems][kNQ];
  double part[kNQ];
#pragma unroll
  for (int q = 0; q < kNQ; ++q) part[q] = 0.0;
#pragma unroll
  for (int k = 0; k < kSwItems; ++k) {
    const long long e = e0 + k;
    if (e < E) endpoint_terms(pay[e], P, X, R, t[k]);
    else
#pragma unroll
      for (int q = 0; q < kNQ; ++q) t[k][q] = 0.0;
#pragma unroll
    for (int q = 0; q < kNQ; ++q) part[q] += t[k][q];
  }
  double tot[kNQ];
  block_exclusive_scan(part, tot, sm);
  if (!kFinal) {
    if (threadIdx.x == 0)
#pragma unroll
      for (int q = 0; q < kNQ; ++q) tile_sums[(size_t)blockIdx.x * kNQ + q] = tot[q];
    return;
  }
  // tile_sums now holds exclusive prefixes; slot [n_tiles] holds the grand totals
  double run[kNQ];
#pragma unroll
  for (int q = 0; q < kNQ; ++q) run[q] = tile_sums[(size_t)blockIdx.x * kNQ + q] + part[q];
  const double ranges_all = tile_sums[(size_t)n_tiles * kNQ + 6];
  double bc = CUDART_INF, bx = 0.0, be = 9.0e18;
#pragma unroll
  for (int k = 0; k < kSwItems; ++k) {
    const long long e = e0 + k;
    if (e < E) {
#pragma unroll
      for (int q = 0; q < 6; ++q) run[q] += t[k][q];
      const double xh = run[2] / run[1];
      const double cost = (run[0] * xh * xh + run[5] - 2 * run[4] * xh) + (ranges_all - run[3]);
      if (cost < bc) { bc = cost; bx = xh; be = (double)e; }
    }
  }
  // block argmin: lowest cost, ties -> lowest position (first minimum)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double oc = __shfl_xor_sync(0xffffffffu, bc, o), ox = __shfl_xor_sync(0xffffffffu, bx, o), oe = __shfl_xor_sync(0xffffffffu, be, o);
    if (oc < bc || (oc == bc && oe < be)) { bc = oc; bx = ox; be = oe; }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { sbest[warp] = bc; sbest[8 + warp] = bx; sbest[16 + warp] = be; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w)
      if (sbest[w] < bc || (sbest[w] == bc && sbest[16 + w] < be)) { bc = sbest[w]; bx = sbest[8 + w]; be = sbest[16 + w]; }
    best[(size_t)blockIdx.x * 3] = bc; best[(size_t)blockIdx.x * 3 + 1] = bx; best[(size_t)blockIdx.x * 3 + 2] = be;
  }
}

// single CTA: exclusive scan of the tile sums in place (tile order), grand totals to slot [n_tiles]
__global__ void __launch_bounds__(512) tls_scan_tiles_kernel(const ProbDesc* __restrict__ descs, char* __restrict__ ws, size_t ws_stride, ScaleLayout L) {
  const int n = descs[blockIdx.x].n;
  const long long E = (long long)n * (n - 1);
  const int n_tiles = (int)((E + kSwTile - 1) / kSwTile);
  double* __restrict__ tile_sums = reinterpret_cast<double*>(ws + (size_t)blockIdx.x * ws_stride + L.tile_sums);
  __shared__ double part[kNQ][512];
  const int tid = threadIdx.x;
  const int chunk = (n_tiles + 511) / 512;
  const int b = min(tid * chunk, n_tiles), e = min(b + chunk, n_tiles);
  double s[kNQ];
#pragma unroll
  for (int q = 0; q < kNQ; ++q) s[q] = 0.0;
  for (int i = b; i < e; ++i)
#pragma unroll
    for (int q = 0; q < kNQ; ++q) s[q] += tile_sums[(size_t)i * kNQ + q];
#pragma unroll
  for (int q = 0; q < kNQ; ++q) part[q][tid] = s[q];
  __syncthreads();
  if (tid < kNQ) {
    double run = 0;
    for (int i = 0; i < 512; ++i) { const double t = part[tid][i]; part[tid][i] = run; run += t; }
    tile_sums[(size_t)n_tiles * kNQ + tid] = run;
  }
  __syncthreads();
#pragma unroll
  for (int q = 0; q < kNQ; ++q) s[q] = part[q][tid];
  for (int i = b; i < e; ++i)
#pragma unroll
    for (int q = 0; q < kNQ; ++q) { const double t = tile_sums[(size_t)i * kNQ + q]; tile_sums[(size_t)i * kNQ + q] = s[q]; s[q] += t; }
}

// single CTA: argmin over the tiles -> ProbState::scale
__global__ void __launch_bounds__(256) tls_pick_kernel(const ProbDesc* __restrict__ descs, const char* __restrict__ ws, size_t ws_stride, ScaleLayout L) {
  const ProbDesc d = descs[blockIdx.x];
  ProbState* __restrict__ st = d.st;
  const long long E = (long long)d.n * (d.n - 1);
  const int n_tiles = (int)((E + kSwTile - 1) / kSwTile);
  const double* __restrict__ best = reinterpret_cast<const double*>(ws + (size_t)blockIdx.x * ws_stride + L.best);
  __shared__ double sb[3 * 256];
  double bc = CUDART_INF, bx = 0.0, be = 9.0e18;
  for (int i = threadIdx.x; i < n_tiles; i += 256) {
    const double c = best[(size_t)i * 3], x = best[(size_t)i * 3 + 1], e = best[(size_t)i * 3 + 2];
    if (c < bc || (c == bc && e < be)) { bc = c; bx = x; be = e; }
  }
  sb[threadIdx.x] = bc; sb[256 + threadIdx.x] = bx; sb[512 + threadIdx.x] = be;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) {
      const double c = sb[threadIdx.x + s], e = sb[512 + threadIdx.x + s];
      if (c < sb[threadIdx.x] || (c == sb[threadIdx.x] && e < sb[512 + threadIdx.x])) {
        sb[threadIdx.x] = c; sb[256 + threadIdx.x] = sb[256 + threadIdx.x + s]; sb[512 + threadIdx.x] = e;
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) st->scale = (sb[512] < 8.0e18) ? sb[256] : CUDART_NAN;
}

}  // namespace

size_t scale_workspace_bytes(int n) { return scale_layout(n).total; }
int scale_max_chunk() { return kScaleMaxChunk; }

// A batch of problems (device descriptors dd[0 .. batch)): every problem's scale into its ProbState::scale, in ONE set of launches
// (grid dimension = problem) instead of ~40 launches per problem.  ws: batch * scale_workspace_bytes(n_max) bytes, 256-byte aligned.
int launch_scale_estimate_batch(const ProbDesc* dd, int batch, int n_max, double beta, void* ws, cudaStream_t stream) {
  if (n_max < 2 || batch <= 0) return 0;
  const ScaleLayout L = scale_layout(n_max);
  const size_t stride = L.total;
  const size_t E = (size_t)n_max * (n_max - 1);
  char* base = (char*)ws;
  int launches = 0;
  dim3 grid((n_max + 255) / 256, n_max - 1, batch);
  scale_tims_kernel<<<grid, 256, 0, stream>>>(dd, beta, base, stride, L);
  ++launches;
  bool in_alt = false;
  // keys (8 B) and payloads (4 B) of consecutive problems are `stride` BYTES apart: stride / 8 keys, stride / 4 payloads and table words
  launches += radix_sort_pairs_batched(reinterpret_cast<uint64_t*>(base + L.keys), reinterpret_cast<uint32_t*>(base + L.pay),
                                        reinterpret_cast<uint64_t*>(base + L.keys_alt), reinterpret_cast<uint32_t*>(base + L.pay_alt),
                                        reinterpret_cast<uint32_t*>(base + L.table), E, reinterpret_cast<const unsigned long long*>(base + L.e_count),
                                        stride / 8, stride / 4, stride / 4, batch, 0, 64, stream, &in_alt);
  dim3 sgrid(L.n_tiles_max, batch);
  tls_sweep_kernel<false><<<sgrid, kSwThreads, 0, stream>>>(dd, base, stride, L, in_alt ? 1 : 0);
  tls_scan_tiles_kernel<<<batch, 512, 0, stream>>>(dd, base, stride, L);
  tls_sweep_kernel<true><<<sgrid, kSwThreads, 0, stream>>>(dd, base, stride, L, in_alt ? 1 : 0);
  tls_pick_kernel<<<batch, 256, 0, stream>>>(dd, base, stride, L);
  return launches + 4;
}

}  // namespace b2r
