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
// trusted only when |F| > tau * (A + B + beta^2)^2 (tau = 1e-9, seven orders above the rounding error of F and
// of the reference's own sqrt/sub); the rest (about 1e-8 of the pairs) take the exact two-sqrt path.
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

template <bool kExactOnly>
__device__ __forceinline__ bool verdict(double A, double B, const KernelParams& kp, unsigned& slow) {
  if (kExactOnly) return verdict_exact(A, B, kp.beta);
  const double S1 = A + B;
  const double H = fma(0.5, S1, -kp.half_beta2);       // D / 2
  const double AB = A * B;
  const double F = fma(H, H, -AB);                     // (D^2 - 4AB) / 4
  const double Sg = fma(kp.stau, S1, kp.stau_beta2);   // sqrt(tau)/2 * (A + B + beta^2)
  bool in = (H <= 0.0) | (F <= 0.0);
  if (!(fabs(F) > Sg * Sg)) {                          // ambiguous (also catches NaN): exact path
    in = verdict_exact(A, B, kp.beta);
    ++slow;
  }
  return in;
}

struct TilePts {
  double2 a[kTile];  // (sx, sy)
  double2 b[kTile];  // (sz, dx)
  double2 c[kTile];  // (dy, dz)
};

// Computes one 128x128 tile: rows i0.., columns j0...  Thread t owns column j0+t.
// dir[t] (shared) receives the 4 words of row i0+t for the column block; tw[0..3] (registers) the 4 words of
// row j0+t for the row block (transposed tile).
template <bool kExactOnly>
__device__ __forceinline__ void compute_tile(const double* __restrict__ src, const double* __restrict__ dst, int n,
                                             int i0, int j0, const KernelParams& kp, TilePts& sp,
                                             uint32_t (*dir)[4], uint32_t tw[4], unsigned& slow) {
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  {  // stage the row points (coalesced 24-byte AoS reads), zero-fill past n
    const int i = i0 + t;
    double v[6] = {0, 0, 0, 0, 0, 0};
    if (i < n) {
      v[0] = src[3 * (size_t)i]; v[1] = src[3 * (size_t)i + 1]; v[2] = src[3 * (size_t)i + 2];
      v[3] = dst[3 * (size_t)i]; v[4] = dst[3 * (size_t)i + 1]; v[5] = dst[3 * (size_t)i + 2];
    }
    sp.a[t] = make_double2(v[0], v[1]);
    sp.b[t] = make_double2(v[2], v[3]);
    sp.c[t] = make_double2(v[4], v[5]);
  }
  const int j = j0 + t;
  double xj = 0, yj = 0, zj = 0, uj = 0, vj = 0, wj = 0;
  const bool jok = j < n;
  if (jok) {
    xj = src[3 * (size_t)j]; yj = src[3 * (size_t)j + 1]; zj = src[3 * (size_t)j + 2];
    uj = dst[3 * (size_t)j]; vj = dst[3 * (size_t)j + 1]; wj = dst[3 * (size_t)j + 2];
  }
  __syncthreads();
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint32_t tword = 0, dword = 0;
#pragma unroll 8
    for (int k = 0; k < 32; ++k) {
      const int il = 32 * c + k;
      const double2 pa = sp.a[il], pb = sp.b[il], pc = sp.c[il];
      const double A = sq3_ref(__dsub_rn(xj, pa.x), __dsub_rn(yj, pa.y), __dsub_rn(zj, pb.x));
      const double B = sq3_ref(__dsub_rn(uj, pb.y), __dsub_rn(vj, pc.x), __dsub_rn(wj, pc.y));
      const int i = i0 + il;
      bool in = verdict<kExactOnly>(A, B, kp, slow);
      in = in && jok && (i < n) && (i != j);
      const uint32_t bal = __ballot_sync(0xffffffffu, in);
      if (lane == k) dword = bal;
      tword |= (in ? 1u : 0u) << k;
    }
    dir[32 * c + lane][warp] = dword;
    tw[c] = tword;
  }
  __syncthreads();
}

// Batched, symmetric: grid = (T(T+1)/2 for the largest problem, batch).
template <bool kExactOnly>
__global__ void __launch_bounds__(kTile) tim_graph_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
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
  uint32_t tw[4];
  unsigned slow = 0;
  compute_tile<kExactOnly>(d.src, d.dst, n, ti * kTile, tj * kTile, kp, sp, dir, tw, slow);
  const int t = threadIdx.x;
  const int i = ti * kTile + t;
  if (i < n) *reinterpret_cast<uint4*>(d.adj + (size_t)i * d.pitch + 4 * tj) = *reinterpret_cast<const uint4*>(dir[t]);
  if (ti != tj) {
    const int j = tj * kTile + t;
    if (j < n) *reinterpret_cast<uint4*>(d.adj + (size_t)j * d.pitch + 4 * ti) = make_uint4(tw[0], tw[1], tw[2], tw[3]);
  }
  if (!kExactOnly && slow) atomicAdd(&d.st->slow_pairs, (unsigned long long)slow);
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
  uint32_t tw[4];
  unsigned slow = 0;
  // rows beyond row_end are computed but not stored (n_eff trick keeps the ballot uniform)
  compute_tile<kExactOnly>(src, dst, n, i0, tj * kTile, kp, sp, dir, tw, slow);
  const int i = i0 + threadIdx.x;
  if (i < row_end && i < n)
    *reinterpret_cast<uint4*>(adj + (size_t)(i - row_begin) * pitch + 4 * tj) = *reinterpret_cast<const uint4*>(dir[threadIdx.x]);
  if (!kExactOnly && slow && slow_pairs) atomicAdd(slow_pairs, (unsigned long long)slow);
}

// Degrees: one warp per row, popcount over the row's words; also max degree and 2*edges per problem.
__global__ void __launch_bounds__(256) degree_kernel(const ProbDesc* __restrict__ descs) {
  const ProbDesc d = descs[blockIdx.y];
  const int lane = threadIdx.x & 31;
  const int v = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (v >= d.n) return;
  const int W = 4 * ((d.n + kTile - 1) / kTile);
  const uint32_t* row = d.adj + (size_t)v * d.pitch;
  int c = 0;
  for (int w = lane; w < W; w += 32) c += __popc(row[w]);
  c = __reduce_add_sync(0xffffffffu, c);
  if (lane == 0) {
    d.deg[v] = c;
    atomicMax(&d.st->max_deg, c);
    atomicAdd(&d.st->deg_sum, (unsigned long long)c);
  }
}

int launch_tim_graph(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, cudaStream_t st) {
  const int T = (n_max + kTile - 1) / kTile;
  dim3 grid(T * (T + 1) / 2, batch);
  if (kp.exact_only) tim_graph_kernel<true><<<grid, kTile, 0, st>>>(d_desc, kp);
  else tim_graph_kernel<false><<<grid, kTile, 0, st>>>(d_desc, kp);
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
  dim3 grid((n_max + 7) / 8, batch);
  degree_kernel<<<grid, 256, 0, st>>>(d_desc);
  return 1;
}

}  // namespace b2r
