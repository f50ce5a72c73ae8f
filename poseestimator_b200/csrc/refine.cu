// Refinement and scoring right after TEASER++: what the reference runs per template
//   registration_icp(src, dst, 0.05, H, PointToPoint, max_iteration=1000)   PoseEstimator.py:97-101, main_realsense.py:112-115
//   alignment_score / nn_residuals / voxel_coverage                         HelpersRealtime.py:100-130
//   chamfer_distance                                                        registration_utils.py:276-280
//   cloud_resolution                                                        registration_utils.py:229-240
//   pca_axes / initial_align_centroid_pca                                   registration_utils.py:290-343
// (paths under /root/reference/src/pose_estimator/EstimHelpers/).
//
// ICP keeps the whole loop on the device: per iteration three launches (nearest neighbours of the moved source, a one-CTA
// statistics + Umeyama step in a fixed summation order, the in-place move), every kernel looks at a device-side `done` flag
// first, and the host only reads that flag once per chunk of iterations.  Distances are the exact fp64 expression of the
// oracle ((dx*dx + dy*dy) + dz*dz, no FMA: this file is compiled with -fmad=false), ties go to the lowest index.
#include <math_constants.h>

#include "linalg.cuh"
#include "solver.cuh"

namespace b2r {

b2r_status internal_knn(b2r_solver* s, const double* pts, int n, const double* queries, int nq, int K, double radius, int* idx,
                        double* d2, int* cnt);  // presteps.cu

namespace {

#define FULL 0xffffffffu
constexpr int kNnWarps = 8;
constexpr int kNnTile = 1024;

struct IcpState {
  double T[16];        // accumulated transformation (row-major 4x4)
  double U[16];        // update of the current iteration
  double fitness, rmse, prev_fitness, prev_rmse;
  int n_corr, iterations, done, converged;
};

__device__ __forceinline__ double sq3r(double dx, double dy, double dz) {
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

// p' = T * (x, y, z, 1): ((T0 x + T1 y) + T2 z) + T3, each operation rounded (Open3D PointCloud::Transform)
__device__ __forceinline__ void xform(const double* T, double x, double y, double z, double& ox, double& oy, double& oz) {
  ox = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[0], x), __dmul_rn(T[1], y)), __dmul_rn(T[2], z)), T[3]);
  oy = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[4], x), __dmul_rn(T[5], y)), __dmul_rn(T[6], z)), T[7]);
  oz = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(T[8], x), __dmul_rn(T[9], y)), __dmul_rn(T[10], z)), T[11]);
}

struct Mat12 { double m[12]; };

__global__ void __launch_bounds__(256) transform_kernel(const double* __restrict__ in, int n, Mat12 T, double* __restrict__ out) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  double x, y, z;
  xform(T.m, in[3 * (size_t)i], in[3 * (size_t)i + 1], in[3 * (size_t)i + 2], x, y, z);
  out[3 * (size_t)i] = x; out[3 * (size_t)i + 1] = y; out[3 * (size_t)i + 2] = z;
}

// Exact nearest neighbour, one warp per query, the reference cloud streamed through shared memory in tiles shared by the CTA's
// 8 warps.  r2 > 0: index -1 when the nearest point is not strictly inside the radius (SearchHybrid(point, r, 1)).
// done != NULL && *done: nothing to do (ICP finished).
__global__ void __launch_bounds__(kNnWarps * 32) nn1_kernel(const double* __restrict__ q, int nq, const double* __restrict__ pts, int n, double r2,
                                                            int* __restrict__ idx_out, double* __restrict__ d2_out, const int* __restrict__ done) {
  if (done && *done) return;
  __shared__ double tile[3 * kNnTile];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = blockIdx.x * kNnWarps + warp;
  const bool active = i < nq;
  double qx = 0, qy = 0, qz = 0;
  if (active) { qx = q[3 * (size_t)i]; qy = q[3 * (size_t)i + 1]; qz = q[3 * (size_t)i + 2]; }
  double best = CUDART_INF; int bj = 0x7fffffff;
  for (int base = 0; base < n; base += kNnTile) {
    __syncthreads();
    const int m = min(kNnTile, n - base);
    for (int e = threadIdx.x; e < 3 * m; e += kNnWarps * 32) tile[e] = pts[3 * (size_t)base + e];
    __syncthreads();
    if (!active) continue;
    for (int jl = lane; jl < m; jl += 32) {
      const double d2 = sq3r(__dsub_rn(tile[3 * jl], qx), __dsub_rn(tile[3 * jl + 1], qy), __dsub_rn(tile[3 * jl + 2], qz));
      if (d2 < best) { best = d2; bj = base + jl; }   // ascending j per lane: the first minimum is kept
    }
  }
  if (!active) return;
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    const double ob = __shfl_xor_sync(FULL, best, o); const int oj = __shfl_xor_sync(FULL, bj, o);
    if (ob < best || (ob == best && oj < bj)) { best = ob; bj = oj; }
  }
  if (lane == 0) {
    const bool hit = bj != 0x7fffffff && (!(r2 > 0) || best < r2);
    if (idx_out) idx_out[i] = hit ? bj : -1;
    if (d2_out) d2_out[i] = best;
  }
}

// K = 1 output of the grid search (idx, d2 per query) -> the same (idx or -1, d2) convention as nn1_kernel
__global__ void __launch_bounds__(256) nn1_from_knn_kernel(int nq, double r2, int* __restrict__ idx, const double* __restrict__ d2, const int* __restrict__ cnt) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= nq) return;
  if (cnt[i] < 1 || (r2 > 0 && !(d2[i] < r2))) idx[i] = -1;
}

template <int K>
__device__ __forceinline__ void block_sum256(double (&v)[K], double* red /* K * 8 */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double x = v[k];
#pragma unroll
    for (int o = 16; o; o >>= 1) x += __shfl_xor_sync(FULL, x, o);
    if (lane == 0) red[k * 8 + warp] = x;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double t = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += red[k * 8 + w];
    v[k] = t;
  }
  __syncthreads();
}

// One CTA: statistics of the current correspondences, convergence test, Umeyama update (GetRegistrationResultAndCorrespondences
// + the loop body of RegistrationICP).  Fixed summation order (thread-strided partial sums, shuffle tree, 8 warp sums).
__global__ void __launch_bounds__(256) icp_step_kernel(const double* __restrict__ cur, int ns, const double* __restrict__ dst,
                                                       const int* __restrict__ corr, const double* __restrict__ d2, int max_iter,
                                                       double rel_fitness, double rel_rmse, IcpState* __restrict__ st) {
  if (st->done) return;
  __shared__ double red[10 * 8];
  __shared__ double s_mean[6];
  double v[7] = {0, 0, 0, 0, 0, 0, 0};  // count, err, sum src xyz ... (dst sums in a second array to keep registers low)
  double w[3] = {0, 0, 0};
  for (int i = threadIdx.x; i < ns; i += 256) {
    const int j = corr[i];
    if (j >= 0) {
      v[0] += 1.0; v[1] += d2[i];
      v[2] += cur[3 * (size_t)i]; v[3] += cur[3 * (size_t)i + 1]; v[4] += cur[3 * (size_t)i + 2];
      w[0] += dst[3 * (size_t)j]; w[1] += dst[3 * (size_t)j + 1]; w[2] += dst[3 * (size_t)j + 2];
    }
  }
  block_sum256<7>(v, red);
  block_sum256<3>(w, red);
  const double m = v[0];
  const double fitness = m > 0 ? m / (double)ns : 0.0;
  const double rmse = m > 0 ? sqrt(v[1] / m) : 0.0;
  bool stop = false;
  if (st->iterations > 0 && fabs(st->prev_fitness - fitness) < rel_fitness && fabs(st->prev_rmse - rmse) < rel_rmse) stop = true;
  const bool conv = stop;
  if (st->iterations >= max_iter) stop = true;
  if (threadIdx.x < 3) { s_mean[threadIdx.x] = m > 0 ? v[2 + threadIdx.x] / m : 0.0; s_mean[3 + threadIdx.x] = m > 0 ? w[threadIdx.x] / m : 0.0; }
  __syncthreads();
  if (stop) {
    if (threadIdx.x == 0) { st->fitness = fitness; st->rmse = rmse; st->n_corr = (int)m; st->done = 1; st->converged = conv ? 1 : 0; }
    return;
  }
  // H = sum (src - ms)(dst - md)^T  (transpose of Umeyama's sigma; the 1/m factor does not change the rotation)
  double h[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  const double ms0 = s_mean[0], ms1 = s_mean[1], ms2 = s_mean[2], md0 = s_mean[3], md1 = s_mean[4], md2 = s_mean[5];
  for (int i = threadIdx.x; i < ns; i += 256) {
    const int j = corr[i];
    if (j >= 0) {
      const double a0 = cur[3 * (size_t)i] - ms0, a1 = cur[3 * (size_t)i + 1] - ms1, a2 = cur[3 * (size_t)i + 2] - ms2;
      const double b0 = dst[3 * (size_t)j] - md0, b1 = dst[3 * (size_t)j + 1] - md1, b2 = dst[3 * (size_t)j + 2] - md2;
      h[0] += a0 * b0; h[1] += a0 * b1; h[2] += a0 * b2;
      h[3] += a1 * b0; h[4] += a1 * b1; h[5] += a1 * b2;
      h[6] += a2 * b0; h[7] += a2 * b1; h[8] += a2 * b2;
    }
  }
  block_sum256<9>(h, red);
  if (threadIdx.x == 0) {
    double U[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    if (m > 0) {
      double R[9];
      rotation_from_H(h, R);   // R = V diag(1,1,det) U^T of H = U S V^T  ==  Umeyama's U_s S' V_s^T of sigma = H^T
      for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) U[4 * r + c] = R[3 * r + c];
        U[4 * r + 3] = s_mean[3 + r] - (R[3 * r] * ms0 + R[3 * r + 1] * ms1 + R[3 * r + 2] * ms2);
      }
    }
    double Tn[16];
    for (int r = 0; r < 4; ++r)
      for (int c = 0; c < 4; ++c) {
        double acc = 0;
        for (int k = 0; k < 4; ++k) acc += U[4 * r + k] * st->T[4 * k + c];
        Tn[4 * r + c] = acc;
      }
    for (int k = 0; k < 16; ++k) { st->T[k] = Tn[k]; st->U[k] = U[k]; }
    st->prev_fitness = fitness; st->prev_rmse = rmse;
    st->fitness = fitness; st->rmse = rmse; st->n_corr = (int)m;
    st->iterations += 1;
  }
}

__global__ void __launch_bounds__(256) icp_apply_kernel(double* __restrict__ cur, int ns, const IcpState* __restrict__ st) {
  if (st->done) return;
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= ns) return;
  double x, y, z;
  xform(st->U, cur[3 * (size_t)i], cur[3 * (size_t)i + 1], cur[3 * (size_t)i + 2], x, y, z);
  cur[3 * (size_t)i] = x; cur[3 * (size_t)i + 1] = y; cur[3 * (size_t)i + 2] = z;
}

// ------------------------------------------------------------------------------------------------------------------
// Order statistics without a sort: rank of every value = #{j : (v_j, j) < (v_i, i)}; the values whose rank is one of the
// (<= 4) requested positions are written out.  O(n^2) compares, one launch, deterministic; n <= kSelectMax.
// sqrt_in: values are squared distances, compare and output their square roots (monotone, so ranks are unchanged).
// ------------------------------------------------------------------------------------------------------------------
constexpr int kSelectMax = 1 << 17;
struct Ranks { int k[4]; };

__global__ void __launch_bounds__(256) select_kernel(const double* __restrict__ v, int n, Ranks want, int sqrt_out, double* __restrict__ out4) {
  __shared__ double tile[2048];
  const int i = blockIdx.x * 256 + threadIdx.x;
  const double mine = i < n ? v[i] : 0.0;
  int rank = 0;
  for (int base = 0; base < n; base += 2048) {
    __syncthreads();
    const int m = min(2048, n - base);
    for (int e = threadIdx.x; e < m; e += 256) tile[e] = v[base + e];
    __syncthreads();
    if (i < n) {
      for (int e = 0; e < m; ++e) {
        const double o = tile[e];
        rank += (o < mine) || (o == mine && base + e < i);
      }
    }
  }
  if (i < n) {
#pragma unroll
    for (int q = 0; q < 4; ++q)
      if (want.k[q] == rank) out4[q] = sqrt_out ? sqrt(mine) : mine;
  }
}

// voxel_coverage: number of distinct (floor(x/v), floor(y/v), floor(z/v)) int32 triples.  Open-addressing table of point
// ids: a slot is claimed with one CAS; a later point compares its key with the claimant's (recomputed from the input).
__device__ __forceinline__ void voxel_key(const double* __restrict__ p, int i, double voxel, int& x, int& y, int& z) {
  x = (int)floor(__ddiv_rn(p[3 * (size_t)i], voxel)); y = (int)floor(__ddiv_rn(p[3 * (size_t)i + 1], voxel)); z = (int)floor(__ddiv_rn(p[3 * (size_t)i + 2], voxel));
}
__global__ void __launch_bounds__(256) voxel_count_kernel(const double* __restrict__ pts, int n, double voxel, int* __restrict__ table,
                                                          unsigned mask, int* __restrict__ count) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  bool fresh = false;
  if (i < n) {
    int x, y, z;
    voxel_key(pts, i, voxel, x, y, z);
    unsigned h = (unsigned)x * 73856093u ^ (unsigned)y * 19349663u ^ (unsigned)z * 83492791u;
    h ^= h >> 15; h *= 0x2c1b3c6du; h ^= h >> 12;
    for (unsigned s = h & mask;; s = (s + 1) & mask) {
      const int old = atomicCAS(&table[s], 0, i + 1);
      if (old == 0) { fresh = true; break; }
      int ox, oy, oz;
      voxel_key(pts, old - 1, voxel, ox, oy, oz);
      if (ox == x && oy == y && oz == z) break;
    }
  }
  const unsigned bal = __ballot_sync(FULL, fresh);
  if ((threadIdx.x & 31) == 0 && bal) atomicAdd(count, __popc(bal));
}

// sum of sqrt(d2[i]) in a fixed order (one CTA) -> out[slot]
__global__ void __launch_bounds__(256) sum_sqrt_kernel(const double* __restrict__ d2, int n, double* __restrict__ out, int slot) {
  __shared__ double red[8];
  double v[1] = {0};
  for (int i = threadIdx.x; i < n; i += 256) v[0] += sqrt(d2[i]);
  block_sum256<1>(v, red);
  if (threadIdx.x == 0) out[slot] = v[0];
}

// flatten the k nearest OTHER points' squared distances of the sub-sampled queries (cloud_resolution: dd[1:])
__global__ void __launch_bounds__(256) drop_self_kernel(const double* __restrict__ d2, const int* __restrict__ cnt, int nq, int K, double* __restrict__ out,
                                                        int* __restrict__ total) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= nq) return;
  // every query has exactly min(K, n) neighbours (no radius), so the layout is dense: (K-1) values per query
  for (int t = 1; t < K; ++t) out[(size_t)i * (K - 1) + (t - 1)] = d2[(size_t)i * K + t];
  if (i == 0) *total = nq * (K - 1);
  (void)cnt;
}

__global__ void __launch_bounds__(256) stride_gather_kernel(const double* __restrict__ pts, int nq, int step, double* __restrict__ out) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= nq) return;
  for (int k = 0; k < 3; ++k) out[3 * (size_t)i + k] = pts[3 * (size_t)i * step + k];
}

// pca_axes of one cloud per CTA (grid = number of clouds): centroid, covariance / max(n-1, 1), Jacobi eigenvectors ordered
// by decreasing eigenvalue, det = +1.  out per cloud: R[9] (columns = axes), s[3], centroid[3] = 15 doubles.
struct PcaIn { const double* pts[2]; int n[2]; };
__global__ void __launch_bounds__(256) pca_kernel(PcaIn in, double* __restrict__ out) {
  __shared__ double red[9 * 8];
  __shared__ double s_c[3];
  const double* p = in.pts[blockIdx.x];
  const int n = in.n[blockIdx.x];
  double c[3] = {0, 0, 0};
  for (int i = threadIdx.x; i < n; i += 256) { c[0] += p[3 * (size_t)i]; c[1] += p[3 * (size_t)i + 1]; c[2] += p[3 * (size_t)i + 2]; }
  block_sum256<3>(c, red);
  if (threadIdx.x < 3) s_c[threadIdx.x] = c[threadIdx.x] / (double)n;
  __syncthreads();
  double m[6] = {0, 0, 0, 0, 0, 0};
  for (int i = threadIdx.x; i < n; i += 256) {
    const double x = p[3 * (size_t)i] - s_c[0], y = p[3 * (size_t)i + 1] - s_c[1], z = p[3 * (size_t)i + 2] - s_c[2];
    m[0] += x * x; m[1] += x * y; m[2] += x * z; m[3] += y * y; m[4] += y * z; m[5] += z * z;
  }
  block_sum256<6>(m, red);
  if (threadIdx.x == 0) {
    const double den = (double)max(n - 1, 1);
    const double C[9] = {m[0] / den, m[1] / den, m[2] / den, m[1] / den, m[3] / den, m[4] / den, m[2] / den, m[4] / den, m[5] / den};
    double w[3], Q[9];
    eig3_sym(C, w, Q);
    int o0 = 0, o1 = 1, o2 = 2;
    if (w[o0] < w[o1]) { int t = o0; o0 = o1; o1 = t; }
    if (w[o1] < w[o2]) { int t = o1; o1 = o2; o2 = t; }
    if (w[o0] < w[o1]) { int t = o0; o0 = o1; o1 = t; }
    const int ord[3] = {o0, o1, o2};
    double* o = out + 15 * blockIdx.x;
    for (int jj = 0; jj < 3; ++jj) {
      for (int i = 0; i < 3; ++i) o[3 * i + jj] = Q[3 * i + ord[jj]];
      o[9 + jj] = sqrt(fmax(w[ord[jj]], 0.0));
    }
    if (det3(o) < 0) { o[2] = -o[2]; o[5] = -o[5]; o[8] = -o[8]; }
    o[12] = s_c[0]; o[13] = s_c[1]; o[14] = s_c[2];
  }
}

// uniform_downsample_farthest_point (registration_utils.py:98-128) from a given first index: one persistent CTA, the running
// minimum distances stay in global memory (L2), every round is one fused pass (distance to the last pick, running minimum,
// arg-max with the lowest index on ties) + one block reduction.
__global__ void __launch_bounds__(1024) fps_kernel(const double* __restrict__ pts, int n, int target, int first, double* __restrict__ dist,
                                                   int* __restrict__ idx_out) {
  __shared__ double s_val[32];
  __shared__ int s_idx[32];
  __shared__ int s_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int j = tid; j < n; j += 1024) dist[j] = CUDART_INF;
  if (tid == 0) { idx_out[0] = first; s_last = first; }
  __syncthreads();
  for (int i = 1; i < target; ++i) {
    const int last = s_last;
    const double lx = pts[3 * (size_t)last], ly = pts[3 * (size_t)last + 1], lz = pts[3 * (size_t)last + 2];
    double bv = -1.0; int bj = 0x7fffffff;
    for (int j = tid; j < n; j += 1024) {
      const double d = __dsqrt_rn(sq3r(__dsub_rn(pts[3 * (size_t)j], lx), __dsub_rn(pts[3 * (size_t)j + 1], ly), __dsub_rn(pts[3 * (size_t)j + 2], lz)));
      double m = dist[j];
      if (d < m) { m = d; dist[j] = d; }
      if (m > bv) { bv = m; bj = j; }          // ascending j per thread: the first maximum is kept
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      const double ov = __shfl_xor_sync(FULL, bv, o); const int oj = __shfl_xor_sync(FULL, bj, o);
      if (ov > bv || (ov == bv && oj < bj)) { bv = ov; bj = oj; }
    }
    if (lane == 0) { s_val[warp] = bv; s_idx[warp] = bj; }
    __syncthreads();
    if (warp == 0) {
      bv = s_val[lane]; bj = s_idx[lane];
#pragma unroll
      for (int o = 16; o; o >>= 1) {
        const double ov = __shfl_xor_sync(FULL, bv, o); const int oj = __shfl_xor_sync(FULL, bj, o);
        if (ov > bv || (ov == bv && oj < bj)) { bv = ov; bj = oj; }
      }
      if (lane == 0) { idx_out[i] = bj; dist[bj] = 0.0; s_last = bj; }
    }
    __syncthreads();
  }
}

}  // namespace
}  // namespace b2r

using namespace b2r;

namespace {

constexpr int kNnBruteMax = 16384;  // above this many reference points the uniform-grid search of presteps.cu takes over

struct Ws {
  char* p;
  template <typename T> T* take(size_t count) { char* r = p; p += (count * sizeof(T) + 255) / 256 * 256; return reinterpret_cast<T*>(r); }
};

b2r_status refine_ws(b2r_solver* s, size_t bytes, Ws* w) {
  CK(cudaSetDevice(s->device));
  CK(s->d_refine.ensure(bytes + 8192));
  CK(s->h_refine.ensure(4096));
  w->p = (char*)s->d_refine.p;
  return B2R_OK;
}

// nearest reference point of every query: idx (or -1 outside r2 > 0) and squared distance
b2r_status run_nn1(b2r_solver* s, const double* q, int nq, const double* pts, int n, double r2, int* idx, double* d2, int* cnt_scratch,
                   const int* done) {
  if (nq <= 0) return B2R_OK;
  if (n <= kNnBruteMax || !idx || !d2 || !cnt_scratch) {
    nn1_kernel<<<(nq + kNnWarps - 1) / kNnWarps, kNnWarps * 32, 0, s->stream>>>(q, nq, pts, n, r2, idx, d2, done);
    s->launches += 1;
    return B2R_OK;
  }
  b2r_status r = internal_knn(s, pts, n, q, nq, 1, 0.0, idx, d2, cnt_scratch);
  if (r != B2R_OK) return r;
  nn1_from_knn_kernel<<<(nq + 255) / 256, 256, 0, s->stream>>>(nq, r2, idx, d2, cnt_scratch);
  s->launches += 1;
  return B2R_OK;
}

// numpy.percentile(method="linear") / numpy.median from the order statistics a[i0], a[i1] (numpy's two-sided _lerp)
double np_lerp(double A, double B, double t) { const double d = B - A; return t >= 0.5 ? B - d * (1 - t) : A + d * t; }

// order statistics at (<= 4) ranks of v[0..n) -> host out4
b2r_status select_ranks(b2r_solver* s, const double* v, int n, const int* ranks, int nr, int sqrt_out, double* dev_out4, double* host_out4) {
  if (n > kSelectMax) return fail(s, B2R_ERR_LIMIT, "order statistics over more than 131072 values are not supported");
  Ranks w; for (int q = 0; q < 4; ++q) w.k[q] = q < nr ? ranks[q] : -1;
  select_kernel<<<(n + 255) / 256, 256, 0, s->stream>>>(v, n, w, sqrt_out, dev_out4);
  s->launches += 1;
  CK(cudaMemcpyAsync(host_out4, dev_out4, 4 * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return B2R_OK;
}

b2r_status voxel_coverage_dev(b2r_solver* s, const double* pts, int n, double voxel, int* table, unsigned cap, int* d_count) {
  CK(cudaMemsetAsync(table, 0, (size_t)cap * sizeof(int), s->stream));
  CK(cudaMemsetAsync(d_count, 0, sizeof(int), s->stream));
  if (n > 0) { voxel_count_kernel<<<(n + 255) / 256, 256, 0, s->stream>>>(pts, n, voxel, table, cap - 1, d_count); s->launches += 1; }
  return B2R_OK;
}

unsigned table_cap(int n) { unsigned c = 256; while (c < 2u * (unsigned)(n > 0 ? n : 1)) c <<= 1; return c; }

}  // namespace

extern "C" {

b2r_status b2r_transform_points(b2r_solver* s, const double* T16, const double* in_dev, int n, double* out_dev) {
  B2R_NVTX("b2r:transform_points");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!T16 || !in_dev || !out_dev || n < 0) return fail(s, B2R_ERR_INVALID_ARG, "transform_points arguments");
  CK(cudaSetDevice(s->device));
  Mat12 T; for (int k = 0; k < 12; ++k) T.m[k] = T16[k];
  if (n > 0) { transform_kernel<<<(n + 255) / 256, 256, 0, s->stream>>>(in_dev, n, T, out_dev); s->launches += 1; }
  CK(cudaGetLastError());
  return B2R_OK;
}

b2r_status b2r_nn_distance(b2r_solver* s, const double* q_dev, int nq, const double* ref_dev, int nr, double* dist2_dev, int32_t* idx_dev) {
  B2R_NVTX("b2r:nn_distance");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!q_dev || !ref_dev || nq < 0 || nr < 1 || (!dist2_dev && !idx_dev)) return fail(s, B2R_ERR_INVALID_ARG, "nn_distance arguments");
  Ws w; b2r_status r = refine_ws(s, (size_t)nq * 16 + 1024, &w);
  if (r != B2R_OK) return r;
  int* cnt = w.take<int>(nq);
  int* idx = idx_dev ? idx_dev : w.take<int>(nq);
  double* d2 = dist2_dev ? dist2_dev : w.take<double>(nq);
  r = run_nn1(s, q_dev, nq, ref_dev, nr, 0.0, idx, d2, cnt, nullptr);
  if (r != B2R_OK) return r;
  CK(cudaGetLastError());
  return B2R_OK;
}

b2r_status b2r_icp(b2r_solver* s, const double* src_dev, int ns, const double* dst_dev, int nd, double max_dist, const double* init16,
                   int max_iter, double rel_fitness, double rel_rmse, b2r_icp_result* out, int32_t* corr_dev) {
  B2R_NVTX("b2r:icp");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!src_dev || !dst_dev || !out || ns < 1 || nd < 1 || !(max_dist > 0) || max_iter < 0) return fail(s, B2R_ERR_INVALID_ARG, "icp arguments");
  Ws w; b2r_status r = refine_ws(s, (size_t)ns * (24 + 8 + 4 + 4) + 4096, &w);
  if (r != B2R_OK) return r;
  cudaStream_t st = s->stream;
  double* cur = w.take<double>((size_t)3 * ns);
  double* d2 = w.take<double>(ns);
  int* corr = w.take<int>(ns);
  int* cnt = w.take<int>(ns);
  IcpState* state = w.take<IcpState>(1);
  IcpState* h = (IcpState*)s->h_refine.p;
  memset(h, 0, sizeof(IcpState));
  bool ident = true;
  for (int k = 0; k < 16; ++k) { h->T[k] = init16 ? init16[k] : ((k % 5 == 0) ? 1.0 : 0.0); ident = ident && h->T[k] == ((k % 5 == 0) ? 1.0 : 0.0); }
  CK(cudaMemcpyAsync(state, h, sizeof(IcpState), cudaMemcpyHostToDevice, st));
  if (ident) CK(cudaMemcpyAsync(cur, src_dev, sizeof(double) * 3 * ns, cudaMemcpyDeviceToDevice, st));
  else { Mat12 T; for (int k = 0; k < 12; ++k) T.m[k] = h->T[k]; transform_kernel<<<(ns + 255) / 256, 256, 0, st>>>(src_dev, ns, T, cur); s->launches += 1; }
  const double r2 = max_dist * max_dist;
  const int* done = &state->done;
  // evaluation, then chunks of [step, move, evaluate]; the flag is read back once per chunk
  r = run_nn1(s, cur, ns, dst_dev, nd, r2, corr, d2, cnt, done);
  if (r != B2R_OK) return r;
  int enq = 0;
  for (;;) {
    const int chunk = enq == 0 ? 8 : 16;
    for (int c = 0; c < chunk; ++c) {
      icp_step_kernel<<<1, 256, 0, st>>>(cur, ns, dst_dev, corr, d2, max_iter, rel_fitness, rel_rmse, state);
      icp_apply_kernel<<<(ns + 255) / 256, 256, 0, st>>>(cur, ns, state);
      s->launches += 2;
      r = run_nn1(s, cur, ns, dst_dev, nd, r2, corr, d2, cnt, nd <= kNnBruteMax ? done : nullptr);
      if (r != B2R_OK) return r;
    }
    enq += chunk;
    CK(cudaMemcpyAsync(h, state, sizeof(IcpState), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (h->done) break;
    if (enq > max_iter + 2) return fail(s, B2R_ERR_STATE, "internal error: ICP did not terminate");
  }
  memcpy(out->transformation, h->T, sizeof(double) * 16);
  out->fitness = h->fitness; out->inlier_rmse = h->rmse; out->n_corr = h->n_corr; out->iterations = h->iterations; out->converged = h->converged;
  out->pad_ = 0;
  if (corr_dev) CK(cudaMemcpyAsync(corr_dev, corr, sizeof(int) * ns, cudaMemcpyDeviceToDevice, st));
  CK(cudaGetLastError());
  return B2R_OK;
}

b2r_status b2r_voxel_coverage(b2r_solver* s, const double* pts_dev, int n, double voxel, int* count_out) {
  B2R_NVTX("b2r:voxel_coverage");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!pts_dev || n < 0 || !(voxel > 0) || !count_out) return fail(s, B2R_ERR_INVALID_ARG, "voxel_coverage arguments");
  const unsigned cap = table_cap(n);
  Ws w; b2r_status r = refine_ws(s, (size_t)cap * 4 + 1024, &w);
  if (r != B2R_OK) return r;
  int* table = w.take<int>(cap); int* d_count = w.take<int>(1);
  r = voxel_coverage_dev(s, pts_dev, n, voxel, table, cap, d_count);
  if (r != B2R_OK) return r;
  CK(cudaMemcpyAsync(count_out, d_count, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return B2R_OK;
}

b2r_status b2r_alignment_score(b2r_solver* s, const double* src_aligned_dev, int ns, const double* src_down_dev, int ns2,
                               const double* dst_dev, int nd, double voxel, double* out5) {
  B2R_NVTX("b2r:alignment_score");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!src_aligned_dev || !src_down_dev || !dst_dev || ns < 1 || ns2 < 0 || nd < 1 || !(voxel > 0) || !out5)
    return fail(s, B2R_ERR_INVALID_ARG, "alignment_score arguments");
  const unsigned cap = table_cap(ns > ns2 ? ns : ns2);
  Ws w; b2r_status r = refine_ws(s, (size_t)ns * 16 + 2 * (size_t)cap * 4 + 4096, &w);
  if (r != B2R_OK) return r;
  double* d2 = w.take<double>(ns);
  int* idx = w.take<int>(ns);
  int* cnt = w.take<int>(ns);
  int* table = w.take<int>(cap);
  int* table2 = w.take<int>(cap);
  int* d_counts = w.take<int>(2);
  double* d_out4 = w.take<double>(4);
  // nn_residuals: cKDTree(dst).query(src_aligned, k=1) distances
  r = run_nn1(s, src_aligned_dev, ns, dst_dev, nd, 0.0, idx, d2, cnt, nullptr);
  if (r != B2R_OK) return r;
  r = voxel_coverage_dev(s, src_aligned_dev, ns, voxel, table, cap, d_counts);
  if (r != B2R_OK) return r;
  r = voxel_coverage_dev(s, src_down_dev, ns2, voxel, table2, cap, d_counts + 1);
  if (r != B2R_OK) return r;
  // np.median and np.percentile(., 90): order statistics at n/2-1, n/2 and floor(0.9 (n-1)), +1
  const double vi = (double)(ns - 1) * 0.9;
  double lo = floor(vi); if (lo > ns - 1) lo = ns - 1;
  const int i0 = (int)lo, i1 = i0 + 1 < ns ? i0 + 1 : ns - 1;
  const int ranks[4] = {(ns & 1) ? ns / 2 : ns / 2 - 1, ns / 2, i0, i1};
  double h4[4];
  r = select_ranks(s, d2, ns, ranks, 4, 1, d_out4, h4);   // also synchronises the stream
  if (r != B2R_OK) return r;
  int hc[2];
  CK(cudaMemcpy(hc, d_counts, sizeof(hc), cudaMemcpyDeviceToHost));
  const double med = (ns & 1) ? h4[1] : (h4[0] + h4[1]) / 2.0;
  const double p90 = np_lerp(h4[2], h4[3], vi - lo);
  const double cov_norm = (double)hc[0] / (double)(hc[1] > 1 ? hc[1] : 1);
  out5[0] = med; out5[1] = p90; out5[2] = hc[0]; out5[3] = hc[1];
  out5[4] = (med + 0.3 * p90) + 0.5 * (1.0 - cov_norm);   // HelpersRealtime.py:129
  return B2R_OK;
}

b2r_status b2r_chamfer(b2r_solver* s, const double* a_dev, int na, const double* b_dev, int nb, double* out) {
  B2R_NVTX("b2r:chamfer");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!a_dev || !b_dev || na < 1 || nb < 1 || !out) return fail(s, B2R_ERR_INVALID_ARG, "chamfer arguments");
  const int nm = na > nb ? na : nb;
  Ws w; b2r_status r = refine_ws(s, (size_t)nm * 16 + 4096, &w);
  if (r != B2R_OK) return r;
  double* d2 = w.take<double>(nm); int* idx = w.take<int>(nm); int* cnt = w.take<int>(nm); double* sums = w.take<double>(2);
  r = run_nn1(s, a_dev, na, b_dev, nb, 0.0, idx, d2, cnt, nullptr);
  if (r != B2R_OK) return r;
  sum_sqrt_kernel<<<1, 256, 0, s->stream>>>(d2, na, sums, 0);
  r = run_nn1(s, b_dev, nb, a_dev, na, 0.0, idx, d2, cnt, nullptr);
  if (r != B2R_OK) return r;
  sum_sqrt_kernel<<<1, 256, 0, s->stream>>>(d2, nb, sums, 1);
  s->launches += 2;
  double h[2];
  CK(cudaMemcpyAsync(h, sums, sizeof(h), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  *out = h[0] / na + h[1] / nb;   // np.mean(d1) + np.mean(d2), registration_utils.py:280
  return B2R_OK;
}

b2r_status b2r_cloud_resolution(b2r_solver* s, const double* pts_dev, int n, int k, double* out) {
  B2R_NVTX("b2r:cloud_resolution");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!pts_dev || n < 0 || k < 1 || k > 127 || !out) return fail(s, B2R_ERR_INVALID_ARG, "cloud_resolution arguments");
  if (n < 2) { *out = 0.005; return B2R_OK; }          // registration_utils.py:231-232
  const int step = n / 5000 > 1 ? n / 5000 : 1;       // pts[::max(1, len(pts)//5000)]
  const int nq = (n + step - 1) / step;
  const int K = k + 1 < n ? k + 1 : n;
  Ws w; b2r_status r = refine_ws(s, (size_t)nq * (24 + (size_t)K * 12 + 4 + (size_t)K * 8) + 4096, &w);
  if (r != B2R_OK) return r;
  double* q = w.take<double>((size_t)3 * nq);
  int* idx = w.take<int>((size_t)nq * K); double* d2 = w.take<double>((size_t)nq * K); int* cnt = w.take<int>(nq);
  double* flat = w.take<double>((size_t)nq * (K - 1)); int* total = w.take<int>(1); double* d_out4 = w.take<double>(4);
  stride_gather_kernel<<<(nq + 255) / 256, 256, 0, s->stream>>>(pts_dev, nq, step, q);
  r = internal_knn(s, pts_dev, n, q, nq, K, 0.0, idx, d2, cnt);
  if (r != B2R_OK) return r;
  drop_self_kernel<<<(nq + 255) / 256, 256, 0, s->stream>>>(d2, cnt, nq, K, flat, total);
  s->launches += 2;
  const int m = nq * (K - 1);
  if (m < 1) { *out = 0.005; return B2R_OK; }
  const int ranks[2] = {(m & 1) ? m / 2 : m / 2 - 1, m / 2};
  double h4[4];
  r = select_ranks(s, flat, m, ranks, 2, 1, d_out4, h4);
  if (r != B2R_OK) return r;
  *out = (m & 1) ? h4[1] : (h4[0] + h4[1]) / 2.0;
  return B2R_OK;
}

b2r_status b2r_pca_axes(b2r_solver* s, const double* pts_dev, int n, double* R9, double* s3, double* c3) {
  B2R_NVTX("b2r:pca_axes");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!pts_dev || n < 1 || !R9) return fail(s, B2R_ERR_INVALID_ARG, "pca_axes arguments");
  Ws w; b2r_status r = refine_ws(s, 4096, &w);
  if (r != B2R_OK) return r;
  double* d = w.take<double>(30);
  PcaIn in; in.pts[0] = pts_dev; in.pts[1] = pts_dev; in.n[0] = n; in.n[1] = n;
  pca_kernel<<<1, 256, 0, s->stream>>>(in, d);
  s->launches += 1;
  double h[15];
  CK(cudaMemcpyAsync(h, d, sizeof(h), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  memcpy(R9, h, 9 * sizeof(double));
  if (s3) memcpy(s3, h + 9, 3 * sizeof(double));
  if (c3) memcpy(c3, h + 12, 3 * sizeof(double));
  return B2R_OK;
}

b2r_status b2r_pca_align(b2r_solver* s, const double* src_dev, int ns, const double* dst_dev, int nd, double* T16) {
  B2R_NVTX("b2r:pca_align");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!src_dev || !dst_dev || ns < 1 || nd < 1 || !T16) return fail(s, B2R_ERR_INVALID_ARG, "pca_align arguments");
  Ws w; b2r_status r = refine_ws(s, 4096, &w);
  if (r != B2R_OK) return r;
  double* d = w.take<double>(30);
  PcaIn in; in.pts[0] = src_dev; in.pts[1] = dst_dev; in.n[0] = ns; in.n[1] = nd;
  pca_kernel<<<2, 256, 0, s->stream>>>(in, d);
  s->launches += 1;
  double h[30];
  CK(cudaMemcpyAsync(h, d, sizeof(h), cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  // initial_align_centroid_pca (registration_utils.py:314-343): sign-align the source axes with the target's, keep det = +1
  const double* Rs = h; const double* cs = h + 12; const double* Rd = h + 15; const double* cd = h + 27;
  double Ra[9];
  memcpy(Ra, Rs, sizeof(Ra));
  for (int j = 0; j < 3; ++j) {
    const double dot = Rs[j] * Rd[j] + Rs[3 + j] * Rd[3 + j] + Rs[6 + j] * Rd[6 + j];
    if (dot < 0) for (int i = 0; i < 3; ++i) Ra[3 * i + j] = -Ra[3 * i + j];
  }
  const double det = Ra[0] * (Ra[4] * Ra[8] - Ra[5] * Ra[7]) - Ra[1] * (Ra[3] * Ra[8] - Ra[5] * Ra[6]) + Ra[2] * (Ra[3] * Ra[7] - Ra[4] * Ra[6]);
  if (det < 0) for (int i = 0; i < 3; ++i) Ra[3 * i + 2] = -Ra[3 * i + 2];
  for (int k = 0; k < 16; ++k) T16[k] = (k % 5 == 0) ? 1.0 : 0.0;
  for (int r3 = 0; r3 < 3; ++r3) {
    for (int c = 0; c < 3; ++c) T16[4 * r3 + c] = Rd[3 * r3] * Ra[3 * c] + Rd[3 * r3 + 1] * Ra[3 * c + 1] + Rd[3 * r3 + 2] * Ra[3 * c + 2];
    T16[4 * r3 + 3] = cd[r3] - (T16[4 * r3] * cs[0] + T16[4 * r3 + 1] * cs[1] + T16[4 * r3 + 2] * cs[2]);
  }
  return B2R_OK;
}

b2r_status b2r_farthest_point_sample(b2r_solver* s, const double* pts_dev, int n, int target, int first_index, int32_t* idx_dev) {
  B2R_NVTX("b2r:farthest_point_sample");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!pts_dev || !idx_dev || n < 1 || target < 1 || target > n || first_index < 0 || first_index >= n)
    return fail(s, B2R_ERR_INVALID_ARG, "farthest_point_sample arguments");
  Ws w; b2r_status r = refine_ws(s, (size_t)n * 8 + 1024, &w);
  if (r != B2R_OK) return r;
  double* dist = w.take<double>(n);
  fps_kernel<<<1, 1024, 0, s->stream>>>(pts_dev, n, target, first_index, dist, idx_dev);
  s->launches += 1;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(s->stream));
  return B2R_OK;
}

// ---- b2r_voxel_coverage_noalloc: the voxel-coverage count with its hash table in a file-scope device array (no allocation) ----
// No counterpart in the reference; see include/b2r.h.  The table of b2r_voxel_coverage lives in the solver's workspace, which grows
// with cudaMalloc; this variant keeps it in static device memory (4 MiB on every GPU that loads the library, until exit).
constexpr int kNoAllocMaxPoints = B2R_VOXEL_NOALLOC_MAX_POINTS;
constexpr unsigned kNoAllocTableCap = 2u * (unsigned)kNoAllocMaxPoints;   // table_cap(n) <= 2^20 slots for n <= 2^19
static_assert((kNoAllocTableCap & (kNoAllocTableCap - 1)) == 0, "the table size must be a power of two");
static __device__ __align__(256) int g_noalloc_table[kNoAllocTableCap];

// In host code g_noalloc_table is a host-side stand-in for the symbol, not a device address: the address the kernels are given comes
// from cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kNoAllocMaxDevices = 64;
static int* g_noalloc_ptr[kNoAllocMaxDevices];   // zero-initialised: not yet resolved
static std::mutex g_noalloc_mu;

b2r_status b2r_voxel_coverage_noalloc(b2r_solver* s, const double* pts_dev, int n, double voxel, int* count_dev) {
  B2R_NVTX("b2r:voxel_coverage_noalloc");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!pts_dev || n < 0 || !(voxel > 0) || !count_dev) return fail(s, B2R_ERR_INVALID_ARG, "voxel_coverage_noalloc arguments");
  if (n > kNoAllocMaxPoints)
    return fail(s, B2R_ERR_LIMIT, "voxel_coverage_noalloc: " + std::to_string(n) + " points exceed the static table's limit of " +
                                      std::to_string(kNoAllocMaxPoints) + " (B2R_VOXEL_NOALLOC_MAX_POINTS); use b2r_voxel_coverage");
  if (s->device < 0 || s->device >= kNoAllocMaxDevices) return fail(s, B2R_ERR_LIMIT, "voxel_coverage_noalloc: device index beyond 63");
  CK(cudaSetDevice(s->device));
  int* table = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[s->device]) {
      void* p = nullptr;
      CK(cudaGetSymbolAddress(&p, g_noalloc_table));
      g_noalloc_ptr[s->device] = static_cast<int*>(p);
    }
    table = g_noalloc_ptr[s->device];
  }
  // every call shares the table, which still holds the last call's point ids: clear the slots this call can touch, and the count
  const unsigned cap = table_cap(n);
  CK(cudaMemsetAsync(table, 0, (size_t)cap * sizeof(int), s->stream));
  CK(cudaMemsetAsync(count_dev, 0, sizeof(int), s->stream));
  if (n > 0) {
    voxel_count_kernel<<<(n + 255) / 256, 256, 0, s->stream>>>(pts_dev, n, voxel, table, cap - 1, count_dev);
    s->launches += 1;
  }
  CK(cudaGetLastError());
  return B2R_OK;
}

}  // extern "C"
