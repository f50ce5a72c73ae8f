// Rotation (GNC-TLS) and translation (adaptive voting) on the pruned correspondences.
//
// Replaces upstream's chain/complete TIMs over the clique + GNCTLSRotationSolver::solveForRotation (A.6) and
// TLSTranslationSolver / ScalarTLSEstimator::estimate (A.7), reached from `solver.solve`
// (registration_utils.py:255-261 sets GNC_TLS, factor 1.4, 100 iterations, cost threshold 1e-12).
// The problems are tiny (a few hundred TIMs), so the design goal is latency and batching: one CTA per
// registration, the whole GNC loop and the three per-axis votes run inside one launch each, reductions by warp
// shuffle, nothing returns to the host in between.
#include <math_constants.h>

#include "common.cuh"
#include "linalg.cuh"

namespace b2r {

#define FULL 0xffffffffu
constexpr int kRtThreads = 256;
constexpr int kRtWarps = kRtThreads / 32;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(FULL, v, o));
  return v;
}

// block-wide sum of K doubles per thread; result valid in every thread
template <int K>
__device__ void block_sum(double (&v)[K], double* red /* K * kRtWarps */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = warp_sum(v[k]);
  __syncthreads();
  if (lane == 0)
#pragma unroll
    for (int k = 0; k < K; ++k) red[k * kRtWarps + warp] = v[k];
  __syncthreads();
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double s = 0;
#pragma unroll
    for (int w = 0; w < kRtWarps; ++w) s += red[k * kRtWarps + w];
    v[k] = s;
  }
}

__device__ double block_max(double v, double* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double m = red[0];
#pragma unroll
  for (int w = 1; w < kRtWarps; ++w) m = fmax(m, red[w]);
  return m;
}


__global__ void __launch_bounds__(kRtThreads) rotation_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  const ProbDesc d = descs[blockIdx.x];
  ProbState* st = d.st;
  __shared__ double red[9 * kRtWarps];
  __shared__ double sR[9];
  __shared__ int sh_cnt;
  const int tid = threadIdx.x;
  const int C = st->clique_size;
  if (tid == 0) {
    if (!kp.estimate_scaling) st->scale = 1.0;  // else: written by the scale stage (scale.cu) before the graph kernel
    for (int i = 0; i < 9; ++i) st->R[i] = (i % 4 == 0) ? 1.0 : 0.0;
    st->t[0] = st->t[1] = st->t[2] = 0.0;
    st->valid = 0; st->n_rot_tims = 0; st->n_rot_inl = 0; st->n_trans_inl = 0; st->gnc_iters = 0; st->gnc_cost = 0.0;
  }
  if (C <= 1) return;  // A.5-6: solution stays invalid
  const double scale = kp.estimate_scaling ? st->scale : 1.0;
  const double inv_s = 1.0 / scale;
  long long Tll = kp.tim_graph == B2R_TIM_CHAIN ? (long long)C : (long long)C * (C - 1) / 2;
  if (Tll > d.tim_cap) {  // COMPLETE graph over a huge clique: report instead of truncating silently
    if (tid == 0) atomicOr(&st->flags, 1u << 4);
    return;
  }
  const int T = (int)Tll;
  const double* src = d.src; const double* dst = d.dst;
  // A.6-1 pruned TIMs
  if (kp.tim_graph == B2R_TIM_CHAIN) {
    for (int i = tid; i < T; i += kRtThreads) {
      const int root = d.clique[i], leaf = d.clique[(i + 1 == C) ? 0 : i + 1];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        d.timx[3 * (size_t)i + k] = src[3 * (size_t)leaf + k] - src[3 * (size_t)root + k];
        d.timy[3 * (size_t)i + k] = (dst[3 * (size_t)leaf + k] - dst[3 * (size_t)root + k]) * inv_s;
      }
      d.weights[i] = 1.0;
    }
  } else {
    for (int i = tid; i < C - 1; i += kRtThreads) {
      long long k = (long long)i * C - (long long)i * (i + 1) / 2;
      const int a = d.clique[i];
      for (int j = i + 1; j < C && k < T; ++j, ++k) {
        const int b = d.clique[j];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
          d.timx[3 * k + q] = src[3 * (size_t)b + q] - src[3 * (size_t)a + q];
          d.timy[3 * k + q] = (dst[3 * (size_t)b + q] - dst[3 * (size_t)a + q]) * inv_s;
        }
        d.weights[k] = 1.0;
      }
    }
  }
  __syncthreads();
  // A.6-2/3 GNC-TLS
  const double nb_base = kp.rot_noise_bound > 0 ? kp.rot_noise_bound : kp.noise_bound;
  const double nb = nb_base * (2.0 / scale);
  double nb2 = nb * nb;
  if (nb2 < 1e-16) nb2 = 1e-2;
  double mu = 1.0, prev_cost = CUDART_INF, cost = 0.0;
  double* r2buf = d.raw;  // scratch: T doubles
  int iters = 0;
  for (int it = 0; it < kp.max_iterations; ++it) {
    double h[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = tid; i < T; i += kRtThreads) {
      const double w = d.weights[i];
      const double x0 = d.timx[3 * (size_t)i], x1 = d.timx[3 * (size_t)i + 1], x2 = d.timx[3 * (size_t)i + 2];
      const double y0 = d.timy[3 * (size_t)i], y1 = d.timy[3 * (size_t)i + 1], y2 = d.timy[3 * (size_t)i + 2];
      h[0] += w * x0 * y0; h[1] += w * x0 * y1; h[2] += w * x0 * y2;
      h[3] += w * x1 * y0; h[4] += w * x1 * y1; h[5] += w * x1 * y2;
      h[6] += w * x2 * y0; h[7] += w * x2 * y1; h[8] += w * x2 * y2;
    }
    block_sum<9>(h, red);
    if (tid == 0) rotation_from_H(h, sR);
    __syncthreads();
    double maxr = 0.0;
    for (int i = tid; i < T; i += kRtThreads) {
      const double x0 = d.timx[3 * (size_t)i], x1 = d.timx[3 * (size_t)i + 1], x2 = d.timx[3 * (size_t)i + 2];
      const double e0 = d.timy[3 * (size_t)i] - ((sR[0] * x0 + sR[1] * x1) + sR[2] * x2);
      const double e1 = d.timy[3 * (size_t)i + 1] - ((sR[3] * x0 + sR[4] * x1) + sR[5] * x2);
      const double e2 = d.timy[3 * (size_t)i + 2] - ((sR[6] * x0 + sR[7] * x1) + sR[8] * x2);
      const double r2 = (e0 * e0 + e1 * e1) + e2 * e2;
      r2buf[i] = r2;
      maxr = fmax(maxr, r2);
    }
    if (it == 0) {
      maxr = block_max(maxr, red);
      mu = 1.0 / (2.0 * maxr / nb2 - 1.0);
      if (mu <= 0.0) break;  // near-noiseless: weights stay 1 (A.6-3c)
    }
    const double th1 = (mu + 1.0) / mu * nb2, th2 = mu / (mu + 1.0) * nb2;
    double cs[1] = {0.0};
    for (int i = tid; i < T; i += kRtThreads) {
      const double r2 = r2buf[i];
      double w = d.weights[i];
      cs[0] += w * r2;  // cost uses the weights of the previous iteration (A.6-3d)
      if (r2 >= th1) w = 0.0;
      else if (r2 <= th2) w = 1.0;
      else w = sqrt(nb2 * mu * (mu + 1.0) / r2) - mu;
      d.weights[i] = w;
    }
    block_sum<1>(cs, red);
    cost = cs[0];
    const double cost_diff = fabs(cost - prev_cost);
    mu *= kp.gnc_factor;
    prev_cost = cost;
    iters = it + 1;
    if (cost_diff < kp.cost_threshold) break;
  }
  __syncthreads();
  // rotation inliers (w >= 0.5), ascending positions
  if (tid == 0) sh_cnt = 0;
  __syncthreads();
  for (int i0 = 0; i0 < T; i0 += kRtThreads) {
    const int i = i0 + tid;
    const bool in = i < T && d.weights[i] >= 0.5;
    if (i < T) d.rot_mask[i] = in ? 1 : 0;
    const uint32_t bal = __ballot_sync(FULL, in);
    const int lane = tid & 31, warp = tid >> 5;
    __shared__ int wcnt[kRtWarps];
    if (lane == 0) wcnt[warp] = __popc(bal);
    __syncthreads();
    int base = sh_cnt;
    for (int w = 0; w < warp; ++w) base += wcnt[w];
    if (in) d.rot_inl[base + __popc(bal & ((1u << lane) - 1u))] = i;
    __syncthreads();
    if (tid == 0) { int s = 0; for (int w = 0; w < kRtWarps; ++w) s += wcnt[w]; sh_cnt += s; }
    __syncthreads();
  }
  if (tid == 0) {
    for (int i = 0; i < 9; ++i) st->R[i] = sR[i];
    st->n_rot_tims = T; st->n_rot_inl = sh_cnt; st->gnc_iters = iters; st->gnc_cost = cost;
  }
}

// ---------------------------------------------------------------------------------------------------------
// Adaptive voting (A.7): endpoints sorted by (value, entry before exit, index) with a bitonic network, running
// sums by a two-level scan, first minimum of the cost.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool ep_less(double va, int ia, double vb, int ib) {
  if (va != vb) return va < vb;
  const bool ae = ia > 0, be = ib > 0;
  if (ae != be) return ae;
  return abs(ia) < abs(ib);
}

__device__ void bitonic_sort(double* val, int* id, int Epad) {
  for (int k = 2; k <= Epad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      __syncthreads();
      for (int i = threadIdx.x; i < Epad; i += blockDim.x) {
        const int l = i ^ j;
        if (l > i) {
          const double vi = val[i], vl = val[l];
          const int ii = id[i], il = id[l];
          const bool up = (i & k) == 0;
          const bool swap = up ? ep_less(vl, il, vi, ii) : ep_less(vi, ii, vl, il);
          if (swap) { val[i] = vl; val[l] = vi; id[i] = il; id[l] = ii; }
        }
      }
    }
  }
  __syncthreads();
}

// One TLS estimate over X[0..M) with a constant range r.  Returns the estimate in every thread.
__device__ double tls_estimate_const_range(const double* X, int M, double r, double* val, int* id, double* sm /*6*kRtThreads+8*/) {
  const int tid = threadIdx.x;
  const int E = 2 * M;
  int Epad = 2;
  while (Epad < E) Epad <<= 1;
  for (int e = tid; e < Epad; e += kRtThreads) {
    if (e < E) {
      const int i = e >> 1;
      if (e & 1) { val[e] = X[i] + r; id[e] = -(i + 1); }
      else { val[e] = X[i] - r; id[e] = i + 1; }
    } else { val[e] = CUDART_INF; id[e] = 0x7fffffff; }
  }
  bitonic_sort(val, id, Epad);
  const double w = 1.0 / (r * r);
  const int chunk = (E + kRtThreads - 1) / kRtThreads;
  const int e0 = min(tid * chunk, E), e1 = min(e0 + chunk, E);
  double s_w = 0, s_xw = 0, s_r = 0, s_x = 0, s_xx = 0, s_c = 0;
  for (int e = e0; e < e1; ++e) {
    const int idx = abs(id[e]) - 1;
    const double eps = id[e] > 0 ? 1.0 : -1.0;
    const double x = X[idx];
    s_c += eps; s_w += eps * w; s_xw += eps * w * x; s_r += eps * r; s_x += eps * x; s_xx += eps * x * x;
  }
  __syncthreads();
  sm[0 * kRtThreads + tid] = s_c; sm[1 * kRtThreads + tid] = s_w; sm[2 * kRtThreads + tid] = s_xw;
  sm[3 * kRtThreads + tid] = s_r; sm[4 * kRtThreads + tid] = s_x; sm[5 * kRtThreads + tid] = s_xx;
  __syncthreads();
  if (tid < 6) {  // exclusive scan of one quantity per thread
    double run = 0;
    double* a = sm + tid * kRtThreads;
    for (int i = 0; i < kRtThreads; ++i) { const double t = a[i]; a[i] = run; run += t; }
  }
  __syncthreads();
  double card = sm[0 * kRtThreads + tid], dw = sm[1 * kRtThreads + tid], dxw = sm[2 * kRtThreads + tid];
  double rsum = (double)M * r - sm[3 * kRtThreads + tid];
  double sx = sm[4 * kRtThreads + tid], sxx = sm[5 * kRtThreads + tid];
  double best_cost = CUDART_INF, best_x = 0.0;
  int best_e = 0x7fffffff;
  for (int e = e0; e < e1; ++e) {
    const int idx = abs(id[e]) - 1;
    const double eps = id[e] > 0 ? 1.0 : -1.0;
    const double x = X[idx];
    card += eps; dw += eps * w; dxw += eps * w * x; rsum -= eps * r; sx += eps * x; sxx += eps * x * x;
    const double xh = dxw / dw;
    const double cost = (card * xh * xh + sxx - 2 * sx * xh) + rsum;
    if (cost < best_cost) { best_cost = cost; best_x = xh; best_e = e; }
  }
  // argmin over threads: lowest cost, ties -> lowest endpoint position (first minimum)
  __syncthreads();
  sm[tid] = best_cost; sm[kRtThreads + tid] = best_x; sm[2 * kRtThreads + tid] = (double)best_e;
  __syncthreads();
  for (int s = kRtThreads / 2; s > 0; s >>= 1) {
    if (tid < s) {
      const double ca = sm[tid], cb = sm[tid + s];
      const double ea = sm[2 * kRtThreads + tid], eb = sm[2 * kRtThreads + tid + s];
      if (cb < ca || (cb == ca && eb < ea)) { sm[tid] = cb; sm[kRtThreads + tid] = sm[kRtThreads + tid + s]; sm[2 * kRtThreads + tid] = eb; }
    }
    __syncthreads();
  }
  const double est = (sm[2 * kRtThreads] < 2.0e9) ? sm[kRtThreads] : CUDART_NAN;
  __syncthreads();
  return est;
}

__global__ void __launch_bounds__(kRtThreads) translation_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  extern __shared__ double shd[];  // [6*kRtThreads + 8] scan scratch, then optional endpoint staging
  const ProbDesc d = descs[blockIdx.x];
  ProbState* st = d.st;
  const int tid = threadIdx.x;
  const int C = st->clique_size;
  if (C <= 1) return;
  const int M = st->n_rot_inl;
  const double scale = st->scale;
  double R[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = st->R[i];
  double* raw = d.raw;  // 3 x M (row r at raw + r*M)
  for (int i = tid; i < M; i += kRtThreads) {
    const int pt = d.clique[d.rot_inl[i] % C];  // upstream indexes max_clique_ by TIM position (A.8)
    const double x0 = d.src[3 * (size_t)pt], x1 = d.src[3 * (size_t)pt + 1], x2 = d.src[3 * (size_t)pt + 2];
#pragma unroll
    for (int r = 0; r < 3; ++r)
      raw[(size_t)r * M + i] = d.dst[3 * (size_t)pt + r] - scale * ((R[3 * r] * x0 + R[3 * r + 1] * x1) + R[3 * r + 2] * x2);
    d.trans_mask[i] = 1;
  }
  __syncthreads();
  double t[3] = {0, 0, 0};
  if (M >= 2) {
    const double r = kp.noise_bound * sqrt(kp.cbar2);  // A.7-4: 1x, not 2x
    int Epad = 2;
    while (Epad < 2 * M) Epad <<= 1;
    double* sm = shd;
    double* val = d.tls_val; int* id = d.tls_id;
    if (Epad <= 2048) {  // endpoints fit in shared memory (24 KB)
      val = shd + 6 * kRtThreads + 8;
      id = (int*)(val + 2048);
    }
    for (int ax = 0; ax < 3; ++ax) {
      const double* X = raw + (size_t)ax * M;
      const double est = tls_estimate_const_range(X, M, r, val, id, sm);
      t[ax] = est;
      for (int i = tid; i < M; i += kRtThreads)
        if (!(fabs(X[i] - est) <= r)) d.trans_mask[i] = 0;
      __syncthreads();
    }
  } else if (M == 1) {
    t[0] = raw[0]; t[1] = raw[1]; t[2] = raw[2];
  }
  // count translation inliers
  __shared__ int sh_c;
  if (tid == 0) sh_c = 0;
  __syncthreads();
  int c = 0;
  for (int i = tid; i < M; i += kRtThreads) c += d.trans_mask[i];
  c = __reduce_add_sync(FULL, c);
  if ((tid & 31) == 0) atomicAdd(&sh_c, c);
  __syncthreads();
  if (tid == 0) {
    st->t[0] = t[0]; st->t[1] = t[1]; st->t[2] = t[2];
    st->n_trans_inl = sh_c;
    st->valid = 1;
  }
}

int launch_rotation_translation(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, cudaStream_t st,
                                cudaEvent_t* ev) {
  static bool attr = false;
  const size_t tsmem = (6 * kRtThreads + 8) * sizeof(double) + 2048 * (sizeof(double) + sizeof(int));
  if (!attr) { cudaFuncSetAttribute(translation_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsmem); attr = true; }
  rotation_kernel<<<batch, kRtThreads, 0, st>>>(d_desc, kp);
  if (ev) cudaEventRecord(ev[0], st);
  translation_kernel<<<batch, kRtThreads, tsmem, st>>>(d_desc, kp);
  (void)n_max;
  return 2;
}

}  // namespace b2r
