// Device code of the rotation (GNC-TLS) and translation (adaptive voting) stages, shared by rt.cu (one
// kernel per stage) and tail.cu (fused tail): a group of kRtThreads threads runs the rotation on a named barrier, three such
// groups vote the three translation axes concurrently.
//
// Replaces upstream's chain/complete TIMs over the clique + GNCTLSRotationSolver::solveForRotation (A.6) and
// TLSTranslationSolver / ScalarTLSEstimator::estimate (A.7), reached from `solver.solve`
// (registration_utils.py:255-261 sets GNC_TLS, factor 1.4, 100 iterations, cost threshold 1e-12).
// The problems are tiny (a few hundred TIMs), so the design goal is latency and batching: one CTA per
// registration, the whole GNC loop and the three per-axis votes run inside one launch each, reductions by warp
// shuffle, nothing returns to the host in between.
#pragma once
#include <math_constants.h>

#include "common.cuh"
#include "linalg.cuh"

namespace b2r {

#ifndef FULL
#define FULL 0xffffffffu
#endif
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

// a group of kRtThreads threads synchronises on its own named barrier (0 = the CTA barrier of a kRtThreads-thread kernel)
__device__ __forceinline__ void grp_sync(int bar) { asm volatile("bar.sync %0, %1;" ::"r"(bar), "n"(kRtThreads) : "memory"); }

// group-wide sum of K doubles per thread; result valid in every thread
template <int K>
__device__ void block_sum(double (&v)[K], double* red /* K * kRtWarps */, int bar) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = warp_sum(v[k]);
  grp_sync(bar);
  if (lane == 0)
#pragma unroll
    for (int k = 0; k < K; ++k) red[k * kRtWarps + warp] = v[k];
  grp_sync(bar);
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double s = 0;
#pragma unroll
    for (int w = 0; w < kRtWarps; ++w) s += red[k * kRtWarps + w];
    v[k] = s;
  }
}

__device__ inline double block_max(double v, double* red, int bar) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_max(v);
  grp_sync(bar);
  if (lane == 0) red[warp] = v;
  grp_sync(bar);
  double m = red[0];
#pragma unroll
  for (int w = 1; w < kRtWarps; ++w) m = fmax(m, red[w]);
  return m;
}


// The rotation stage of one registration on threads 0 .. kRtThreads-1 of the CTA (named barrier `bar`).  Returns false when the
// solution stays invalid (clique <= 1, COMPLETE TIM graph beyond the workspace).
// cache (optional): cache_doubles doubles of shared memory for the pruned TIMs, their weights and residuals (8 T doubles); the GNC loop then
// never touches global memory (each of its ~10 iterations made two dependent passes over arrays in L2).
__device__ inline bool rotation_body(const ProbDesc& d, const KernelParams& kp, int bar, TailMarks tm = TailMarks(), double* cache = nullptr,
                                     size_t cache_doubles = 0) {
  ProbState* st = d.st;
  __shared__ double red[9 * kRtWarps];
  __shared__ double sR[9];
  __shared__ double sV[9];
  __shared__ int sh_cnt;
  const int tid = threadIdx.x;
  const int C = st->clique_size;
  if (tid == 0) {
    if (!kp.estimate_scaling) st->scale = 1.0;  // else: written by the scale stage (scale.cu) before the graph kernel
    for (int i = 0; i < 9; ++i) st->R[i] = (i % 4 == 0) ? 1.0 : 0.0;
    st->t[0] = st->t[1] = st->t[2] = 0.0;
    st->valid = 0; st->n_rot_tims = 0; st->n_rot_inl = 0; st->n_trans_inl = 0; st->gnc_iters = 0; st->gnc_cost = 0.0;
  }
  if (C <= 1) return false;  // A.5-6: solution stays invalid
  const double scale = kp.estimate_scaling ? st->scale : 1.0;
  const double inv_s = 1.0 / scale;
  long long Tll = kp.tim_graph == B2R_TIM_CHAIN ? (long long)C : (long long)C * (C - 1) / 2;
  if (Tll > d.tim_cap) {  // COMPLETE graph over a huge clique: report instead of truncating silently
    if (tid == 0) atomicOr(&st->flags, 1u << 4);
    return false;
  }
  const int T = (int)Tll;
  const double* src = d.src; const double* dst = d.dst;
  // where the TIMs live during the GNC loop: shared memory when the caller has room for 8 T doubles, else the global workspace
  const bool cached = cache != nullptr && (size_t)8 * (size_t)T <= cache_doubles;
  double* timx = cached ? cache : d.timx;
  double* timy = cached ? cache + 3 * (size_t)T : d.timy;
  double* wts = cached ? cache + 6 * (size_t)T : d.weights;
  if (tid < 9) sV[tid] = (tid % 4 == 0) ? 1.0 : 0.0;
  // A.6-1 pruned TIMs
  if (kp.tim_graph == B2R_TIM_CHAIN) {
    for (int i = tid; i < T; i += kRtThreads) {
      const int root = d.clique[i], leaf = d.clique[(i + 1 == C) ? 0 : i + 1];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        timx[3 * (size_t)i + k] = src[3 * (size_t)leaf + k] - src[3 * (size_t)root + k];
        timy[3 * (size_t)i + k] = (dst[3 * (size_t)leaf + k] - dst[3 * (size_t)root + k]) * inv_s;
      }
      wts[i] = 1.0;
    }
  } else {
    for (int i = tid; i < C - 1; i += kRtThreads) {
      long long k = (long long)i * C - (long long)i * (i + 1) / 2;
      const int a = d.clique[i];
      for (int j = i + 1; j < C && k < T; ++j, ++k) {
        const int b = d.clique[j];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
          timx[3 * k + q] = src[3 * (size_t)b + q] - src[3 * (size_t)a + q];
          timy[3 * k + q] = (dst[3 * (size_t)b + q] - dst[3 * (size_t)a + q]) * inv_s;
        }
        wts[k] = 1.0;
      }
    }
  }
  grp_sync(bar);
  tm.mark(28);   // TIMs built
  // A.6-2/3 GNC-TLS
  const double nb_base = kp.rot_noise_bound > 0 ? kp.rot_noise_bound : kp.noise_bound;
  const double nb = nb_base * (2.0 / scale);
  double nb2 = nb * nb;
  if (nb2 < 1e-16) nb2 = 1e-2;
  double mu = 1.0, prev_cost = CUDART_INF, cost = 0.0;
  double* r2buf = cached ? cache + 7 * (size_t)T : d.raw;  // scratch: T doubles
  int iters = 0;
  // Few TIMs (the usual case on real correspondences: a clique of 3 .. 200): the whole loop runs inside WARP 0 with no barrier at all.
  // Every lane keeps its <= 8 TIMs in registers, the sums are warp shuffles, and every lane evaluates the 3 x 3 rotation itself
  // (same inputs, same result) instead of one thread computing it for 255 waiting ones.  With tiny cliques GNC runs tens of
  // iterations (cost threshold 1e-12), and at 6 us per iteration of the block-wide loop that was most of a config-2 solve.
  __shared__ int s_iters;
  __shared__ double s_cost;
  const bool warp_path = cached && T <= 8 * 32;
  if (warp_path) {
    if (tid < 32) {
      const int lane = tid;
      // TIMs are read from the shared-memory cache every pass (the 48 doubles per lane would not fit the fused tail's 85 registers);
      // weights and residuals live in registers
      double w[8], r2v[8];
      const int nq = (T - lane + 31) / 32;          // TIMs of this lane: lane, lane + 32, ...
#pragma unroll
      for (int q = 0; q < 8; ++q) { w[q] = q < nq ? 1.0 : 0.0; r2v[q] = 0.0; }
      double Rw[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
      double Vw[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
      for (int it = 0; it < kp.max_iterations; ++it) {
        double h[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (q < nq) {
            const double* xp = timx + 3 * (size_t)(lane + 32 * q);
            const double* yp = timy + 3 * (size_t)(lane + 32 * q);
            const double y0 = yp[0], y1 = yp[1], y2 = yp[2];
#pragma unroll
            for (int a = 0; a < 3; ++a) { const double wx = w[q] * xp[a]; h[3 * a] += wx * y0; h[3 * a + 1] += wx * y1; h[3 * a + 2] += wx * y2; }
          }
        }
#pragma unroll
        // the nine butterfly reductions step by step, not sum by sum: the same additions in the same order (bit-identical), but
        // nine independent shuffle + add pairs per step instead of 45 dependent ones (2.2 k -> ~0.6 k cycles, measured with cycle marks)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          double t9[9];
#pragma unroll
          for (int k = 0; k < 9; ++k) t9[k] = __shfl_xor_sync(FULL, h[k], o);
#pragma unroll
          for (int k = 0; k < 9; ++k) h[k] += t9[k];
        }
        if (!rotation_from_H_polar(h, Rw)) rotation_from_H(h, Rw, Vw);
        double maxr = 0.0;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (q < nq) {
            const double* xp = timx + 3 * (size_t)(lane + 32 * q);
            const double* yp = timy + 3 * (size_t)(lane + 32 * q);
            const double x0 = xp[0], x1 = xp[1], x2 = xp[2];
            const double e0 = yp[0] - ((Rw[0] * x0 + Rw[1] * x1) + Rw[2] * x2);
            const double e1 = yp[1] - ((Rw[3] * x0 + Rw[4] * x1) + Rw[5] * x2);
            const double e2 = yp[2] - ((Rw[6] * x0 + Rw[7] * x1) + Rw[8] * x2);
            r2v[q] = (e0 * e0 + e1 * e1) + e2 * e2;
            maxr = fmax(maxr, r2v[q]);
          }
        }
        if (it == 0) {
          maxr = warp_max(maxr);
          mu = 1.0 / (2.0 * maxr / nb2 - 1.0);
          if (mu <= 0.0) break;  // near-noiseless: weights stay 1 (A.6-3c)
        }
        const double th1 = (mu + 1.0) / mu * nb2, th2 = mu / (mu + 1.0) * nb2;
        double cs = 0.0;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const bool ok = q < nq;
          cs += w[q] * r2v[q];
          double nw;
          if (r2v[q] >= th1) nw = 0.0;
          else if (r2v[q] <= th2) nw = 1.0;
          else nw = sqrt(nb2 * mu * (mu + 1.0) / r2v[q]) - mu;
          w[q] = ok ? nw : 0.0;
        }
        cost = warp_sum(cs);
        const double cost_diff = fabs(cost - prev_cost);
        mu *= kp.gnc_factor;
        prev_cost = cost;
        iters = it + 1;
        if (cost_diff < kp.cost_threshold) break;
      }
#pragma unroll
      for (int q = 0; q < 8; ++q) { const int i = lane + 32 * q; if (i < T) wts[i] = w[q]; }
      if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 9; ++k) sR[k] = Rw[k];
        s_iters = iters; s_cost = cost;
      }
    }
    grp_sync(bar);
    iters = s_iters; cost = s_cost;
  }
  for (int it = 0; !warp_path && it < kp.max_iterations; ++it) {
    double h[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int i = tid; i < T; i += kRtThreads) {
      const double w = wts[i];
      const double x0 = timx[3 * (size_t)i], x1 = timx[3 * (size_t)i + 1], x2 = timx[3 * (size_t)i + 2];
      const double y0 = timy[3 * (size_t)i], y1 = timy[3 * (size_t)i + 1], y2 = timy[3 * (size_t)i + 2];
      h[0] += w * x0 * y0; h[1] += w * x0 * y1; h[2] += w * x0 * y2;
      h[3] += w * x1 * y0; h[4] += w * x1 * y1; h[5] += w * x1 * y2;
      h[6] += w * x2 * y0; h[7] += w * x2 * y1; h[8] += w * x2 * y2;
    }
    block_sum<9>(h, red, bar);
    if (tid == 0 && !rotation_from_H_polar(h, sR)) rotation_from_H(h, sR, sV);
    grp_sync(bar);
    double maxr = 0.0;
    for (int i = tid; i < T; i += kRtThreads) {
      const double x0 = timx[3 * (size_t)i], x1 = timx[3 * (size_t)i + 1], x2 = timx[3 * (size_t)i + 2];
      const double e0 = timy[3 * (size_t)i] - ((sR[0] * x0 + sR[1] * x1) + sR[2] * x2);
      const double e1 = timy[3 * (size_t)i + 1] - ((sR[3] * x0 + sR[4] * x1) + sR[5] * x2);
      const double e2 = timy[3 * (size_t)i + 2] - ((sR[6] * x0 + sR[7] * x1) + sR[8] * x2);
      const double r2 = (e0 * e0 + e1 * e1) + e2 * e2;
      r2buf[i] = r2;
      maxr = fmax(maxr, r2);
    }
    if (it == 0) {
      maxr = block_max(maxr, red, bar);
      mu = 1.0 / (2.0 * maxr / nb2 - 1.0);
      if (mu <= 0.0) break;  // near-noiseless: weights stay 1 (A.6-3c)
    }
    const double th1 = (mu + 1.0) / mu * nb2, th2 = mu / (mu + 1.0) * nb2;
    double cs[1] = {0.0};
    for (int i = tid; i < T; i += kRtThreads) {
      const double r2 = r2buf[i];
      double w = wts[i];
      cs[0] += w * r2;  // cost uses the weights of the previous iteration (A.6-3d)
      if (r2 >= th1) w = 0.0;
      else if (r2 <= th2) w = 1.0;
      else w = sqrt(nb2 * mu * (mu + 1.0) / r2) - mu;
      wts[i] = w;
    }
    block_sum<1>(cs, red, bar);
    cost = cs[0];
    const double cost_diff = fabs(cost - prev_cost);
    mu *= kp.gnc_factor;
    prev_cost = cost;
    iters = it + 1;
    if (cost_diff < kp.cost_threshold) break;
  }
  grp_sync(bar);
  if (cached) {
    for (int i = tid; i < T; i += kRtThreads) d.weights[i] = wts[i];   // getters (b2r_get_gnc_weights) and the inlier list read the global copy
    grp_sync(bar);
  }
  tm.mark(29);   // GNC done
  // rotation inliers (w >= 0.5), ascending positions
  if (tid == 0) sh_cnt = 0;
  grp_sync(bar);
  for (int i0 = 0; i0 < T; i0 += kRtThreads) {
    const int i = i0 + tid;
    const bool in = i < T && d.weights[i] >= 0.5;
    if (i < T) d.rot_mask[i] = in ? 1 : 0;
    const uint32_t bal = __ballot_sync(FULL, in);
    const int lane = tid & 31, warp = tid >> 5;
    __shared__ int wcnt[kRtWarps];
    if (lane == 0) wcnt[warp] = __popc(bal);
    grp_sync(bar);
    int base = sh_cnt;
    for (int w = 0; w < warp; ++w) base += wcnt[w];
    if (in) d.rot_inl[base + __popc(bal & ((1u << lane) - 1u))] = i;
    grp_sync(bar);
    if (tid == 0) { int s = 0; for (int w = 0; w < kRtWarps; ++w) s += wcnt[w]; sh_cnt += s; }
    grp_sync(bar);
  }
  if (tid == 0) {
    for (int i = 0; i < 9; ++i) st->R[i] = sR[i];
    st->n_rot_tims = T; st->n_rot_inl = sh_cnt; st->gnc_iters = iters; st->gnc_cost = cost;
  }
  return true;
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

// The three axes of the translation vote run concurrently: thread group g = threadIdx.x / kRtThreads owns axis g and
// synchronises on its own named barrier (1 + g); tid is the index inside the group.

__device__ inline void bitonic_sort(double* val, int* id, int Epad, int tid, int bar) {
  for (int k = 2; k <= Epad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      grp_sync(bar);
      for (int i = tid; i < Epad; i += kRtThreads) {
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
  grp_sync(bar);
}

// The same sort for Epad = kRtThreads * R endpoints in shared memory (R = 1, 2, 4, 8), elements in REGISTERS: thread t owns
// e = t + kRtThreads * r.  Of the log2(Epad) (log2(Epad) + 1) / 2 compare-exchange stages only those with a partner in another
// warp (strides 32 .. 128) go through shared memory and the group barrier; strides >= kRtThreads pair two registers of one
// thread and strides <= 16 are warp shuffles -- at 512 endpoints 9 barrier stages instead of 45, every thread busy in all of them.
// The stage is bound by instruction issue (3 axes x 512 endpoints x 45 stages on one SM), so the keys are integers: the value
// as an order-preserving 64-bit pattern (-0 folded into +0 first, as `va != vb` does) and (exit before entry flag, |id|) as 32
// bits -- a compare is 3 integer compares instead of two fp64 compares and two branches.  ep_less is a strict total order (ids
// are unique), so the result is the one bitonic_sort leaves.
struct EpKey { unsigned long long hi; unsigned lo; };
__device__ __forceinline__ EpKey ep_key(double v, int id) {
  const long long b = __double_as_longlong(v + 0.0);
  EpKey k;
  k.hi = (unsigned long long)b ^ ((unsigned long long)(b >> 63) | 0x8000000000000000ull);
  k.lo = id > 0 ? (unsigned)id : (0x80000000u | (unsigned)(-id));
  if (id == 0x7fffffff) { k.hi = ~0ull; k.lo = ~0u; }   // padding: behind every real endpoint, NaN values included
  return k;
}
__device__ __forceinline__ bool key_less(const EpKey& a, const EpKey& b) { return (a.hi < b.hi) | ((a.hi == b.hi) & (a.lo < b.lo)); }

template <int R>
__device__ __forceinline__ void bitonic_sort_regs(double* val, int* id, int tid, int bar) {
  constexpr int Epad = kRtThreads * R;
  EpKey key[R];
  grp_sync(bar);
#pragma unroll
  for (int r = 0; r < R; ++r) key[r] = ep_key(val[tid + kRtThreads * r], id[tid + kRtThreads * r]);
  unsigned long long* kh = reinterpret_cast<unsigned long long*>(val);   // the exchange stages reuse the arrays for the keys
  unsigned* kl = reinterpret_cast<unsigned*>(id);
#pragma unroll 1
  for (int k = 2; k <= Epad; k <<= 1) {
#pragma unroll 1
    for (int j = k >> 1; j > 0; j >>= 1) {
      if (j >= kRtThreads) {  // partner in another register of this thread (register indices stay compile-time constants)
#pragma unroll
        for (int dr = R / 2; dr >= 1; dr >>= 1) {
          if (j != dr * kRtThreads) continue;
#pragma unroll
          for (int r = 0; r < R; ++r) {
            if ((r & dr) != 0) continue;
            const int r2 = r | dr;
            const bool up = ((tid + kRtThreads * r) & k) == 0;
            // ascending: the lower register keeps the smaller element
            const bool swap = key_less(key[r2], key[r]) == up;
            const EpKey a = key[r], b = key[r2];
            key[r].hi = swap ? b.hi : a.hi; key[r].lo = swap ? b.lo : a.lo;
            key[r2].hi = swap ? a.hi : b.hi; key[r2].lo = swap ? a.lo : b.lo;
          }
        }
      } else {
        EpKey pk[R];
        if (j >= 32) {        // partner in another warp: through shared memory
#pragma unroll
          for (int r = 0; r < R; ++r) { kh[tid + kRtThreads * r] = key[r].hi; kl[tid + kRtThreads * r] = key[r].lo; }
          grp_sync(bar);
#pragma unroll
          for (int r = 0; r < R; ++r) { pk[r].hi = kh[(tid ^ j) + kRtThreads * r]; pk[r].lo = kl[(tid ^ j) + kRtThreads * r]; }
          grp_sync(bar);
        } else {              // partner in another lane
#pragma unroll
          for (int r = 0; r < R; ++r) { pk[r].hi = __shfl_xor_sync(FULL, key[r].hi, j); pk[r].lo = __shfl_xor_sync(FULL, key[r].lo, j); }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int e = tid + kRtThreads * r;
          const bool keep_min = ((e & k) == 0) == ((e & j) == 0);
          const bool take = key_less(pk[r], key[r]) == keep_min;   // keys are distinct: !less(p, k) <=> less(k, p); no divergent branch
          key[r].hi = take ? pk[r].hi : key[r].hi;
          key[r].lo = take ? pk[r].lo : key[r].lo;
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const unsigned long long h = key[r].hi;
    const long long b = (long long)((h >> 63) ? (h ^ 0x8000000000000000ull) : ~h);
    const bool pad = key[r].lo == ~0u;
    val[tid + kRtThreads * r] = pad ? CUDART_INF : __longlong_as_double(b);
    id[tid + kRtThreads * r] = pad ? 0x7fffffff : ((key[r].lo & 0x80000000u) ? -(int)(key[r].lo & 0x7fffffffu) : (int)key[r].lo);
  }
  grp_sync(bar);
}

// exclusive prefix over the group's kRtThreads per-thread partials of K quantities (warp shuffles + one pass over the warp
// totals); sm: K * kRtWarps doubles
template <int K>
__device__ __forceinline__ void grp_exclusive_scan(double (&v)[K], double* sm, int tid, int bar) {
  const int lane = tid & 31, warp = tid >> 5;
  double inc[K];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double x = v[k];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const double y = __shfl_up_sync(FULL, x, o);
      if (lane >= o) x += y;
    }
    inc[k] = x;
  }
  grp_sync(bar);
  if (lane == 31)
#pragma unroll
    for (int k = 0; k < K; ++k) sm[k * kRtWarps + warp] = inc[k];
  grp_sync(bar);
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double base = 0.0;
    for (int w = 0; w < warp; ++w) base += sm[k * kRtWarps + w];
    v[k] = base + (inc[k] - v[k]);
  }
  grp_sync(bar);
}

// Constant range: all intervals have the same half-width r, so the entry endpoints fl(X - r) and the exit endpoints fl(X + r) are
// each ordered like X itself.  Sort the M values ONCE by (X, index) -- half the keys, 2.4x fewer compare-exchanges than sorting the
// 2 M endpoints -- and merge the two ordered sequences by rank: an entry goes to i + #{exits < it}, an exit to j + #{entries <= it}
// (entries before exits at equal values, as ep_less orders them).  Two different X may round to the same endpoint value, where the
// reference orders by index, not by X: such a pair (and any NaN) makes the shortcut decline and the full sort runs instead.
// On success id[0 .. 2 M) holds the endpoint ids in exactly the order bitonic_sort_regs leaves; val is scratch.
__device__ inline bool sort_endpoints_const_range(const double* X, int M, double r, double* val, int* id, double* sm, int tid, int bar) {
  int Mpad = kRtThreads;
  while (Mpad < M) Mpad <<= 1;
  grp_sync(bar);
  for (int i = tid; i < Mpad; i += kRtThreads) {
    if (i < M) { val[i] = X[i]; id[i] = i + 1; }
    else { val[i] = CUDART_INF; id[i] = 0x7fffffff; }
  }
  if (Mpad == kRtThreads) bitonic_sort_regs<1>(val, id, tid, bar);
  else if (Mpad == 2 * kRtThreads) bitonic_sort_regs<2>(val, id, tid, bar);
  else bitonic_sort_regs<4>(val, id, tid, bar);
  // sorted by (X, index): val[i], id[i] = index + 1.  Check the two endpoint sequences, then rank.
  int bad = 0;
  for (int i = tid; i < M; i += kRtThreads) {
    const double x = val[i];
    if (x != x) bad = 1;
    if (i + 1 < M) {
      const double y = val[i + 1];
      if (x != y && id[i] > id[i + 1] && ((x - r) == (y - r) || (x + r) == (y + r))) bad = 1;
    }
  }
  int* flag = reinterpret_cast<int*>(sm);
  if (tid == 0) *flag = 0;
  grp_sync(bar);
  if (bad) *flag = 1;
  grp_sync(bar);
  const bool declined = *flag != 0;
  grp_sync(bar);
  if (declined) return false;
  int pe[4], px[4], ix[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const int i = tid + kRtThreads * q;
    pe[q] = px[q] = -1; ix[q] = 0;
    if (i < M) {
      const double a = val[i] - r, b = val[i] + r;
      ix[q] = id[i];
      int lo = 0, hi = M;                       // #{j : fl(X_j + r) < a}
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (val[mid] + r < a) lo = mid + 1; else hi = mid; }
      pe[q] = i + lo;
      lo = 0; hi = M;                           // #{k : fl(X_k - r) <= b}
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (val[mid] - r <= b) lo = mid + 1; else hi = mid; }
      px[q] = i + lo;
    }
  }
  grp_sync(bar);
#pragma unroll
  for (int q = 0; q < 4; ++q)
    if (pe[q] >= 0) { id[pe[q]] = ix[q]; id[px[q]] = -ix[q]; }
  grp_sync(bar);
  return true;
}

// One TLS estimate over X[0..M) with a constant range r.  Returns the estimate in every thread of the group.
__device__ inline double tls_estimate_const_range(const double* X, int M, double r, double* val, int* id, double* sm /*3*kRtThreads*/, int tid, int bar, bool fast,
                                                  TailMarks tm = TailMarks()) {
  const int E = 2 * M;
  int Epad = fast ? kRtThreads : 2;   // fast: val / id are shared memory with room for kTrEndpoints entries
  while (Epad < E) Epad <<= 1;
  for (int e = tid; e < Epad; e += kRtThreads) {
    if (e < E) {
      const int i = e >> 1;
      if (e & 1) { val[e] = X[i] + r; id[e] = -(i + 1); }
      else { val[e] = X[i] - r; id[e] = i + 1; }
    } else { val[e] = CUDART_INF; id[e] = 0x7fffffff; }
  }
  tm.mark(21);   // endpoints written
  bool sorted_fast = false;
  if (fast && M <= 4 * kRtThreads) sorted_fast = sort_endpoints_const_range(X, M, r, val, id, sm, tid, bar);
  if (!sorted_fast && fast && M <= 4 * kRtThreads) {   // the shortcut declined (it used val / id as scratch): write the endpoints again
    for (int e = tid; e < Epad; e += kRtThreads) {
      if (e < E) {
        const int i = e >> 1;
        if (e & 1) { val[e] = X[i] + r; id[e] = -(i + 1); }
        else { val[e] = X[i] - r; id[e] = i + 1; }
      } else { val[e] = CUDART_INF; id[e] = 0x7fffffff; }
    }
  }
  if (sorted_fast) { /* id[] holds the 2 M endpoint ids in the reference's order */ }
  else if (fast && Epad == kRtThreads) bitonic_sort_regs<1>(val, id, tid, bar);
  else if (fast && Epad == 2 * kRtThreads) bitonic_sort_regs<2>(val, id, tid, bar);
  else if (fast && Epad == 4 * kRtThreads) bitonic_sort_regs<4>(val, id, tid, bar);
  else if (fast && Epad == 8 * kRtThreads) bitonic_sort_regs<8>(val, id, tid, bar);
  else bitonic_sort(val, id, Epad, tid, bar);
  tm.mark(22);   // sorted
  const double w = 1.0 / (r * r);
  const int chunk = (E + kRtThreads - 1) / kRtThreads;
  const int e0 = min(tid * chunk, E), e1 = min(e0 + chunk, E);
  double pre[6] = {0, 0, 0, 0, 0, 0};   // card, w, xw, r, x, xx
  for (int e = e0; e < e1; ++e) {
    const int idx = abs(id[e]) - 1;
    const double eps = id[e] > 0 ? 1.0 : -1.0;
    const double x = X[idx];
    pre[0] += eps; pre[1] += eps * w; pre[2] += eps * w * x; pre[3] += eps * r; pre[4] += eps * x; pre[5] += eps * x * x;
  }
  grp_exclusive_scan<6>(pre, sm, tid, bar);
  tm.mark(23);   // prefix sums
  double card = pre[0], dw = pre[1], dxw = pre[2];
  double rsum = (double)M * r - pre[3];
  double sx = pre[4], sxx = pre[5];
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
  grp_sync(bar);
  tm.mark(24);   // costs evaluated
  sm[tid] = best_cost; sm[kRtThreads + tid] = best_x; sm[2 * kRtThreads + tid] = (double)best_e;
  grp_sync(bar);
  for (int s2 = kRtThreads / 2; s2 > 0; s2 >>= 1) {
    if (tid < s2) {
      const double ca = sm[tid], cb = sm[tid + s2];
      const double ea = sm[2 * kRtThreads + tid], eb = sm[2 * kRtThreads + tid + s2];
      if (cb < ca || (cb == ca && eb < ea)) { sm[tid] = cb; sm[kRtThreads + tid] = sm[kRtThreads + tid + s2]; sm[2 * kRtThreads + tid] = eb; }
    }
    grp_sync(bar);
  }
  const double est = (sm[2 * kRtThreads] < 2.0e9) ? sm[kRtThreads] : CUDART_NAN;
  grp_sync(bar);
  return est;
}

constexpr int kTrEndpoints = 2048;                                        // endpoints per axis that fit in shared memory
constexpr size_t kTrGroupDoubles = 3 * kRtThreads + kTrEndpoints + kTrEndpoints / 2;   // scratch + val + id (ints)

// The translation stage of one registration on a CTA of exactly 3 * kRtThreads threads; shd: 3 * kTrGroupDoubles doubles of shared
// memory (per axis group: [3*kRtThreads] scratch, [2048] endpoint values, [2048] endpoint ids).
__device__ inline void translation_body(const ProbDesc& d, const KernelParams& kp, double* shd, TailMarks tm = TailMarks()) {
  ProbState* st = d.st;
  const int gtid = threadIdx.x;
  const int grp = gtid / kRtThreads, tid = gtid % kRtThreads;
  const int C = st->clique_size;
  if (C <= 1) return;
  const int M = st->n_rot_inl;
  const double scale = st->scale;
  double R[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = st->R[i];
  double* raw = d.raw;  // 3 x M (row r at raw + r*M)
  for (int i = gtid; i < M; i += 3 * kRtThreads) {
    const int pt = d.clique[d.rot_inl[i] % C];  // upstream indexes max_clique_ by TIM position (A.8)
    const double x0 = d.src[3 * (size_t)pt], x1 = d.src[3 * (size_t)pt + 1], x2 = d.src[3 * (size_t)pt + 2];
#pragma unroll
    for (int r = 0; r < 3; ++r)
      raw[(size_t)r * M + i] = d.dst[3 * (size_t)pt + r] - scale * ((R[3 * r] * x0 + R[3 * r + 1] * x1) + R[3 * r + 2] * x2);
    d.trans_mask[i] = 1;
  }
  __shared__ double s_t[3];
  __shared__ int sh_c;
  if (gtid == 0) sh_c = 0;
  if (gtid < 3) s_t[gtid] = 0.0;   // M == 0 (every GNC weight below 0.5, or the rotation stage bailed out): t = 0, valid = 0
  __syncthreads();
  tm.mark(20);   // raw translations written
  if (M >= 2) {
    const double r = kp.noise_bound * sqrt(kp.cbar2);  // A.7-4: 1x, not 2x
    int Epad = 2;
    while (Epad < 2 * M) Epad <<= 1;
    double* sm = shd + (size_t)grp * kTrGroupDoubles;
    double* val = sm + 3 * kRtThreads;
    int* id = (int*)(val + kTrEndpoints);
    if (Epad <= kTrEndpoints) {
      const double est = tls_estimate_const_range(raw + (size_t)grp * M, M, r, val, id, sm, tid, 1 + grp, true, tm);
      if (tid == 0) s_t[grp] = est;
    } else if (grp == 0) {  // large cliques: endpoints in the global scratch (4 * tim_cap entries), one axis after the other
      for (int ax = 0; ax < 3; ++ax) {
        const double est = tls_estimate_const_range(raw + (size_t)ax * M, M, r, d.tls_val, d.tls_id, sm, tid, 1, false);
        if (tid == 0) s_t[ax] = est;
      }
    }
    __syncthreads();
    tm.mark(25);   // three estimates
    for (int i = gtid; i < M; i += 3 * kRtThreads) {
      const bool in = (fabs(raw[i] - s_t[0]) <= r) && (fabs(raw[(size_t)M + i] - s_t[1]) <= r) && (fabs(raw[2 * (size_t)M + i] - s_t[2]) <= r);
      if (!in) d.trans_mask[i] = 0;
    }
  } else if (M == 1) {
    if (gtid < 3) s_t[gtid] = raw[gtid];
  }
  __syncthreads();
  // count translation inliers
  int c = 0;
  for (int i = gtid; i < M; i += 3 * kRtThreads) c += d.trans_mask[i];
  c = __reduce_add_sync(FULL, c);
  if ((gtid & 31) == 0) atomicAdd(&sh_c, c);
  __syncthreads();
  if (gtid == 0) {
    st->t[0] = s_t[0]; st->t[1] = s_t[1]; st->t[2] = s_t[2];
    st->n_trans_inl = sh_c;
    st->valid = M >= 1 ? 1 : 0;
  }
}

}  // namespace b2r
