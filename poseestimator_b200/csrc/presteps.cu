// Pre-steps of the registration path (SURVEY.md 8a-1a .. 8a-1d): everything the reference does with Open3D between
// the detector's mask and `solver.solve`:
//   back-projection        RealSenseClass.py:56-67, registration_utils.py:54-90, PoseEstimator.py:130-135   (B.1)
//   statistical outliers   RealSenseClass.py:68, registration_utils.py:93                                   (B.2)
//   voxel_down_sample      PoseEstimator.py:87,93                                                           (B.3)
//   estimate_normals       HelpersRealtime.py:215-219                                                       (B.4)
//   compute_fpfh_feature   HelpersRealtime.py:221-224                                                       (B.5)
// (paths under /root/reference/src/pose_estimator/EstimHelpers/).
//
// All arithmetic is IEEE fp64 in the reference's operation order; this translation unit is compiled with -fmad=false so
// no multiply-add is ever contracted.  Results that are order dependent (point order after compaction, voxel order,
// neighbour order, sequential sums) follow the canonical orders of the oracle (oracle/presteps_oracle.cpp header).
// The transcendental functions (acos, cos, atan2) come from include/b2r_exactmath.h, the double-double implementation the
// oracle includes verbatim, so histogram bins and eigen-solver outputs are identical bits on both sides.  Where only a DECISION
// depends on the value (the acos comparison and the atan2 bin of the SPFH pair features) CUDA's libm decides first and the exact
// function is evaluated only inside a guard band around the decision boundary -- the result is that of the exact function always.
//
// Kernels are HBM/L2-bound byte and index work: coalesced and vectorised loads (8 depth pixels = one 16-byte load),
// shared-memory staging of the point tiles, ordered stream compaction by a single-pass decoupled look-back scan.
#include <math_constants.h>

#include "scan.cuh"
#include "solver.cuh"
#include "../../include/b2r_exactmath.h"

namespace b2r {

namespace {

// ------------------------------------------------------------------------------------------------------------------
// a-1a  depth image -> point cloud (B.1), fused: scale, mask, z-range, truncation, pinhole back-projection and
// row-major ordered compaction.  One thread = one aligned group of 8 pixels (16-byte depth load, 8-byte mask load).
// ------------------------------------------------------------------------------------------------------------------
struct BackprojArgs {
  const uint16_t* depth; const uint8_t* mask;
  int H, W, u0, v0, u1, v1;
  double fx, fy, cx, cy, zmin, zmax;
  float scale, trunc;
  int use_range, use_trunc, vec_ok;
  double* pts; int* pix; int cap;
  unsigned long long* ws; int* count;
  unsigned int* tile_counts;            // two-pass form: points per tile (MODE 1 writes, tile_scan_kernel scans in place)
};

// MODE 0: single pass, ordered compaction by decoupled look-back (small images: one launch).
// MODE 1 / 2: two passes for large images -- count the valid pixels of every tile, scan the 8 k tile counts in one small kernel,
// then recompute and write with the tile's base known up front: no inter-CTA dependency chain at all, at the price of reading the
// 3 bytes per pixel twice (the output is 24 bytes per valid point).
// Output: every warp stages its <= 256 points in shared memory and writes them as contiguous doubles (the per-lane 8-byte stores this
// replaces touched 32 different sectors per instruction).
constexpr int kBpSmemBytes = 8 * (256 * 24 + 256 * 4);   // per warp: 256 points + 256 pixel ids

template <int MODE>
__global__ void __launch_bounds__(256) backproject_kernel(BackprojArgs a) {
  extern __shared__ __align__(16) unsigned char bp_smem[];
  __shared__ int s_tile;
  __shared__ unsigned s_warp[8];
  __shared__ unsigned long long s_base;
  const int tile = MODE == 0 ? take_ticket(a.ws, &s_tile) : (int)blockIdx.x;
  const int g0 = a.u0 >> 3;
  const int G = ((a.u1 + 7) >> 3) - g0;
  const long long total = (long long)(a.v1 - a.v0) * G;
  const long long g = (long long)tile * 256 + threadIdx.x;
  unsigned bits = 0;
  float dv[8];
  int row = 0, gu = 0;
  if (g < total) {
    row = a.v0 + (int)(g / G);
    gu = (g0 + (int)(g % G)) << 3;
    unsigned short dep[8];
    unsigned char mk[8];
    const size_t off = (size_t)row * a.W + gu;
    if (a.vec_ok) {
      const uint4 q = *reinterpret_cast<const uint4*>(a.depth + off);
      dep[0] = q.x & 0xffff; dep[1] = q.x >> 16; dep[2] = q.y & 0xffff; dep[3] = q.y >> 16;
      dep[4] = q.z & 0xffff; dep[5] = q.z >> 16; dep[6] = q.w & 0xffff; dep[7] = q.w >> 16;
      if (a.mask) {
        const uint2 m = *reinterpret_cast<const uint2*>(a.mask + off);
#pragma unroll
        for (int k = 0; k < 4; ++k) { mk[k] = (m.x >> (8 * k)) & 0xff; mk[4 + k] = (m.y >> (8 * k)) & 0xff; }
      }
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const bool in = gu + k < a.W;
        dep[k] = in ? a.depth[off + k] : 0;
        mk[k] = (in && a.mask) ? a.mask[off + k] : 0;
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int u = gu + k;
      float d = __fmul_rn((float)dep[k], a.scale);              // numpy float32 * float32(depth_scale)
      if (a.mask && mk[k] == 0) d = 0.f;
      if (a.use_range) { const double dd = (double)d; if (!(dd > a.zmin && dd < a.zmax)) d = 0.f; }
      if (a.use_trunc && d >= a.trunc) d = 0.f;
      dv[k] = d;
      if (u >= a.u0 && u < a.u1 && d > 0.f) bits |= 1u << k;
    }
  }
  unsigned blk_total;
  const unsigned local = block_scan_256(__popc(bits), s_warp, &blk_total);
  if (MODE == 1) {
    if (threadIdx.x == 0) a.tile_counts[tile] = blk_total;
    return;
  }
  if (threadIdx.x == 0) {
    if (MODE == 0) {
      s_base = lookback(a.ws, tile, blk_total);
      if (tile == (int)gridDim.x - 1) *a.count = (int)(s_base + blk_total);
    } else {
      s_base = a.tile_counts[tile];     // exclusive prefix after tile_scan_kernel
    }
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double* sp = reinterpret_cast<double*>(bp_smem + (size_t)warp * (256 * 24));
  int* sx = reinterpret_cast<int*>(bp_smem + 8 * (256 * 24) + (size_t)warp * (256 * 4));
  const unsigned wfirst = __shfl_sync(FULL, local, 0);
  const unsigned wtotal = __shfl_sync(FULL, local + (unsigned)__popc(bits), 31) - wfirst;
  unsigned c = local - wfirst;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    if (bits & (1u << k)) {
      const int u = gu + k;
      const double z = (double)dv[k];
      sp[3 * c] = __ddiv_rn(__dmul_rn((double)u - a.cx, z), a.fx);
      sp[3 * c + 1] = __ddiv_rn(__dmul_rn((double)row - a.cy, z), a.fy);
      sp[3 * c + 2] = z;
      sx[c] = row * a.W + u;
      ++c;
    }
  }
  __syncwarp();
  const long long gbase = (long long)s_base + wfirst;
  long long room = (long long)a.cap - gbase;          // points of this warp that still fit the output
  if (room > (long long)wtotal) room = wtotal;
  for (long long i = lane; i < 3 * room; i += 32) a.pts[3 * gbase + i] = sp[i];
  if (a.pix) for (long long i = lane; i < room; i += 32) a.pix[gbase + i] = sx[i];
}

// exclusive scan of the tile counts in place (one CTA), total to *count
__global__ void __launch_bounds__(1024) tile_scan_kernel(unsigned int* __restrict__ counts, int tiles, int* __restrict__ count) {
  __shared__ unsigned int wsum[32];
  __shared__ unsigned int carry_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int i0 = 0; i0 < tiles; i0 += 1024) {
    const int i = i0 + threadIdx.x;
    const unsigned v = i < tiles ? counts[i] : 0u;
    unsigned x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(FULL, x, o); if (lane >= o) x += y; }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    unsigned wb = 0;
    for (int w = 0; w < warp; ++w) wb += wsum[w];
    const unsigned carry = carry_s;
    if (i < tiles) counts[i] = carry + wb + x - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry_s = carry + wb + x;
    __syncthreads();
  }
  if (threadIdx.x == 0) *count = (int)carry_s;
}

constexpr int kCompactTile = 1024;
// ordered compaction of 3-vectors by a keep mask (four points per thread); idx_out optional
__global__ void __launch_bounds__(256) compact_points_kernel(const double* __restrict__ pts, const unsigned char* __restrict__ keep,
                                                             int n, double* __restrict__ out, int* __restrict__ idx_out,
                                                             unsigned long long* ws, int* count) {
  __shared__ int s_tile;
  __shared__ unsigned s_warp[8];
  __shared__ unsigned long long s_base;
  // tiles of 1024 points, four consecutive points per thread: the look-back is a serial chain over the tiles (~50 ns a link), a
  // 56 k-point frame has 56 links instead of 221
  const int tile = take_ticket(ws, &s_tile);
  const int i0 = tile * kCompactTile + 4 * threadIdx.x;
  unsigned kk[4], k = 0;
#pragma unroll
  for (int u = 0; u < 4; ++u) { kk[u] = (i0 + u < n && keep[i0 + u]) ? 1u : 0u; k += kk[u]; }
  unsigned blk_total;
  const unsigned local = block_scan_256(k, s_warp, &blk_total);
  if (threadIdx.x == 0) {
    s_base = lookback(ws, tile, blk_total);
    if (tile == (int)gridDim.x - 1) *count = (int)(s_base + blk_total);
  }
  __syncthreads();
  size_t m = (size_t)s_base + local;
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    if (kk[u]) {
      const int i = i0 + u;
      out[3 * m] = pts[3 * (size_t)i]; out[3 * m + 1] = pts[3 * (size_t)i + 1]; out[3 * m + 2] = pts[3 * (size_t)i + 2];
      if (idx_out) idx_out[m] = i;
      ++m;
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------
// kNN / hybrid search (B.4: the max_nn nearest points, the query itself included, then d2 < radius^2), brute force,
// one warp per query, neighbours in ascending (d2, index).  The cloud streams through shared memory in tiles shared
// by the 8 warps of a CTA; a warp keeps its best list sorted in shared memory and merges qualifying candidates in
// batches (bitonic sort of the batch + merge path), so the common case per 32 candidates is: 32 distance evaluations,
// one compare against the current K-th key, one ballot.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kKnnWarps = 8;
constexpr int kKnnTile = 512;
constexpr int kKnnMaxK = 128;
constexpr int kKnnBuf = 64;
constexpr int kKnnList = kKnnMaxK + kKnnBuf;  // 192

struct KnnWarpSmem {
  unsigned long long lk[2][kKnnList];
  int li[2][kKnnList];
  unsigned long long ck[kKnnBuf];
  int ci[kKnnBuf];
};

__device__ __forceinline__ bool key_less(unsigned long long k1, int i1, unsigned long long k2, int i2) {
  return k1 < k2 || (k1 == k2 && i1 < i2);
}

__device__ __forceinline__ double sq3_ref(double dx, double dy, double dz) {
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

// One compare-exchange step of a bitonic network held in registers: element e = lane + 32 h of the 64-slot batch is register h of
// `lane`; the partner e ^ j (j < 32) is the same register of lane ^ j.  keep_min: this side keeps the smaller of the two.
__device__ __forceinline__ void knn_cx(unsigned long long& k, int& i, int j, bool keep_min) {
  const unsigned long long ok = __shfl_xor_sync(0xffffffffu, k, j);
  const int oi = __shfl_xor_sync(0xffffffffu, i, j);
  const bool other_less = key_less(ok, oi, k, i);
  if (other_less == keep_min) { k = ok; i = oi; }   // (equal entries are padding: either copy will do)
}

// ascending sort of the batch w.ck / w.ci [0, nb), padded to 32 or 64 slots with (~0, int max): a bitonic network on two registers per
// lane and shuffles.  (Round 2: the shared-memory network it replaces -- 21 stages, two barriers-separated passes each with half the
// lanes idle -- was 40 % of knn_grid_kernel's instructions on a config-2 frame.)  (key, index) is a strict total order, so the sorted
// batch is the same whatever network produces it.
__device__ __forceinline__ void knn_sort_batch(KnnWarpSmem& w, int nb, int lane) {
  unsigned long long k0 = lane < nb ? w.ck[lane] : ~0ull;
  int i0 = lane < nb ? w.ci[lane] : 0x7fffffff;
  if (nb <= 32) {
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
      const bool up = k == 32 || (lane & k) == 0;
#pragma unroll
      for (int j = k >> 1; j > 0; j >>= 1) knn_cx(k0, i0, j, ((lane & j) == 0) == up);
    }
    __syncwarp();
    w.ck[lane] = k0; w.ci[lane] = i0;
  } else {
    unsigned long long k1 = lane + 32 < nb ? w.ck[lane + 32] : ~0ull;
    int i1 = lane + 32 < nb ? w.ci[lane + 32] : 0x7fffffff;
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
      // direction of element e = lane + 32 h in stage k: ascending when (e & k) == 0; for k == 32 that is h == 0
      const bool up0 = k == 32 || (lane & k) == 0;
      const bool up1 = k == 32 ? false : (lane & k) == 0;
#pragma unroll
      for (int j = k >> 1; j > 0; j >>= 1) {
        knn_cx(k0, i0, j, ((lane & j) == 0) == up0);
        knn_cx(k1, i1, j, ((lane & j) == 0) == up1);
      }
    }
    // k == 64: ascending everywhere; j == 32 pairs the two registers of a lane
    if (key_less(k1, i1, k0, i0)) { const unsigned long long tk = k0; k0 = k1; k1 = tk; const int t = i0; i0 = i1; i1 = t; }
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
      knn_cx(k0, i0, j, (lane & j) == 0);
      knn_cx(k1, i1, j, (lane & j) == 0);
    }
    __syncwarp();
    w.ck[lane] = k0; w.ci[lane] = i0;
    w.ck[lane + 32] = k1; w.ci[lane + 32] = i1;
  }
  __syncwarp();
}

// merge the (unsorted) candidate batch of size nb into the sorted list of size a; returns the new size (<= K)
__device__ int knn_flush(KnnWarpSmem& w, int& cur, int a, int nb, int K, int lane) {
  knn_sort_batch(w, nb, lane);
  const unsigned long long* A = w.lk[cur]; const int* Ai = w.li[cur];
  unsigned long long* O = w.lk[cur ^ 1]; int* Oi = w.li[cur ^ 1];
  for (int e = lane; e < a; e += 32) {   // rank of A[e] = e + |{B < A[e]}|
    const unsigned long long k = A[e]; const int id = Ai[e];
    int lo = 0, hi = nb;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (key_less(w.ck[mid], w.ci[mid], k, id)) lo = mid + 1; else hi = mid; }
    const int r = e + lo;
    if (r < K) { O[r] = k; Oi[r] = id; }
  }
  for (int e = lane; e < nb; e += 32) {  // rank of B[e] = e + |{A < B[e]}|
    const unsigned long long k = w.ck[e]; const int id = w.ci[e];
    int lo = 0, hi = a;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (key_less(A[mid], Ai[mid], k, id)) lo = mid + 1; else hi = mid; }
    const int r = e + lo;
    if (r < K) { O[r] = k; Oi[r] = id; }
  }
  __syncwarp();
  cur ^= 1;
  return min(a + nb, K);
}

__global__ void __launch_bounds__(kKnnWarps * 32) knn_kernel(const double* __restrict__ pts, int n, const double* __restrict__ queries,
                                                             int nq, int K, double radius, int* __restrict__ idx_out,
                                                             double* __restrict__ d2_out, int* __restrict__ cnt_out) {
  extern __shared__ __align__(16) unsigned char knn_smem[];
  double* tile = reinterpret_cast<double*>(knn_smem);                                   // kKnnTile x 3
  KnnWarpSmem* ws = reinterpret_cast<KnnWarpSmem*>(knn_smem + sizeof(double) * 3 * kKnnTile);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  KnnWarpSmem& w = ws[warp];
  const int q = blockIdx.x * kKnnWarps + warp;
  const bool active = q < nq;
  double qx = 0, qy = 0, qz = 0;
  if (active) { qx = queries[3 * (size_t)q]; qy = queries[3 * (size_t)q + 1]; qz = queries[3 * (size_t)q + 2]; }
  const bool use_r = radius > 0;
  const double r2 = __dmul_rn(radius, radius);
  int cur = 0, a = 0, nb = 0;
  unsigned long long tk = ~0ull; int ti = 0x7fffffff;  // key of the K-th best once the list is full
  for (int base = 0; base < n; base += kKnnTile) {
    __syncthreads();
    const int m = min(kKnnTile, n - base);
    for (int e = threadIdx.x; e < 3 * m; e += kKnnWarps * 32) tile[e] = pts[3 * (size_t)base + e];
    __syncthreads();
    if (!active) continue;
    for (int s = 0; s < m; s += 32) {
      const int jl = s + lane;
      bool qual = false;
      unsigned long long key = 0;
      if (jl < m) {
        const double d2 = sq3_ref(__dsub_rn(tile[3 * jl], qx), __dsub_rn(tile[3 * jl + 1], qy), __dsub_rn(tile[3 * jl + 2], qz));
        key = (unsigned long long)__double_as_longlong(d2);
        qual = (!use_r || d2 < r2) && key_less(key, base + jl, tk, ti);
      }
      const unsigned bal = __ballot_sync(FULL, qual);
      if (bal) {
        if (qual) { const int p = nb + __popc(bal & ((1u << lane) - 1u)); w.ck[p] = key; w.ci[p] = base + jl; }
        nb += __popc(bal);
        __syncwarp();
        if (nb > 32) {
          a = knn_flush(w, cur, a, nb, K, lane);
          nb = 0;
          if (a == K) { tk = w.lk[cur][K - 1]; ti = w.li[cur][K - 1]; }
        }
      }
    }
  }
  if (!active) return;
  if (nb > 0) a = knn_flush(w, cur, a, nb, K, lane);
  for (int t = lane; t < K; t += 32) {
    idx_out[(size_t)q * K + t] = t < a ? w.li[cur][t] : -1;
    d2_out[(size_t)q * K + t] = t < a ? __longlong_as_double((long long)w.lk[cur][t]) : 0.0;
  }
  if (lane == 0) cnt_out[q] = a;
}

// ------------------------------------------------------------------------------------------------------------------
// a-1a (tail)  remove_statistical_outlier (B.2)
// ------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) avg_dist_kernel(const double* __restrict__ d2, const int* __restrict__ cnt, int n, int K,
                                                       double* __restrict__ avg) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const int c = cnt[i];
  double mean = -1.0;
  if (c > 0) {
    double s = 0;
    for (int t = 0; t < c; ++t) s = __dadd_rn(s, __dsqrt_rn(d2[(size_t)i * K + t]));  // sequential, ascending distance
    mean = __ddiv_rn(s, (double)c);
  }
  avg[i] = mean;
}

template <int NT>
__device__ double block_sum_tree(double v, double* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0;
#pragma unroll
  for (int w = 0; w < NT / 32; ++w) s += red[w];
  return s;
}

// single CTA.  The reference sums sequentially (std::accumulate / inner_product); a tree sum differs in the last
// bits, which can only matter for a point whose average distance is within a hair of the threshold.  The tree result
// is used when no point lies in that band (|avg - thr| <= 1e-8 |thr|, five orders above the summation error for
// n <= 1e6), otherwise one thread replays the reference's sequential sums -- either way the verdicts are the reference's.
__global__ void __launch_bounds__(1024) outlier_threshold_kernel(const double* __restrict__ avg, int n, double std_ratio,
                                                                 double* __restrict__ thr_out /*[0] thr [1] mean [2] std*/) {
  __shared__ double red[32];
  __shared__ int s_amb;
  const int tid = threadIdx.x;
  // (every pass: eight independent loads in flight per thread -- one CTA walking 56 k values one L2 latency at a time took 22 us)
  double s = 0; int valid_t = 0;
  for (int i0 = tid; i0 < n; i0 += 8 * 1024) {
    double v8[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) { const int i = i0 + 1024 * u; v8[u] = i < n ? avg[i] : -1.0; }
#pragma unroll
    for (int u = 0; u < 8; ++u) { if (v8[u] > 0) s += v8[u]; if (v8[u] >= 0) ++valid_t; }
  }
  const double sum = block_sum_tree<1024>(s, red);
  const double valid = block_sum_tree<1024>((double)valid_t, red);
  const double mean = sum / valid;
  double q = 0;
  for (int i0 = tid; i0 < n; i0 += 8 * 1024) {
    double v8[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) { const int i = i0 + 1024 * u; v8[u] = i < n ? avg[i] : -1.0; }
#pragma unroll
    for (int u = 0; u < 8; ++u) if (v8[u] > 0) q += (v8[u] - mean) * (v8[u] - mean);
  }
  const double sq = block_sum_tree<1024>(q, red);
  double stdv = sqrt(sq / (valid - 1.0));
  double thr = mean + std_ratio * stdv;
  if (tid == 0) s_amb = 0;
  __syncthreads();
  int amb = 0;
  const double band = 1e-8 * fabs(thr);
  for (int i0 = tid; i0 < n; i0 += 8 * 1024) {
    double v8[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) { const int i = i0 + 1024 * u; v8[u] = i < n ? avg[i] : -1.0; }
#pragma unroll
    for (int u = 0; u < 8; ++u) if (v8[u] > 0 && fabs(v8[u] - thr) <= band) amb = 1;
  }
  if (amb) s_amb = 1;
  __syncthreads();
  if (tid == 0) {
    double m2 = mean;
    if (s_amb || !(thr == thr)) {
      double ss = 0;
      for (int i = 0; i < n; ++i) if (avg[i] > 0) ss = __dadd_rn(ss, avg[i]);
      m2 = __ddiv_rn(ss, valid);
      double qq = 0;
      for (int i = 0; i < n; ++i) {
        const double v = avg[i];
        if (v > 0) { const double dlt = __dsub_rn(v, m2); qq = __dadd_rn(qq, __dmul_rn(dlt, dlt)); } else qq = __dadd_rn(qq, 0.0);
      }
      stdv = __dsqrt_rn(__ddiv_rn(qq, valid - 1.0));
      thr = __dadd_rn(m2, __dmul_rn(std_ratio, stdv));
    }
    thr_out[0] = thr; thr_out[1] = m2; thr_out[2] = stdv;
  }
}

__global__ void __launch_bounds__(256) outlier_keep_kernel(const double* __restrict__ avg, int n, const double* __restrict__ thr,
                                                           unsigned char* __restrict__ keep) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n) { const double v = avg[i]; keep[i] = (v > 0 && v < thr[0]) ? 1 : 0; }
}

// ------------------------------------------------------------------------------------------------------------------
// a-1b  voxel_down_sample (B.3)
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long ordered_key(double v) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__host__ __device__ inline double ordered_key_inv(unsigned long long k) {
  const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
  double d;
#ifdef __CUDA_ARCH__
  d = __longlong_as_double((long long)b);
#else
  memcpy(&d, &b, sizeof(d));
#endif
  return d;
}

// bounds[0..2] = min keys (init ~0), bounds[3..5] = max keys (init 0)
__global__ void __launch_bounds__(256) bounds_kernel(const double* __restrict__ pts, int n, unsigned long long* __restrict__ bounds) {
  __shared__ unsigned long long s[6];
  if (threadIdx.x < 3) s[threadIdx.x] = ~0ull; else if (threadIdx.x < 6) s[threadIdx.x] = 0ull;
  __syncthreads();
  unsigned long long mn[3] = {~0ull, ~0ull, ~0ull}, mx[3] = {0, 0, 0};
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256)
#pragma unroll
    for (int a = 0; a < 3; ++a) { const unsigned long long k = ordered_key(pts[3 * (size_t)i + a]); mn[a] = min(mn[a], k); mx[a] = max(mx[a], k); }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { mn[a] = min(mn[a], __shfl_xor_sync(FULL, mn[a], o)); mx[a] = max(mx[a], __shfl_xor_sync(FULL, mx[a], o)); }
    if ((threadIdx.x & 31) == 0) { atomicMin(&s[a], mn[a]); atomicMax(&s[3 + a], mx[a]); }
  }
  __syncthreads();
  if (threadIdx.x < 3) atomicMin(&bounds[threadIdx.x], s[threadIdx.x]);
  else if (threadIdx.x < 6) atomicMax(&bounds[threadIdx.x], s[threadIdx.x]);
}

struct VoxelGrid { double org[3]; double vs; long long dim0, dim1; };

__global__ void __launch_bounds__(256) voxel_key_kernel(const double* __restrict__ pts, int n, VoxelGrid g, unsigned long long* __restrict__ keys,
                                                        uint32_t* __restrict__ vals) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  long long c[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) c[a] = (long long)floor(__ddiv_rn(__dsub_rn(pts[3 * (size_t)i + a], g.org[a]), g.vs));
  keys[i] = (unsigned long long)(c[0] + g.dim0 * (c[1] + g.dim1 * c[2]));
  vals[i] = (uint32_t)i;
}

// segment heads of the sorted keys -> voxel ids; seg_start[v] = first sorted position of voxel v; *count = V
__global__ void __launch_bounds__(256) voxel_heads_kernel(const unsigned long long* __restrict__ keys, int n, int* __restrict__ seg_start,
                                                          unsigned long long* ws, int* count) {
  __shared__ int s_tile;
  __shared__ unsigned s_warp[8];
  __shared__ unsigned long long s_base;
  const int tile = take_ticket(ws, &s_tile);
  const int e0 = tile * kCompactTile + 4 * threadIdx.x;   // four consecutive keys per thread (see compact_points_kernel)
  unsigned hh[4], head = 0;
  unsigned long long prev = e0 > 0 && e0 - 1 < n ? keys[e0 - 1] : 0ull;
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int e = e0 + u;
    const unsigned long long k = e < n ? keys[e] : 0ull;
    hh[u] = (e < n && (e == 0 || k != prev)) ? 1u : 0u;
    head += hh[u];
    prev = k;
  }
  unsigned blk_total;
  const unsigned local = block_scan_256(head, s_warp, &blk_total);
  if (threadIdx.x == 0) {
    s_base = lookback(ws, tile, blk_total);
    if (tile == (int)gridDim.x - 1) *count = (int)(s_base + blk_total);
  }
  __syncthreads();
  unsigned long long m = s_base + local;
#pragma unroll
  for (int u = 0; u < 4; ++u) if (hh[u]) seg_start[m++] = e0 + u;
}

// one thread per voxel: mean of its points in ascending point index (the order the sequential insert produces)
__global__ void __launch_bounds__(128) voxel_mean_kernel(const double* __restrict__ pts, int n, const uint32_t* __restrict__ sorted_idx,
                                                         const int* __restrict__ seg_start, const int* __restrict__ count,
                                                         double* __restrict__ out, int* __restrict__ vox_of_point, int* __restrict__ counts) {
  const int v = blockIdx.x * 128 + threadIdx.x;
  const int V = *count;
  if (v >= V) return;
  const int b = seg_start[v], e = (v + 1 < V) ? seg_start[v + 1] : n;
  double ax = 0, ay = 0, az = 0;
  // eight points at a time: their indices, then their coordinates, are loaded together (two memory latencies per eight points instead
  // of two per point -- a voxel of a depth frame holds ~17 points and the kernel was a chain of dependent L2 round trips); the sums
  // stay sequential in ascending point index
  for (int s0 = b; s0 < e; s0 += 8) {
    uint32_t i8[8];
    double x8[8], y8[8], z8[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) i8[u] = sorted_idx[min(s0 + u, e - 1)];
#pragma unroll
    for (int u = 0; u < 8; ++u) { x8[u] = pts[3 * (size_t)i8[u]]; y8[u] = pts[3 * (size_t)i8[u] + 1]; z8[u] = pts[3 * (size_t)i8[u] + 2]; }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (s0 + u < e) {
        ax = __dadd_rn(ax, x8[u]); ay = __dadd_rn(ay, y8[u]); az = __dadd_rn(az, z8[u]);
        if (vox_of_point) vox_of_point[i8[u]] = v;
      }
    }
  }
  const double c = (double)(e - b);
  out[3 * (size_t)v] = __ddiv_rn(ax, c); out[3 * (size_t)v + 1] = __ddiv_rn(ay, c); out[3 * (size_t)v + 2] = __ddiv_rn(az, c);
  if (counts) counts[v] = e - b;
}

// ------------------------------------------------------------------------------------------------------------------
// a-1c  estimate_normals (B.4): covariance from cumulants in neighbour order + analytic symmetric 3x3 eigen-solver
// (the algorithm of Open3D's FastEigen3x3, after Eberly).  Operation order = oracle/presteps_oracle.cpp.
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void cross3(const double* a, const double* b, double* c) {
  c[0] = a[1] * b[2] - a[2] * b[1];
  c[1] = a[2] * b[0] - a[0] * b[2];
  c[2] = a[0] * b[1] - a[1] * b[0];
}
__device__ __forceinline__ double dot3(const double* a, const double* b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }

__device__ void eigvec0(const double A[9], double ev, double out[3]) {
  const double r0[3] = {A[0] - ev, A[1], A[2]};
  const double r1[3] = {A[1], A[4] - ev, A[5]};
  const double r2[3] = {A[2], A[5], A[8] - ev};
  double c01[3], c02[3], c12[3];
  cross3(r0, r1, c01); cross3(r0, r2, c02); cross3(r1, r2, c12);
  const double d0 = dot3(c01, c01), d1 = dot3(c02, c02), d2 = dot3(c12, c12);
  double dmax = d0; int imax = 0;
  if (d1 > dmax) { dmax = d1; imax = 1; }
  if (d2 > dmax) { imax = 2; }
  double c[3]; double d;
  if (imax == 0) { c[0] = c01[0]; c[1] = c01[1]; c[2] = c01[2]; d = d0; }
  else if (imax == 1) { c[0] = c02[0]; c[1] = c02[1]; c[2] = c02[2]; d = d1; }
  else { c[0] = c12[0]; c[1] = c12[1]; c[2] = c12[2]; d = d2; }
  const double s = sqrt(d);
  out[0] = c[0] / s; out[1] = c[1] / s; out[2] = c[2] / s;
}

__device__ void eigvec1(const double A[9], const double e0[3], double ev1, double out[3]) {
  double U[3], V[3];
  if (fabs(e0[0]) > fabs(e0[1])) {
    const double inv = 1.0 / sqrt(e0[0] * e0[0] + e0[2] * e0[2]);
    U[0] = -e0[2] * inv; U[1] = 0; U[2] = e0[0] * inv;
  } else {
    const double inv = 1.0 / sqrt(e0[1] * e0[1] + e0[2] * e0[2]);
    U[0] = 0; U[1] = e0[2] * inv; U[2] = -e0[1] * inv;
  }
  cross3(e0, U, V);
  const double AU[3] = {(A[0] * U[0] + A[1] * U[1]) + A[2] * U[2], (A[1] * U[0] + A[4] * U[1]) + A[5] * U[2],
                        (A[2] * U[0] + A[5] * U[1]) + A[8] * U[2]};
  const double AV[3] = {(A[0] * V[0] + A[1] * V[1]) + A[2] * V[2], (A[1] * V[0] + A[4] * V[1]) + A[5] * V[2],
                        (A[2] * V[0] + A[5] * V[1]) + A[8] * V[2]};
  double m00 = dot3(U, AU) - ev1, m01 = dot3(U, AV), m11 = dot3(V, AV) - ev1;
  const double a00 = fabs(m00), a01 = fabs(m01), a11 = fabs(m11);
  if (a00 >= a11) {
    const double mx = fmax(a00, a01);
    if (mx > 0) {
      if (a00 >= a01) { m01 /= m00; m00 = 1 / sqrt(1 + m01 * m01); m01 *= m00; }
      else { m00 /= m01; m01 = 1 / sqrt(1 + m00 * m00); m00 *= m01; }
#pragma unroll
      for (int i = 0; i < 3; ++i) out[i] = m01 * U[i] - m00 * V[i];
    } else { out[0] = U[0]; out[1] = U[1]; out[2] = U[2]; }
  } else {
    const double mx = fmax(a11, a01);
    if (mx > 0) {
      if (a11 >= a01) { m01 /= m11; m11 = 1 / sqrt(1 + m01 * m01); m01 *= m11; }
      else { m11 /= m01; m01 = 1 / sqrt(1 + m11 * m11); m11 *= m01; }
#pragma unroll
      for (int i = 0; i < 3; ++i) out[i] = m11 * U[i] - m01 * V[i];
    } else { out[0] = U[0]; out[1] = U[1]; out[2] = U[2]; }
  }
}

__device__ void smallest_eigvec(const double C[9], double out[3]) {
  double A[9];
  double mx = C[0];
#pragma unroll
  for (int i = 1; i < 9; ++i) mx = fmax(mx, C[i]);
  if (mx == 0) { out[0] = out[1] = out[2] = 0; return; }
#pragma unroll
  for (int i = 0; i < 9; ++i) A[i] = C[i] / mx;
  const double norm = (A[1] * A[1] + A[2] * A[2]) + A[5] * A[5];
  if (norm > 0) {
    const double q = ((A[0] + A[4]) + A[8]) / 3.0;
    const double b00 = A[0] - q, b11 = A[4] - q, b22 = A[8] - q;
    const double p = sqrt((((b00 * b00 + b11 * b11) + b22 * b22) + norm * 2.0) / 6.0);
    const double c00 = b11 * b22 - A[5] * A[5];
    const double c01 = A[1] * b22 - A[5] * A[2];
    const double c02 = A[1] * A[5] - b11 * A[2];
    const double det = ((b00 * c00 - A[1] * c01) + A[2] * c02) / ((p * p) * p);
    double half = det * 0.5;
    half = fmin(fmax(half, -1.0), 1.0);
    const double angle = b2r_acos(half) / 3.0;
    const double two_thirds_pi = 2.09439510239319549;
    const double beta2 = b2r_cos(angle) * 2.0;
    const double beta0 = b2r_cos(angle + two_thirds_pi) * 2.0;
    const double beta1 = -(beta0 + beta2);
    const double e0 = q + p * beta0, e1 = q + p * beta1, e2 = q + p * beta2;
    double v0[3], v1[3], v2[3];
    if (half >= 0) {
      eigvec0(A, e2, v2);
      if (e2 < e0 && e2 < e1) { out[0] = v2[0]; out[1] = v2[1]; out[2] = v2[2]; return; }
      eigvec1(A, v2, e1, v1);
      if (e1 < e0 && e1 < e2) { out[0] = v1[0]; out[1] = v1[1]; out[2] = v1[2]; return; }
      cross3(v1, v2, out);
    } else {
      eigvec0(A, e0, v0);
      if (e0 < e1 && e0 < e2) { out[0] = v0[0]; out[1] = v0[1]; out[2] = v0[2]; return; }
      eigvec1(A, v0, e1, v1);
      if (e1 < e0 && e1 < e2) { out[0] = v1[0]; out[1] = v1[1]; out[2] = v1[2]; return; }
      cross3(v0, v1, out);
    }
  } else {
    if (C[0] < C[4] && C[0] < C[8]) { out[0] = 1; out[1] = 0; out[2] = 0; }
    else if (C[4] < C[0] && C[4] < C[8]) { out[0] = 0; out[1] = 1; out[2] = 0; }
    else { out[0] = 0; out[1] = 0; out[2] = 1; }
  }
}

__global__ void __launch_bounds__(128) normals_kernel(const double* __restrict__ pts, int n, const int* __restrict__ idx,
                                                      const int* __restrict__ cnt, int K, int orient, double vx, double vy, double vz,
                                                      double* __restrict__ normals) {
  const int i = blockIdx.x * 128 + threadIdx.x;
  if (i >= n) return;
  double nrm[3] = {0, 0, 1};
  const int c_n = cnt[i];
  if (c_n >= 3) {
    double c[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int t = 0; t < c_n; ++t) {  // cumulants in neighbour order
      const double* p = pts + 3 * (size_t)idx[(size_t)i * K + t];
      const double x = p[0], y = p[1], z = p[2];
      c[0] += x; c[1] += y; c[2] += z;
      c[3] += x * x; c[4] += x * y; c[5] += x * z;
      c[6] += y * y; c[7] += y * z; c[8] += z * z;
    }
    const double m = (double)c_n;
#pragma unroll
    for (int j = 0; j < 9; ++j) c[j] /= m;
    double Cm[9];
    Cm[0] = c[3] - c[0] * c[0]; Cm[4] = c[6] - c[1] * c[1]; Cm[8] = c[8] - c[2] * c[2];
    Cm[1] = Cm[3] = c[4] - c[0] * c[1];
    Cm[2] = Cm[6] = c[5] - c[0] * c[2];
    Cm[5] = Cm[7] = c[7] - c[1] * c[2];
    smallest_eigvec(Cm, nrm);
    const double len = sqrt((nrm[0] * nrm[0] + nrm[1] * nrm[1]) + nrm[2] * nrm[2]);
    if (len == 0.0 || !(len == len)) { nrm[0] = 0; nrm[1] = 0; nrm[2] = 1; }
  }
  if (orient) {
    const double dx = vx - pts[3 * (size_t)i], dy = vy - pts[3 * (size_t)i + 1], dz = vz - pts[3 * (size_t)i + 2];
    double dp = (nrm[0] * dx + nrm[1] * dy) + nrm[2] * dz;
    if (orient == 2) dp = -dp;
    if (dp < 0) { nrm[0] = -nrm[0]; nrm[1] = -nrm[1]; nrm[2] = -nrm[2]; }
  }
  normals[3 * (size_t)i] = nrm[0]; normals[3 * (size_t)i + 1] = nrm[1]; normals[3 * (size_t)i + 2] = nrm[2];
}

// ------------------------------------------------------------------------------------------------------------------
// a-1d  compute_fpfh_feature (B.5).  One warp per point.
//   SPFH: lanes evaluate the pair features of the neighbours, histogram COUNTS go to shared memory; a bin's value is
//         count copies of the increment added one after the other, exactly what the reference's `+= incr` loop yields.
//   FPFH: lane b accumulates bin b over the neighbours in order; the three group sums are accumulated in the
//         reference's (neighbour, bin) order by the lanes 0..2 through shuffles.
// spfh_t / fpfh internal layout: [n][33] (one 264-byte row per point, coalesced gathers); the output is Open3D's
// Feature.data layout [33][n].
// ------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int bin11(double x) {
  const int h = (int)floor(x);
  return h < 0 ? 0 : (h > 10 ? 10 : h);
}

__device__ void pair_features(const double* p1, const double* n1, const double* p2, const double* n2, double f[4]) {
  double dp[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  f[0] = f[1] = f[2] = 0;
  f[3] = sqrt((dp[0] * dp[0] + dp[1] * dp[1]) + dp[2] * dp[2]);
  if (f[3] == 0.0) return;
  const double* a = n1;
  const double* b = n2;
  const double angle1 = dot3(n1, dp) / f[3];
  const double angle2 = dot3(n2, dp) / f[3];
  // acos(|angle1|) > acos(|angle2|) on the exact (correctly rounded) values.  CUDA's acos is within 2 ulp of them (<= 4.5e-16 here,
  // values in [0, pi/2]): when the two libm values differ by more than 4e-15 the exact ones are ordered the same way.
  const double q1 = acos(fabs(angle1)), q2 = acos(fabs(angle2));
  bool swap_roles = q1 > q2;
  if (fabs(q1 - q2) <= 4e-15) swap_roles = b2r_acos(fabs(angle1)) > b2r_acos(fabs(angle2));
  if (swap_roles) {
    a = n2; b = n1;
    dp[0] = -dp[0]; dp[1] = -dp[1]; dp[2] = -dp[2];
    f[2] = -angle2;
  } else {
    f[2] = angle1;
  }
  double v[3];
  cross3(dp, a, v);
  const double vn = sqrt((v[0] * v[0] + v[1] * v[1]) + v[2] * v[2]);
  if (vn == 0.0) { f[0] = f[1] = f[2] = f[3] = 0; return; }
  v[0] /= vn; v[1] /= vn; v[2] /= vn;
  double w[3];
  cross3(a, v, w);
  f[1] = dot3(v, b);
  // Only floor(11 (f0 + pi) / (2 pi)) is used.  CUDA's atan2 is within 2 ulp (<= 9e-16) of the exact value, which moves the bin
  // coordinate by < 2e-15 plus a few roundings of the same size; inside 1e-13 of a bin edge the exact function decides.
  const double wy = dot3(w, b), wx = dot3(a, b);
  double f0 = atan2(wy, wx);
  const double kPi_ = 3.14159265358979323846;
  const double xb = 11 * (f0 + kPi_) / (2.0 * kPi_);
  const double fr = xb - floor(xb);
  if (!(fr > 1e-13 && fr < 1.0 - 1e-13)) f0 = b2r_atan2(wy, wx);
  f[0] = f0;
}

constexpr int kFpWarps = 4;

__global__ void __launch_bounds__(kFpWarps * 32) spfh_kernel(const double* __restrict__ pts, const double* __restrict__ nrm, int n,
                                                             const int* __restrict__ idx, const int* __restrict__ cnt, int K,
                                                             double* __restrict__ spfh_t /*[n][33]*/) {
  __shared__ int hist[kFpWarps][33];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = blockIdx.x * kFpWarps + warp;
  if (i >= n) return;
  hist[warp][lane] = 0;
  if (lane == 0) hist[warp][32] = 0;
  __syncwarp();
  const int c = cnt[i];
  const double kPi = 3.14159265358979323846;
  if (c > 1) {
    const double pi_[3] = {pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2]};
    const double ni[3] = {nrm[3 * (size_t)i], nrm[3 * (size_t)i + 1], nrm[3 * (size_t)i + 2]};
    for (int t = 1 + lane; t < c; t += 32) {
      const int j = idx[(size_t)i * K + t];
      const double pj[3] = {pts[3 * (size_t)j], pts[3 * (size_t)j + 1], pts[3 * (size_t)j + 2]};
      const double nj[3] = {nrm[3 * (size_t)j], nrm[3 * (size_t)j + 1], nrm[3 * (size_t)j + 2]};
      double f[4];
      pair_features(pi_, ni, pj, nj, f);
      atomicAdd(&hist[warp][bin11(11 * (f[0] + kPi) / (2.0 * kPi))], 1);
      atomicAdd(&hist[warp][bin11(11 * (f[1] + 1.0) * 0.5) + 11], 1);
      atomicAdd(&hist[warp][bin11(11 * (f[2] + 1.0) * 0.5) + 22], 1);
    }
  }
  __syncwarp();
  const double incr = c > 1 ? 100.0 / (double)(c - 1) : 0.0;
  for (int b = lane; b < 33; b += 32) {
    double v = 0.0;
    const int k = hist[warp][b];
    for (int r = 0; r < k; ++r) v = __dadd_rn(v, incr);
    spfh_t[(size_t)i * 33 + b] = v;
  }
}

__global__ void __launch_bounds__(kFpWarps * 32) fpfh_kernel(const double* __restrict__ spfh_t, int n, const int* __restrict__ idx,
                                                             const double* __restrict__ d2, const int* __restrict__ cnt, int K,
                                                             double* __restrict__ out /*[33][n]*/) {
  // Neighbours are taken kFpBatch at a time: the batch's SPFH rows are loaded and divided together (independent fp64 divisions
  // overlap; one at a time, the division's ~400 cycles of dependent latency and two warp barriers were paid per neighbour and the
  // kernel ran at ~1000 cycles per neighbour), then added in neighbour order.  A neighbour beyond the count, or at distance zero (the
  // reference skips it), contributes +0.0 to sums that are never negative: the same bits as not adding it.
  constexpr int kFpBatch = 8;
  __shared__ double sv[kFpWarps][kFpBatch][34];
  __shared__ int s_idx[kFpWarps][kKnnMaxK];
  __shared__ double s_d2[kFpWarps][kKnnMaxK];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = blockIdx.x * kFpWarps + warp;
  if (i >= n) return;
  const int c = cnt[i];
  for (int t = lane; t < c; t += 32) { s_idx[warp][t] = idx[(size_t)i * K + t]; s_d2[warp][t] = d2[(size_t)i * K + t]; }
  __syncwarp();
  double acc = 0.0;    // bin `lane`
  double acc32 = 0.0;  // lane 3: bin 32 (the 33rd bin of the batch's neighbours is divided by lanes 0 .. kFpBatch-1 in ONE pass -- a second
                       // division per neighbour for lane 31 alone cost the warp as many instructions as the other 32 bins)
  double gsum = 0.0;   // lanes 0..2: group sums
  for (int t0 = 1; t0 < c; t0 += kFpBatch) {
    double row[kFpBatch], dist[kFpBatch];
#pragma unroll
    for (int b = 0; b < kFpBatch; ++b) {
      const int t = t0 + b;
      const bool ok = t < c;
      const int j = ok ? s_idx[warp][t] : i;
      dist[b] = ok ? s_d2[warp][t] : 0.0;
      row[b] = spfh_t[(size_t)j * 33 + lane];
    }
    {
      const int t = t0 + lane;
      const bool ok = lane < kFpBatch && t < c;
      const double dd = ok ? s_d2[warp][t] : 0.0;
      const double r32 = ok ? spfh_t[(size_t)s_idx[warp][t] * 33 + 32] : 0.0;
      const double v32 = dd != 0.0 ? __ddiv_rn(r32, dd) : 0.0;
      if (lane < kFpBatch) sv[warp][lane][32] = v32;
    }
#pragma unroll
    for (int b = 0; b < kFpBatch; ++b) {
      const double v = dist[b] != 0.0 ? __ddiv_rn(row[b], dist[b]) : 0.0;
      sv[warp][b][lane] = v;
      acc = __dadd_rn(acc, v);
    }
    __syncwarp();
    // group g (lane g < 3) adds its 11 bins in bin order, neighbour after neighbour: bins 11g .. 11g+10; lane 3 adds bin 32
    if (lane < 3) {
#pragma unroll
      for (int b = 0; b < kFpBatch; ++b)
#pragma unroll
        for (int k = 0; k < 11; ++k) gsum = __dadd_rn(gsum, sv[warp][b][11 * lane + k]);
    } else if (lane == 3) {
#pragma unroll
      for (int b = 0; b < kFpBatch; ++b) acc32 = __dadd_rn(acc32, sv[warp][b][32]);
    }
    __syncwarp();
  }
  acc32 = __shfl_sync(FULL, acc32, 3);   // lane 31 writes bin 32 below
  // scale by 100 / group sum, add the point's own SPFH
  double g0 = __shfl_sync(FULL, gsum, 0), g1 = __shfl_sync(FULL, gsum, 1), g2 = __shfl_sync(FULL, gsum, 2);
  if (g0 != 0.0) g0 = __ddiv_rn(100.0, g0);
  if (g1 != 0.0) g1 = __ddiv_rn(100.0, g1);
  if (g2 != 0.0) g2 = __ddiv_rn(100.0, g2);
  double r = 0.0, r32 = 0.0;
  if (c > 1) {
    const double s = lane < 11 ? g0 : (lane < 22 ? g1 : g2);
    r = __dadd_rn(__dmul_rn(acc, s), spfh_t[(size_t)i * 33 + lane]);
    if (lane == 31) r32 = __dadd_rn(__dmul_rn(acc32, g2), spfh_t[(size_t)i * 33 + 32]);
  }
  out[(size_t)lane * n + i] = r;
  if (lane == 31) out[(size_t)32 * n + i] = r32;
}

// transposed copy helper: [33][n] <- [n][33] (SPFH export for tests)
__global__ void __launch_bounds__(256) transpose33_kernel(const double* __restrict__ in_t, int n, double* __restrict__ out) {
  const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
  if (e >= 33ll * n) return;
  const int b = (int)(e / n), i = (int)(e % n);
  out[e] = in_t[(size_t)i * 33 + b];
}

// ------------------------------------------------------------------------------------------------------------------
// a-2 (exact form)  nearest neighbour in feature space: sequential sum_t (a_t - b_t)^2 in fp64, lowest index wins
// ties (B.6).  Brute force; also the re-rank / fallback of the tensor-core shortlist (nn_gemm.cu).
// One thread per query (query in registers), reference rows broadcast from shared memory.
// ------------------------------------------------------------------------------------------------------------------
constexpr int kNnDim = 33;
constexpr int kNnTile = 64;

__global__ void __launch_bounds__(128) nn_exact33_kernel(const double* __restrict__ fq, int nq, const double* __restrict__ fr, int nr,
                                                         int* __restrict__ nn, double* __restrict__ best_d) {
  __shared__ double tile[kNnTile * kNnDim];
  const int q = blockIdx.x * 128 + threadIdx.x;
  double x[kNnDim];
#pragma unroll
  for (int t = 0; t < kNnDim; ++t) x[t] = q < nq ? fq[(size_t)q * kNnDim + t] : 0.0;
  double bd = CUDART_INF; int best = -1;
  for (int base = 0; base < nr; base += kNnTile) {
    const int m = min(kNnTile, nr - base);
    __syncthreads();
    for (int e = threadIdx.x; e < m * kNnDim; e += 128) tile[e] = fr[(size_t)base * kNnDim + e];
    __syncthreads();
    for (int j = 0; j < m; ++j) {
      double s = 0;
#pragma unroll
      for (int t = 0; t < kNnDim; ++t) { const double d = __dsub_rn(x[t], tile[j * kNnDim + t]); s = __dadd_rn(s, __dmul_rn(d, d)); }
      if (s < bd) { bd = s; best = base + j; }
    }
  }
  if (q < nq) { nn[q] = best; if (best_d) best_d[q] = bd; }
}

// generic dimension (slow path, dim <= 256): query row read from global
__global__ void __launch_bounds__(128) nn_exact_generic_kernel(const double* __restrict__ fq, int nq, const double* __restrict__ fr, int nr,
                                                               int dim, int* __restrict__ nn, double* __restrict__ best_d) {
  const int q = blockIdx.x * 128 + threadIdx.x;
  if (q >= nq) return;
  double bd = CUDART_INF; int best = -1;
  for (int j = 0; j < nr; ++j) {
    double s = 0;
    for (int t = 0; t < dim; ++t) { const double d = __dsub_rn(fq[(size_t)q * dim + t], fr[(size_t)j * dim + t]); s = __dadd_rn(s, __dmul_rn(d, d)); }
    if (s < bd) { bd = s; best = j; }
  }
  nn[q] = best; if (best_d) best_d[q] = bd;
}

// mutual filter: pairs (i, nn01[i]) with nn10[nn01[i]] == i, ascending i
__global__ void __launch_bounds__(256) mutual_pairs_kernel(const int* __restrict__ nn01, int na, const int* __restrict__ nn10,
                                                           int* __restrict__ pairs, unsigned long long* ws, int* count) {
  __shared__ int s_tile;
  __shared__ unsigned s_warp[8];
  __shared__ unsigned long long s_base;
  const int tile = take_ticket(ws, &s_tile);
  const int i = tile * 256 + threadIdx.x;
  int j = -1;
  if (i < na) { j = nn01[i]; if (j < 0 || nn10[j] != i) j = -1; }
  const unsigned k = j >= 0 ? 1u : 0u;
  unsigned blk_total;
  const unsigned local = block_scan_256(k, s_warp, &blk_total);
  if (threadIdx.x == 0) {
    s_base = lookback(ws, tile, blk_total);
    if (tile == (int)gridDim.x - 1) *count = (int)(s_base + blk_total);
  }
  __syncthreads();
  if (k) { const size_t m = (size_t)s_base + local; pairs[2 * m] = i; pairs[2 * m + 1] = j; }
}

// gather the 3xN operands of `solver.solve` from the matched pairs (HelpersRealtime.py:142-143)
__global__ void __launch_bounds__(256) gather_corr_kernel(const double* __restrict__ a_pts, const double* __restrict__ b_pts,
                                                          const int* __restrict__ pairs, const int* __restrict__ count,
                                                          double* __restrict__ src, double* __restrict__ dst) {
  const int m = blockIdx.x * 256 + threadIdx.x;
  if (m >= *count) return;
  const int i = pairs[2 * m], j = pairs[2 * m + 1];
#pragma unroll
  for (int k = 0; k < 3; ++k) { src[3 * (size_t)m + k] = a_pts[3 * (size_t)i + k]; dst[3 * (size_t)m + k] = b_pts[3 * (size_t)j + k]; }
}


// ------------------------------------------------------------------------------------------------------------------
// Grid-accelerated exact kNN for large clouds (the raw back-projected cloud has tens of thousands of points; brute
// force is O(n^2)).  The points are binned into a uniform grid (cell edge h from the cloud's extent and K), sorted by
// cell (radix.cu), and every query walks cubic shells of cells around its own cell until the K-th best distance is
// provably final: every point within r*h of the query lies in the cells of shells 0..r.  Same candidate buffer / merge
// as the brute-force kernel, same (d2, index) order, same fp64 distance expression -> identical output.
// Everything is sized on the device (no host round trip): the dense cell table holds at most 2^24 cells.
// ------------------------------------------------------------------------------------------------------------------
constexpr long long kGridMaxCells = 1ll << 24;
inline size_t pad256_(size_t b) { return (b + 255) / 256 * 256; }

struct GridDev {
  double org[3];
  double h;
  int dim[3];
  int pad_;
};

// single thread: grid geometry from the bounds
__global__ void grid_setup_kernel(const unsigned long long* __restrict__ bounds, int n, int K, double radius, GridDev* __restrict__ g, double h_scale) {
  double mn[3], mx[3], L[3];
  for (int a = 0; a < 3; ++a) { mn[a] = ordered_key_inv(bounds[a]); mx[a] = ordered_key_inv(bounds[3 + a]); L[a] = mx[a] - mn[a]; }
  const double A = L[0] * L[1] + L[1] * L[2] + L[0] * L[2];      // surface-like clouds: spacing ~ sqrt(area / n)
  const double Lmax = fmax(L[0], fmax(L[1], L[2]));
  double h;
  // sqrt(A / n) sqrt(K / pi) estimates the distance to the K-th neighbour on a surface of the bounding box's area; the shell walk
  // wants cells of twice that (h_scale = 1).  (A lane-per-query kernel over 3 x 3 x 3 blocks of half-size cells was tried in round 2:
  // 1.14 ms against 0.85 ms for the outlier-removal stage of the config-2 frame -- the ~10 % of queries whose K-th neighbour lies
  // outside the block pay more in the shell walk on the finer grid than the others save.)
  if (A > 0) h = 2.0 * h_scale * sqrt(A / (double)n) * sqrt((double)K / 3.14159265358979);
  else h = Lmax > 0 ? Lmax / 64.0 : 1.0;
  if (radius > 0 && radius < h) h = radius;
  if (!(h > 0) || !(h == h)) h = 1.0;
  long long d[3];
  for (;;) {
    long long cells = 1;
    for (int a = 0; a < 3; ++a) { d[a] = (long long)floor(L[a] / h) + 1; if (d[a] < 1) d[a] = 1; cells = (d[a] > kGridMaxCells) ? kGridMaxCells + 1 : cells * d[a]; if (cells > kGridMaxCells) break; }
    if (cells <= kGridMaxCells) break;
    h *= 1.26;
  }
  for (int a = 0; a < 3; ++a) { g->org[a] = mn[a]; g->dim[a] = (int)d[a]; }
  g->h = h; g->pad_ = 0;
}

__device__ __forceinline__ int grid_cell(double x, double org, double h) { return (int)fmin(fmax(floor((x - org) / h), -2.0e9), 2.0e9); }

__global__ void __launch_bounds__(256) grid_key_kernel(const double* __restrict__ pts, int n, const GridDev* __restrict__ gp,
                                                       unsigned long long* __restrict__ keys, uint32_t* __restrict__ vals) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const GridDev g = *gp;
  int c[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) { c[a] = grid_cell(pts[3 * (size_t)i + a], g.org[a], g.h); c[a] = min(max(c[a], 0), g.dim[a] - 1); }
  keys[i] = (unsigned long long)c[0] + (unsigned long long)g.dim[0] * ((unsigned long long)c[1] + (unsigned long long)g.dim[1] * (unsigned long long)c[2]);
  vals[i] = (uint32_t)i;
}

// sorted copies of the points (coalesced shell scans) and the [start, end) range of every occupied cell
// cell_range is NOT cleared between calls (16 M cells = 128 MB): an entry is valid only if it points into the current sorted array at a
// point of that very cell (cell_points below); stale or uninitialised entries of cells that are empty now fail that test.
__global__ void __launch_bounds__(256) grid_ranges_kernel(const double* __restrict__ pts, int n, const unsigned long long* __restrict__ keys,
                                                          const uint32_t* __restrict__ sidx, double* __restrict__ spts, int2* __restrict__ cell_range) {
  const int e = blockIdx.x * 256 + threadIdx.x;
  if (e >= n) return;
  const uint32_t i = sidx[e];
  spts[3 * (size_t)e] = pts[3 * (size_t)i]; spts[3 * (size_t)e + 1] = pts[3 * (size_t)i + 1]; spts[3 * (size_t)e + 2] = pts[3 * (size_t)i + 2];
  const unsigned long long k = keys[e];
  if (e == 0 || keys[e - 1] != k) cell_range[k].x = e;
  if (e == n - 1 || keys[e + 1] != k) cell_range[k].y = e + 1;
}

// [start, end) of cell k in the sorted array, empty when the entry is not a current one
__device__ __forceinline__ int2 cell_points(const int2* __restrict__ cell_range, const unsigned long long* __restrict__ skeys, int n, unsigned long long k) {
  const int2 cr = cell_range[k];
  if ((unsigned)cr.x < (unsigned)n && cr.y > cr.x && cr.y <= n && skeys[cr.x] == k) return cr;
  return make_int2(0, 0);
}

struct KnnGridSmem { KnnWarpSmem w[kKnnWarps]; };

__global__ void __launch_bounds__(kKnnWarps * 32, 5) knn_grid_kernel(const double* __restrict__ spts, const uint32_t* __restrict__ sidx, int n,
                                                                  const GridDev* __restrict__ gp, const int2* __restrict__ cell_range,
                                                                  const double* __restrict__ queries, int nq, int K, double radius,
                                                                  int* __restrict__ idx_out, double* __restrict__ d2_out, int* __restrict__ cnt_out,
                                                                  const unsigned long long* __restrict__ skeys) {
  extern __shared__ __align__(16) unsigned char knn_smem[];
  KnnWarpSmem& w = reinterpret_cast<KnnWarpSmem*>(knn_smem)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const int q = blockIdx.x * kKnnWarps + (threadIdx.x >> 5);
  if (q >= nq) return;
  const GridDev g = *gp;
  const double qx = queries[3 * (size_t)q], qy = queries[3 * (size_t)q + 1], qz = queries[3 * (size_t)q + 2];
  const int cq0 = grid_cell(qx, g.org[0], g.h), cq1 = grid_cell(qy, g.org[1], g.h), cq2 = grid_cell(qz, g.org[2], g.h);
  const bool use_r = radius > 0;
  const double r2 = __dmul_rn(radius, radius);
  // shells below r_first hold no cell, shells above r_last neither
  const int o0 = max(max(-cq0, cq0 - (g.dim[0] - 1)), 0), o1 = max(max(-cq1, cq1 - (g.dim[1] - 1)), 0), o2 = max(max(-cq2, cq2 - (g.dim[2] - 1)), 0);
  const int r_first = max(o0, max(o1, o2));
  const int r_last = max(max(max(cq0, g.dim[0] - 1 - cq0), max(cq1, g.dim[1] - 1 - cq1)), max(cq2, g.dim[2] - 1 - cq2));
  int cur = 0, a = 0, nb = 0;
  int r_begin = r_first;
  unsigned long long tk = ~0ull; int ti = 0x7fffffff;
  // ---- fast path (K <= 32): the whole 3 x 3 x 3 block of cells at once ------------------------------------------------------------
  // The shell walk below merges candidates into a sorted list batch by batch; its profile is 95 % selection machinery (sorts and merges
  // of 64-slot batches), 1 % distance arithmetic.  Here every lane evaluates its share (<= 16) of ALL candidates of the block into
  // registers, the K-th smallest of the 32 per-lane minima is a threshold that at least K candidates meet (each of those K lanes
  // holds one) and, for evenly spread candidates, not many more; the candidates up to it are collected and sorted ONCE (knn_flush).
  // The answer is final when the list is full and its K-th distance lies within the reach of the block (or the radius does); any other
  // case -- more than 64 candidates up to the threshold -- takes the shell walk; a K-th neighbour outside the block
  // continues the walk at shell 2 from the list found here.
  if (K <= 32) {
    int rb = 0, re = 0;                       // lane t < 9: sorted range of x-run t of the block
    if (lane < 9) {
      const int yy = cq1 + lane % 3 - 1, zz = cq2 + lane / 3 - 1;
      if (yy >= 0 && yy < g.dim[1] && zz >= 0 && zz < g.dim[2]) {
        const unsigned long long kb = (unsigned long long)g.dim[0] * ((unsigned long long)yy + (unsigned long long)g.dim[1] * (unsigned long long)zz);
        int b = 0x7fffffff, e = 0;
        for (int x = max(cq0 - 1, 0); x <= min(cq0 + 1, g.dim[0] - 1); ++x) {
          const int2 cr = cell_points(cell_range, skeys, n, kb + (unsigned long long)x);
          if (cr.y > cr.x) { b = min(b, cr.x); e = max(e, cr.y); }
        }
        if (e > b) { rb = b; re = e; }
      }
    }
    const bool inside = cq0 >= 0 && cq0 < g.dim[0] && cq1 >= 0 && cq1 < g.dim[1] && cq2 >= 0 && cq2 < g.dim[2];
    int len = re - rb, incl = len;            // inclusive prefix of the run lengths over lanes 0 .. 8
#pragma unroll
    for (int o = 1; o < 16; o <<= 1) { const int t = __shfl_up_sync(FULL, incl, o); if (lane >= o) incl += t; }
    const int C = __shfl_sync(FULL, incl, 8);
    if (inside && C > 0) {
      // two passes over the block's candidates, nothing held between them but the threshold (the points are L1-resident for the
      // second pass; holding 12 - 16 candidates per lane in registers cost spills at 4 CTAs per SM and capped the block at 512)
      auto for_each_candidate = [&](auto&& f) {
        int run = 0, run_end = __shfl_sync(FULL, incl, 0), run_b = __shfl_sync(FULL, rb, 0), run_first = 0;
        for (int c0 = 0; c0 < C; c0 += 32) {
          const int c = c0 + lane;
          while (true) {   // (uniform trip count: every lane takes part in the shuffles)
            const bool adv = c < C && c >= run_end;
            if (!__any_sync(FULL, adv)) break;
            const int nr = min(run + 1, 8);
            const int ne = __shfl_sync(FULL, incl, adv ? nr : 0), nbg = __shfl_sync(FULL, rb, adv ? nr : 0);
            if (adv) { run_first = run_end; run = nr; run_end = ne; run_b = nbg; }
          }
          unsigned long long key = ~0ull;
          int sp = 0;
          if (c < C) {
            sp = run_b + (c - run_first);
            const double d2 = sq3_ref(__dsub_rn(spts[3 * (size_t)sp], qx), __dsub_rn(spts[3 * (size_t)sp + 1], qy), __dsub_rn(spts[3 * (size_t)sp + 2], qz));
            if (!use_r || d2 < r2) key = (unsigned long long)__double_as_longlong(d2);
          }
          if (!f(key, sp)) break;
        }
      };
      unsigned long long mymin = ~0ull;
      for_each_candidate([&](unsigned long long key, int) { mymin = key < mymin ? key : mymin; return true; });
      // K-th smallest of the 32 lane minima, ordered by (key, lane): rank by counting (32 shuffles).  Every candidate with a key up to
      // it is kept -- at least K (one per lane of rank < K); ties on the key only add candidates, the sort below decides by (key, id).
      int rank = 0;
      for (int j = 0; j < 32; ++j) {
        const unsigned long long ok = __shfl_sync(FULL, mymin, j);
        rank += (ok < mymin || (ok == mymin && j < lane)) ? 1 : 0;
      }
      const unsigned pick = __ballot_sync(FULL, rank == K - 1 && mymin != ~0ull);
      const int n_valid_lanes = __popc(__ballot_sync(FULL, mymin != ~0ull));
      unsigned long long thr_k = ~0ull - 1ull;                         // fewer than K lanes hold anything: take everything (padding is ~0)
      if (n_valid_lanes >= K && pick) thr_k = __shfl_sync(FULL, mymin, __ffs(pick) - 1);
      int total = 0;
      for_each_candidate([&](unsigned long long key, int sp) {
        const bool keep = key <= thr_k;
        const unsigned bal = __ballot_sync(FULL, keep);
        const int cnt = __popc(bal);
        if (total + cnt > kKnnBuf) { total = kKnnBuf + 1; return false; }   // uniform: more than one batch up to the threshold
        if (keep) { const int pos = total + __popc(bal & ((1u << lane) - 1u)); w.ck[pos] = key; w.ci[pos] = (int)sidx[sp]; }
        total += cnt;
        return true;
      });
      if (total <= kKnnBuf) {
        __syncwarp();
        a = knn_flush(w, cur, 0, total, K, lane);
        const double reach1 = g.h * (1.0 - 1e-9);
        const bool done = (use_r && reach1 >= radius) ||
                          (a == K && __longlong_as_double((long long)w.lk[cur][K - 1]) <= reach1 * reach1);
        if (done) {
          for (int t = lane; t < K; t += 32) {
            idx_out[(size_t)q * K + t] = t < a ? w.li[cur][t] : -1;
            d2_out[(size_t)q * K + t] = t < a ? __longlong_as_double((long long)w.lk[cur][t]) : 0.0;
          }
          if (lane == 0) cnt_out[q] = a;
          return;
        }
        // not final: the list is what the walk would hold after shells 0 and 1 (the best K of every point of the block, same order),
        // so the walk goes on at shell 2 with the K-th key as its threshold -- the block is not scanned again through the merges
        __syncwarp();
        r_begin = 2;
        if (a == K) { tk = w.lk[cur][K - 1]; ti = w.li[cur][K - 1]; }
      }
    }
  }
  for (int r = r_begin; r <= r_last; ++r) {
    for (int dz = -r; dz <= r; ++dz) {
      const int cz = cq2 + dz;
      if (cz < 0 || cz >= g.dim[2]) continue;
      for (int dy = -r; dy <= r; ++dy) {
        const int cy = cq1 + dy;
        if (cy < 0 || cy >= g.dim[1]) continue;
        const bool face = (abs(dz) == r) || (abs(dy) == r);
        // face rows of the shell: the whole x-run; interior rows: only the two end cells
        const int n_runs = (face || r == 0) ? 1 : 2;
        for (int run = 0; run < n_runs; ++run) {
          int xlo, xhi;
          if (face || r == 0) { xlo = cq0 - r; xhi = cq0 + r; }
          else if (run == 0) { xlo = xhi = cq0 - r; }
          else { xlo = xhi = cq0 + r; }
          xlo = max(xlo, 0); xhi = min(xhi, g.dim[0] - 1);
          if (xlo > xhi) continue;
          const unsigned long long kbase = (unsigned long long)g.dim[0] * ((unsigned long long)cy + (unsigned long long)g.dim[1] * (unsigned long long)cz);
          // cells of an x-run are consecutive keys, so their points are one contiguous range of the sorted array
          int b = 0x7fffffff, e = 0;
          for (int x = xlo + lane; x <= xhi; x += 32) {
            const int2 cr = cell_points(cell_range, skeys, n, kbase + (unsigned long long)x);
            if (cr.y > cr.x) { b = min(b, cr.x); e = max(e, cr.y); }
          }
          b = __reduce_min_sync(FULL, b); e = __reduce_max_sync(FULL, e);
          for (int s0 = b; s0 < e; s0 += 32) {
            const int sp = s0 + lane;
            bool qual = false;
            unsigned long long key = 0; int id = 0;
            if (sp < e) {
              const double d2 = sq3_ref(__dsub_rn(spts[3 * (size_t)sp], qx), __dsub_rn(spts[3 * (size_t)sp + 1], qy), __dsub_rn(spts[3 * (size_t)sp + 2], qz));
              key = (unsigned long long)__double_as_longlong(d2);
              id = (int)sidx[sp];
              qual = (!use_r || d2 < r2) && key_less(key, id, tk, ti);
            }
            const unsigned bal = __ballot_sync(FULL, qual);
            if (bal) {
              if (qual) { const int p = nb + __popc(bal & ((1u << lane) - 1u)); w.ck[p] = key; w.ci[p] = id; }
              nb += __popc(bal);
              __syncwarp();
              if (nb > 32) {
                a = knn_flush(w, cur, a, nb, K, lane);
                nb = 0;
                if (a == K) { tk = w.lk[cur][K - 1]; ti = w.li[cur][K - 1]; }
              }
            }
          }
        }
      }
    }
    if (nb > 0) {
      a = knn_flush(w, cur, a, nb, K, lane);
      nb = 0;
      if (a == K) { tk = w.lk[cur][K - 1]; ti = w.li[cur][K - 1]; }
    }
    // everything within reach = r*h (minus a rounding guard for the cell assignment) has been seen
    const double reach = (double)r * g.h * (1.0 - 1e-9);
    if (use_r && reach >= radius) break;
    if (a == K && __longlong_as_double((long long)tk) <= reach * reach) break;
  }
  for (int t = lane; t < K; t += 32) {
    idx_out[(size_t)q * K + t] = t < a ? w.li[cur][t] : -1;
    d2_out[(size_t)q * K + t] = t < a ? __longlong_as_double((long long)w.lk[cur][t]) : 0.0;
  }
  if (lane == 0) cnt_out[q] = a;
}

size_t knn_smem_bytes() { return sizeof(double) * 3 * kKnnTile + sizeof(KnnWarpSmem) * kKnnWarps; }
size_t knn_grid_smem_bytes() { return sizeof(KnnGridSmem); }
size_t knn_grid_ws_bytes(int n) {
  return 2 * pad256_((size_t)n * 8) + 2 * pad256_((size_t)n * 4) + pad256_(radix_table_words(n) * 4) + pad256_((size_t)n * 24) +
         pad256_((size_t)kGridMaxCells * 8) + 1024;
}


}  // namespace

// ------------------------------------------------------------------------------------------------------------------
// launchers used by the C ABI below
// ------------------------------------------------------------------------------------------------------------------
constexpr double kKnnHScale = 0.55;      // cell size relative to twice the estimated K-th neighbour distance (B2R_KNN_HSCALE overrides: A/B runs)
constexpr int kKnnGridMinPoints = 5632;  // below this the brute-force kernel is faster than building a grid: the grid is ~110 us of
                                         // latency-bound sort launches + ~24 us per 1000 points (normals + FPFH searches), brute force 61 us x (n / 2785)^2

static int launch_knn_brute(const double* pts, int n, const double* queries, int nq, int K, double radius, int* idx, double* d2, int* cnt,
                            cudaStream_t st) {
  static PerDeviceOnce attr;
  attr.run([] { return cudaFuncSetAttribute(knn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)knn_smem_bytes()); });
  knn_kernel<<<(nq + kKnnWarps - 1) / kKnnWarps, kKnnWarps * 32, knn_smem_bytes(), st>>>(pts, n, queries, nq, K, radius, idx, d2, cnt);
  return 1;
}

// ws: knn_grid_ws_bytes(n) bytes of device scratch (256-byte aligned)
static int launch_knn_grid(const double* pts, int n, const double* queries, int nq, int K, double radius, int* idx, double* d2, int* cnt,
                           void* ws, cudaStream_t st) {
  static PerDeviceOnce attr;
  attr.run([] { return cudaFuncSetAttribute(knn_grid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)knn_grid_smem_bytes()); });
  char* p = (char*)ws;
  auto take = [&](size_t bytes) { char* r = p; p += pad256_(bytes); return (void*)r; };
  unsigned long long* keys = (unsigned long long*)take((size_t)n * 8);
  unsigned long long* keys_alt = (unsigned long long*)take((size_t)n * 8);
  uint32_t* vals = (uint32_t*)take((size_t)n * 4);
  uint32_t* vals_alt = (uint32_t*)take((size_t)n * 4);
  uint32_t* table = (uint32_t*)take(radix_table_words(n) * 4);
  double* spts = (double*)take((size_t)n * 24);
  int2* cell_range = (int2*)take((size_t)kGridMaxCells * 8);
  unsigned long long* bounds = (unsigned long long*)take(64);
  GridDev* g = (GridDev*)take(sizeof(GridDev));
  const int tiles = (n + 255) / 256;
  cudaMemsetAsync(bounds, 0xff, 24, st);
  cudaMemsetAsync(bounds + 3, 0, 24, st);
  bounds_kernel<<<tiles < 296 ? tiles : 296, 256, 0, st>>>(pts, n, bounds);
  static const double h_scale = [] { const char* e = getenv("B2R_KNN_HSCALE"); const double v = e ? atof(e) : 0.0; return v > 0.05 && v < 10 ? v : kKnnHScale; }();
  grid_setup_kernel<<<1, 1, 0, st>>>(bounds, n, K, radius, g, h_scale);
  grid_key_kernel<<<tiles, 256, 0, st>>>(pts, n, g, keys, vals);
  bool in_alt = false;
  int launches = 3 + radix_sort_pairs((uint64_t*)keys, vals, (uint64_t*)keys_alt, vals_alt, table, (size_t)n, 0, 24, st, &in_alt);
  const unsigned long long* sk = in_alt ? keys_alt : keys;
  const uint32_t* sv = in_alt ? vals_alt : vals;
  grid_ranges_kernel<<<tiles, 256, 0, st>>>(pts, n, sk, sv, spts, cell_range);
  knn_grid_kernel<<<(nq + kKnnWarps - 1) / kKnnWarps, kKnnWarps * 32, knn_grid_smem_bytes(), st>>>(spts, sv, n, g, cell_range, queries, nq, K,
                                                                                                    radius, idx, d2, cnt, sk);
  return launches + 2;
}

}  // namespace b2r

using namespace b2r;

namespace {

// scratch carving from the handle's pre-step workspace
struct Carver {
  char* p; char* end;
  template <typename T> T* take(size_t count) {
    char* r = p;
    p += (count * sizeof(T) + 255) / 256 * 256;
    return reinterpret_cast<T*>(r);
  }
};

b2r_status pre_ws(b2r_solver* s, size_t bytes, Carver* c) {
  CK(cudaSetDevice(s->device));
  CK(s->d_pre.ensure(bytes + 4096));
  c->p = (char*)s->d_pre.p; c->end = c->p + s->d_pre.bytes;
  return B2R_OK;
}

inline size_t pad256(size_t b) { return (b + 255) / 256 * 256; }

// kNN dispatch: brute force for small clouds, uniform grid for large ones (identical results).  The grid scratch lives in
// its own buffer (d_nn is free while the pre-steps run) so that callers keep using d_pre for their own arrays.
b2r_status run_knn(b2r_solver* s, const double* pts, int n, const double* queries, int nq, int K, double radius, int* idx, double* d2,
                   int* cnt, int force /*0 auto, 1 brute, 2 grid*/) {
  const bool grid = force == 2 || (force == 0 && n >= kKnnGridMinPoints);
  if (!grid) { s->launches += launch_knn_brute(pts, n, queries, nq, K, radius, idx, d2, cnt, s->stream); return B2R_OK; }
  CK(s->d_grid.ensure(knn_grid_ws_bytes(n)));
  s->launches += launch_knn_grid(pts, n, queries, nq, K, radius, idx, d2, cnt, s->d_grid.p, s->stream);
  return B2R_OK;
}

}  // namespace

namespace b2r {
// kNN dispatch for the other translation units (refine.cu)
b2r_status internal_knn(b2r_solver* s, const double* pts, int n, const double* queries, int nq, int K, double radius, int* idx,
                        double* d2, int* cnt) {
  return run_knn(s, pts, n, queries, nq, K, radius, idx, d2, cnt, 0);
}
}  // namespace b2r

extern "C" {

b2r_status b2r_backproject(b2r_solver* s, const uint16_t* depth, const uint8_t* mask, int H, int W, int u0, int v0, int u1, int v1,
                           const b2r_camera* cam, int on_device, double* points_dev, int32_t* pixel_ids_dev, int capacity,
                           int* count_out) {
  B2R_NVTX("b2r:backproject");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!depth || !cam || !points_dev || H <= 0 || W <= 0 || capacity < 0 || !count_out) return fail(s, B2R_ERR_INVALID_ARG, "backproject arguments");
  if (!(cam->fx != 0) || !(cam->fy != 0)) return fail(s, B2R_ERR_INVALID_ARG, "fx, fy must be non-zero");
  u0 = u0 < 0 ? 0 : u0; v0 = v0 < 0 ? 0 : v0; u1 = u1 > W ? W : u1; v1 = v1 > H ? H : v1;
  *count_out = 0;
  if (u1 <= u0 || v1 <= v0) return B2R_OK;
  const int G = ((u1 + 7) >> 3) - (u0 >> 3);
  const long long groups = (long long)(v1 - v0) * G;
  const int tiles = (int)((groups + 255) / 256);
  const size_t img = (size_t)H * W;
  Carver c;
  b2r_status r = pre_ws(s, pad256(8 * (size_t)(tiles + 2)) + pad256(4 * (size_t)tiles) + pad256(2 * img) + pad256(img) + 1024, &c);
  if (r != B2R_OK) return r;
  unsigned long long* ws = c.take<unsigned long long>(tiles + 2);
  unsigned int* tile_counts = c.take<unsigned int>(tiles);
  int* d_count = c.take<int>(1);
  const uint16_t* d_depth = depth; const uint8_t* d_mask = mask;
  if (!on_device) {  // host image: one H2D copy each (pageable or pinned)
    uint16_t* dd = c.take<uint16_t>(img);
    CK(cudaMemcpyAsync(dd, depth, 2 * img, cudaMemcpyHostToDevice, s->stream));
    d_depth = dd;
    if (mask) { uint8_t* dm = c.take<uint8_t>(img); CK(cudaMemcpyAsync(dm, mask, img, cudaMemcpyHostToDevice, s->stream)); d_mask = dm; }
  }
  CK(cudaMemsetAsync(ws, 0, 8 * (size_t)(tiles + 2), s->stream));
  BackprojArgs a;
  a.depth = d_depth; a.mask = d_mask; a.H = H; a.W = W; a.u0 = u0; a.v0 = v0; a.u1 = u1; a.v1 = v1;
  a.fx = cam->fx; a.fy = cam->fy; a.cx = cam->cx; a.cy = cam->cy; a.zmin = cam->zmin; a.zmax = cam->zmax;
  a.scale = (float)cam->depth_scale; a.trunc = (float)cam->depth_trunc;
  a.use_range = (cam->zmin > 0 || cam->zmax > 0) ? 1 : 0; a.use_trunc = cam->depth_trunc > 0 ? 1 : 0;
  a.vec_ok = (W % 8 == 0) && ((uintptr_t)d_depth % 16 == 0) && (!d_mask || (uintptr_t)d_mask % 8 == 0);
  a.pts = points_dev; a.pix = pixel_ids_dev; a.cap = capacity; a.ws = ws; a.count = d_count; a.tile_counts = tile_counts;
  static PerDeviceOnce bp_attr;
  CK(bp_attr.run([] {
    cudaError_t e = cudaFuncSetAttribute(backproject_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBpSmemBytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(backproject_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBpSmemBytes);
    return e;
  }));
  static const int two_pass_tiles = [] { const char* e = getenv("B2R_BP_TWO_PASS_TILES"); return e ? atoi(e) : 512; }();   // A/B runs
  if (tiles < two_pass_tiles) {       // a frame-sized bbox: one launch, the look-back chain is a few dozen tiles long
    backproject_kernel<0><<<tiles, 256, kBpSmemBytes, s->stream>>>(a);
    ++s->launches;
  } else {
    backproject_kernel<1><<<tiles, 256, 0, s->stream>>>(a);
    tile_scan_kernel<<<1, 1024, 0, s->stream>>>(tile_counts, tiles, d_count);
    backproject_kernel<2><<<tiles, 256, kBpSmemBytes, s->stream>>>(a);
    s->launches += 3;
  }
  { const b2r_status rr = b2r_internal_read_words(s, d_count, 1, count_out); if (rr != B2R_OK) return rr; }
  CK(cudaGetLastError());
  if (*count_out > capacity) return fail(s, B2R_ERR_LIMIT, "backproject: output capacity too small (count_out holds the needed size)");
  return B2R_OK;
}

b2r_status b2r_knn(b2r_solver* s, const double* pts_dev, int n, const double* queries_dev, int nq, int max_nn, double radius,
                   int32_t* idx_dev, double* d2_dev, int32_t* cnt_dev) {
  B2R_NVTX("b2r:knn");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!pts_dev || !queries_dev || !idx_dev || !d2_dev || !cnt_dev || n < 0 || nq < 0) return fail(s, B2R_ERR_INVALID_ARG, "knn arguments");
  if (max_nn < 1 || max_nn > kKnnMaxK) return fail(s, B2R_ERR_LIMIT, "knn: max_nn must be in [1, 128]");
  if (nq == 0) return B2R_OK;
  CK(cudaSetDevice(s->device));
  {
    const char* f = getenv("B2R_KNN");
    b2r_status r = run_knn(s, pts_dev, n, queries_dev, nq, max_nn, radius, idx_dev, d2_dev, cnt_dev, f ? (f[0] == 'b' ? 1 : (f[0] == 'g' ? 2 : 0)) : 0);
    if (r != B2R_OK) return r;
  }
  CK(cudaGetLastError());
  return B2R_OK;
}

b2r_status b2r_statistical_outlier(b2r_solver* s, const double* pts_dev, int n, int nb_neighbors, double std_ratio, uint8_t* keep_dev,
                                   double* avg_dev, double* pts_out_dev, int32_t* idx_out_dev, int* kept_out) {
  B2R_NVTX("b2r:statistical_outlier");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!pts_dev || n < 0 || nb_neighbors < 1 || !kept_out) return fail(s, B2R_ERR_INVALID_ARG, "statistical_outlier arguments");
  *kept_out = 0;
  if (n == 0) return B2R_OK;
  const int K = nb_neighbors < n ? nb_neighbors : n;
  if (K > kKnnMaxK) return fail(s, B2R_ERR_LIMIT, "statistical_outlier: nb_neighbors must be <= 128");
  const int tiles = (n + 255) / 256;
  Carver c;
  b2r_status r = pre_ws(s, pad256((size_t)n * K * 4) + pad256((size_t)n * K * 8) + 2 * pad256((size_t)n * 4) + pad256((size_t)n * 8) +
                               pad256(n) + pad256(8 * (size_t)(tiles + 2)) + 1024, &c);
  if (r != B2R_OK) return r;
  int* idx = c.take<int>((size_t)n * K);
  double* d2 = c.take<double>((size_t)n * K);
  int* cnt = c.take<int>(n);
  double* avg = avg_dev ? avg_dev : c.take<double>(n);
  uint8_t* keep = keep_dev ? keep_dev : c.take<uint8_t>(n);
  double* thr = c.take<double>(4);
  unsigned long long* ws = c.take<unsigned long long>(tiles + 2);
  int* d_count = c.take<int>(1);
  cudaStream_t st = s->stream;
  r = run_knn(s, pts_dev, n, pts_dev, n, K, 0.0, idx, d2, cnt, 0);
  if (r != B2R_OK) return r;
  avg_dist_kernel<<<tiles, 256, 0, st>>>(d2, cnt, n, K, avg);
  outlier_threshold_kernel<<<1, 1024, 0, st>>>(avg, n, std_ratio, thr);
  outlier_keep_kernel<<<tiles, 256, 0, st>>>(avg, n, thr, keep);
  s->launches += 3;
  if (pts_out_dev) {
    CK(cudaMemsetAsync(ws, 0, 8 * (size_t)(tiles + 2), st));
    compact_points_kernel<<<(n + kCompactTile - 1) / kCompactTile, 256, 0, st>>>(pts_dev, keep, n, pts_out_dev, idx_out_dev, ws, d_count);
    ++s->launches;
    { const b2r_status rr = b2r_internal_read_words(s, d_count, 1, kept_out); if (rr != B2R_OK) return rr; }
  } else {
    std::vector<uint8_t> h(n);
    CK(cudaMemcpyAsync(h.data(), keep, n, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    int k = 0; for (int i = 0; i < n; ++i) k += h[i];
    *kept_out = k;
  }
  CK(cudaGetLastError());
  return B2R_OK;
}

b2r_status b2r_voxel_down(b2r_solver* s, const double* pts_dev, int n, double voxel, double* out_dev, int32_t* voxel_of_point_dev,
                          int32_t* counts_dev, int* n_voxels_out) {
  B2R_NVTX("b2r:voxel_down");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!pts_dev || !out_dev || n < 0 || !(voxel > 0) || !n_voxels_out) return fail(s, B2R_ERR_INVALID_ARG, "voxel_down arguments");
  *n_voxels_out = 0;
  if (n == 0) return B2R_OK;
  const int tiles = (n + 255) / 256;
  Carver c;
  b2r_status r = pre_ws(s, 2 * pad256((size_t)n * 8) + 2 * pad256((size_t)n * 4) + pad256(radix_table_words(n) * 4) + pad256((size_t)n * 4) +
                               pad256(8 * (size_t)(tiles + 2)) + 2048, &c);
  if (r != B2R_OK) return r;
  unsigned long long* keys = c.take<unsigned long long>(n);
  unsigned long long* keys_alt = c.take<unsigned long long>(n);
  uint32_t* vals = c.take<uint32_t>(n);
  uint32_t* vals_alt = c.take<uint32_t>(n);
  uint32_t* table = c.take<uint32_t>(radix_table_words(n));
  int* seg_start = c.take<int>(n);
  unsigned long long* ws = c.take<unsigned long long>(tiles + 2);
  unsigned long long* bounds = c.take<unsigned long long>(8);
  int* d_count = c.take<int>(1);
  cudaStream_t st = s->stream;
  const unsigned long long init[6] = {~0ull, ~0ull, ~0ull, 0ull, 0ull, 0ull};
  CK(cudaMemcpyAsync(bounds, init, sizeof(init), cudaMemcpyHostToDevice, st));
  int gb = tiles < 296 ? tiles : 296;
  bounds_kernel<<<gb, 256, 0, st>>>(pts_dev, n, bounds);
  ++s->launches;
  unsigned long long hb[6];
  { const b2r_status rr = b2r_internal_read_words(s, bounds, 12, hb); if (rr != B2R_OK) return rr; }
  VoxelGrid g;
  long long dim[3];
  for (int a = 0; a < 3; ++a) {
    const double mn = ordered_key_inv(hb[a]), mx = ordered_key_inv(hb[3 + a]);
    g.org[a] = mn - voxel * 0.5;
    dim[a] = (long long)floor((mx - g.org[a]) / voxel) + 1;
    if (!(dim[a] > 0)) return fail(s, B2R_ERR_INVALID_ARG, "voxel_down: non-finite coordinates");
  }
  g.vs = voxel; g.dim0 = dim[0]; g.dim1 = dim[1];
  const long double cells = (long double)dim[0] * (long double)dim[1] * (long double)dim[2];
  if (cells > 9.0e18L) return fail(s, B2R_ERR_LIMIT, "voxel_down: grid has more than 2^63 cells; voxel too small for the cloud's extent");
  int bits = 1;
  while (bits < 64 && (long double)(1ull << bits) < cells) ++bits;
  voxel_key_kernel<<<tiles, 256, 0, st>>>(pts_dev, n, g, keys, vals);
  ++s->launches;
  bool in_alt = false;
  s->launches += radix_sort_pairs((uint64_t*)keys, vals, (uint64_t*)keys_alt, vals_alt, table, (size_t)n, 0, bits, st, &in_alt);
  const unsigned long long* sk = in_alt ? keys_alt : keys;
  const uint32_t* sv = in_alt ? vals_alt : vals;
  CK(cudaMemsetAsync(ws, 0, 8 * (size_t)(tiles + 2), st));
  voxel_heads_kernel<<<(n + kCompactTile - 1) / kCompactTile, 256, 0, st>>>(sk, n, seg_start, ws, d_count);
  voxel_mean_kernel<<<(n + 127) / 128, 128, 0, st>>>(pts_dev, n, sv, seg_start, d_count, out_dev, voxel_of_point_dev, counts_dev);
  s->launches += 2;
  { const b2r_status rr = b2r_internal_read_words(s, d_count, 1, n_voxels_out); if (rr != B2R_OK) return rr; }
  CK(cudaGetLastError());
  return B2R_OK;
}

b2r_status b2r_normals(b2r_solver* s, const double* pts_dev, int n, double radius, int max_nn, int orient, const double* view3,
                       double* normals_dev) {
  B2R_NVTX("b2r:normals");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!pts_dev || !normals_dev || n < 0 || max_nn < 1) return fail(s, B2R_ERR_INVALID_ARG, "normals arguments");
  if (n == 0) return B2R_OK;
  int K = max_nn < n ? max_nn : n;
  if (K < 1) K = 1;
  if (K > kKnnMaxK) return fail(s, B2R_ERR_LIMIT, "normals: max_nn must be <= 128");
  Carver c;
  b2r_status r = pre_ws(s, pad256((size_t)n * K * 4) + pad256((size_t)n * K * 8) + pad256((size_t)n * 4) + 1024, &c);
  if (r != B2R_OK) return r;
  int* idx = c.take<int>((size_t)n * K);
  double* d2 = c.take<double>((size_t)n * K);
  int* cnt = c.take<int>(n);
  r = run_knn(s, pts_dev, n, pts_dev, n, K, radius, idx, d2, cnt, 0);
  if (r != B2R_OK) return r;
  const double vx = view3 ? view3[0] : 0, vy = view3 ? view3[1] : 0, vz = view3 ? view3[2] : 0;
  normals_kernel<<<(n + 127) / 128, 128, 0, s->stream>>>(pts_dev, n, idx, cnt, K, view3 ? orient : 0, vx, vy, vz, normals_dev);
  ++s->launches;
  CK(cudaGetLastError());
  return B2R_OK;
}

b2r_status b2r_fpfh(b2r_solver* s, const double* pts_dev, const double* normals_dev, int n, double radius, int max_nn, double* fpfh_dev,
                    double* spfh_dev) {
  B2R_NVTX("b2r:fpfh");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!pts_dev || !normals_dev || !fpfh_dev || n < 0 || max_nn < 1) return fail(s, B2R_ERR_INVALID_ARG, "fpfh arguments");
  if (n == 0) return B2R_OK;
  int K = max_nn < n ? max_nn : n;
  if (K < 1) K = 1;
  if (K > kKnnMaxK) return fail(s, B2R_ERR_LIMIT, "fpfh: max_nn must be <= 128");
  Carver c;
  b2r_status r = pre_ws(s, pad256((size_t)n * K * 4) + pad256((size_t)n * K * 8) + pad256((size_t)n * 4) + pad256((size_t)n * 33 * 8) + 1024, &c);
  if (r != B2R_OK) return r;
  int* idx = c.take<int>((size_t)n * K);
  double* d2 = c.take<double>((size_t)n * K);
  int* cnt = c.take<int>(n);
  double* spfh_t = c.take<double>((size_t)n * 33);
  cudaStream_t st = s->stream;
  r = run_knn(s, pts_dev, n, pts_dev, n, K, radius, idx, d2, cnt, 0);
  if (r != B2R_OK) return r;
  const int blocks = (n + kFpWarps - 1) / kFpWarps;
  spfh_kernel<<<blocks, kFpWarps * 32, 0, st>>>(pts_dev, normals_dev, n, idx, cnt, K, spfh_t);
  fpfh_kernel<<<blocks, kFpWarps * 32, 0, st>>>(spfh_t, n, idx, d2, cnt, K, fpfh_dev);
  s->launches += 2;
  if (spfh_dev) { transpose33_kernel<<<(int)((33ll * n + 255) / 256), 256, 0, st>>>(spfh_t, n, spfh_dev); ++s->launches; }
  CK(cudaGetLastError());
  return B2R_OK;
}

b2r_status b2r_nn_exact(b2r_solver* s, const double* fq_dev, int nq, const double* fr_dev, int nr, int dim, int32_t* nn_dev,
                        double* dist_dev) {
  B2R_NVTX("b2r:nn_exact");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!fq_dev || !fr_dev || !nn_dev || nq < 0 || nr < 0 || dim < 1) return fail(s, B2R_ERR_INVALID_ARG, "nn_exact arguments");
  if (nq == 0) return B2R_OK;
  CK(cudaSetDevice(s->device));
  if (dim == kNnDim) nn_exact33_kernel<<<(nq + 127) / 128, 128, 0, s->stream>>>(fq_dev, nq, fr_dev, nr, nn_dev, dist_dev);
  else nn_exact_generic_kernel<<<(nq + 127) / 128, 128, 0, s->stream>>>(fq_dev, nq, fr_dev, nr, dim, nn_dev, dist_dev);
  ++s->launches;
  CK(cudaGetLastError());
  return B2R_OK;
}

b2r_status b2r_mutual_pairs(b2r_solver* s, const int32_t* nn01_dev, int na, const int32_t* nn10_dev, int nb, int32_t* pairs_dev,
                            int* n_pairs_out) {
  B2R_NVTX("b2r:mutual_pairs");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!nn01_dev || !nn10_dev || !pairs_dev || na < 0 || nb < 0 || !n_pairs_out) return fail(s, B2R_ERR_INVALID_ARG, "mutual_pairs arguments");
  *n_pairs_out = 0;
  if (na == 0 || nb == 0) return B2R_OK;
  const int tiles = (na + 255) / 256;
  Carver c;
  b2r_status r = pre_ws(s, pad256(8 * (size_t)(tiles + 2)) + 512, &c);
  if (r != B2R_OK) return r;
  unsigned long long* ws = c.take<unsigned long long>(tiles + 2);
  int* d_count = c.take<int>(1);
  CK(cudaMemsetAsync(ws, 0, 8 * (size_t)(tiles + 2), s->stream));
  mutual_pairs_kernel<<<tiles, 256, 0, s->stream>>>(nn01_dev, na, nn10_dev, pairs_dev, ws, d_count);
  ++s->launches;
  { const b2r_status rr = b2r_internal_read_words(s, d_count, 1, n_pairs_out); if (rr != B2R_OK) return rr; }
  CK(cudaGetLastError());
  return B2R_OK;
}

b2r_status b2r_gather_correspondences(b2r_solver* s, const double* a_pts_dev, const double* b_pts_dev, const int32_t* pairs_dev, int n_pairs,
                                      double* src_dev, double* dst_dev) {
  B2R_NVTX("b2r:gather_correspondences");
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!a_pts_dev || !b_pts_dev || !pairs_dev || !src_dev || !dst_dev || n_pairs < 0) return fail(s, B2R_ERR_INVALID_ARG, "gather_correspondences arguments");
  if (n_pairs == 0) return B2R_OK;
  Carver c;
  b2r_status r = pre_ws(s, 512, &c);
  if (r != B2R_OK) return r;
  int* d_count = c.take<int>(1);
  CK(cudaMemcpyAsync(d_count, &n_pairs, sizeof(int), cudaMemcpyHostToDevice, s->stream));
  gather_corr_kernel<<<(n_pairs + 255) / 256, 256, 0, s->stream>>>(a_pts_dev, b_pts_dev, pairs_dev, d_count, src_dev, dst_dev);
  ++s->launches;
  CK(cudaStreamSynchronize(s->stream));  // n_pairs lives on the caller's stack
  CK(cudaGetLastError());
  return B2R_OK;
}

}  // extern "C"
