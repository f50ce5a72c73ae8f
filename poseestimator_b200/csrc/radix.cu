// Device-wide stable LSD radix sort of (u64 key, u32 payload) pairs, 8-bit digits.
//
// Used where the path needs a total order over more items than one CTA holds:
//   * adaptive voting over all N(N-1)/2 TIM scale ratios (upstream std::sorts 2P endpoints);
//   * voxel_down_sample (group points by voxel key, ascending point index inside a voxel).
// HBM-bound: every pass reads the keys twice (histogram, scatter) and moves 12 bytes per item.  Stability is what
// turns "sort by value" into the oracle's (value, entry-before-exit, index) order: the caller lays the items out in
// the order of the secondary key before the first pass.
//
// Per pass: (1) per-tile digit histogram -> table[digit][tile]; (2) exclusive scan of the table in (digit, tile)
// order, one CTA per digit, coalesced; (3) scatter: warp-level multisplit with __match_any_sync gives every item its
// stable rank inside the tile.
#include "common.cuh"

namespace b2r {

namespace {

constexpr int kRxThreads = 256;
constexpr int kRxWarps = kRxThreads / 32;
constexpr int kRxItems = 16;
constexpr int kRxTile = kRxThreads * kRxItems;  // 4096 items per CTA

// Batched form of every kernel: blockIdx.y = segment (one sort per problem), segment s has n_arr[s] items at offset s * seg_stride of
// the key / payload arrays and its own table at s * table_stride; n_arr == nullptr: one segment of n items.
struct RxBatch { const unsigned long long* n_arr; size_t key_stride, val_stride, table_stride; };   // strides in items / words

// ITEMS per thread: 16 (tiles of 4096 items) for large sorts, 4 (tiles of 1024) below 256 k items -- a frame-sized sort (56 k points)
// then runs on 56 CTAs instead of 14 and each of its latency-bound passes takes a third of the time.
template <int ITEMS>
__global__ void __launch_bounds__(kRxThreads) radix_hist_kernel(const uint64_t* __restrict__ keys, size_t n, int shift,
                                                                uint32_t* __restrict__ table, int n_tiles, RxBatch rb) {
  if (rb.n_arr) { n = (size_t)rb.n_arr[blockIdx.y]; keys += blockIdx.y * rb.key_stride; table += blockIdx.y * rb.table_stride; }
  __shared__ uint32_t h[256];
  h[threadIdx.x] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const size_t base = (size_t)blockIdx.x * (kRxThreads * ITEMS);
#pragma unroll 4
  for (int it = 0; it < ITEMS; ++it) {
    const size_t i = base + (size_t)it * kRxThreads + threadIdx.x;
    const bool ok = i < n;
    const unsigned d = ok ? (unsigned)((keys[i] >> shift) & 255ull) : (256u + lane);
    const unsigned m = __match_any_sync(0xffffffffu, d);
    if (ok && lane == __ffs(m) - 1) atomicAdd(&h[d], (uint32_t)__popc(m));
  }
  __syncthreads();
  table[(size_t)threadIdx.x * n_tiles + blockIdx.x] = h[threadIdx.x];
}

// one CTA per digit: sum of its table row
__global__ void __launch_bounds__(256) radix_rowsum_kernel(const uint32_t* __restrict__ table, int n_tiles,
                                                           unsigned long long* __restrict__ totals, RxBatch rb) {
  if (rb.n_arr) { table += blockIdx.y * rb.table_stride; totals = reinterpret_cast<unsigned long long*>(reinterpret_cast<uint32_t*>(totals) + blockIdx.y * rb.table_stride); }
  __shared__ unsigned long long red[8];
  const uint32_t* row = table + (size_t)blockIdx.x * n_tiles;
  unsigned long long s = 0;
  for (int i = threadIdx.x; i < n_tiles; i += 256) s += row[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
    for (int w = 0; w < 8; ++w) t += red[w];
    totals[blockIdx.x] = t;
  }
}

// one CTA per digit: base = sum of the totals of smaller digits; exclusive scan of the row in place.
// Offsets fit 32 bits (n < 2^32).
__global__ void __launch_bounds__(256) radix_rowscan_kernel(uint32_t* __restrict__ table, int n_tiles,
                                                            const unsigned long long* __restrict__ totals, RxBatch rb) {
  if (rb.n_arr) { table += blockIdx.y * rb.table_stride; totals = reinterpret_cast<const unsigned long long*>(reinterpret_cast<const uint32_t*>(totals) + blockIdx.y * rb.table_stride); }
  __shared__ unsigned long long red[8];
  __shared__ uint32_t wsum[8];
  __shared__ uint32_t carry_s;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned long long b = (threadIdx.x < (int)blockIdx.x) ? totals[threadIdx.x] : 0ull;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) b += __shfl_xor_sync(0xffffffffu, b, o);
  if (lane == 0) red[warp] = b;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long t = 0;
    for (int w = 0; w < 8; ++w) t += red[w];
    carry_s = (uint32_t)t;
  }
  __syncthreads();
  uint32_t* row = table + (size_t)blockIdx.x * n_tiles;
  for (int i0 = 0; i0 < n_tiles; i0 += 256) {
    const int i = i0 + threadIdx.x;
    const uint32_t v = i < n_tiles ? row[i] : 0u;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    uint32_t wb = 0;
    for (int w = 0; w < warp; ++w) wb += wsum[w];
    const uint32_t carry = carry_s;
    if (i < n_tiles) row[i] = carry + wb + x - v;
    __syncthreads();
    if (threadIdx.x == 255) carry_s = carry + wb + x;
    __syncthreads();
  }
}

// Small tables (256 * n_tiles <= 32768 entries: sorts below 128 tiles, e.g. the 56 k points of a depth frame): the exclusive scan in
// (digit, tile) order is a scan of the table as one linear array, done by ONE CTA in one launch instead of the row-sum and row-scan
// pair -- at this size every launch of the sort is latency, not bandwidth.
constexpr int kRxScanThreads = 1024;
constexpr int kRxScanMaxTiles = 128;
__global__ void __launch_bounds__(kRxScanThreads) radix_tablescan_kernel(uint32_t* __restrict__ table, int n_tiles, RxBatch rb) {
  if (rb.n_arr) table += blockIdx.y * rb.table_stride;
  extern __shared__ uint32_t rx_tab[];   // the whole table: 256 * n_tiles words (<= 128 KB)
  __shared__ uint32_t wsum[32];
  const int total = 256 * n_tiles;
  // coalesced copy in, every load of a thread independent of the others (one memory latency for the lot)
  for (int e0 = 0; e0 < total; e0 += 8 * kRxScanThreads) {
    uint32_t v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) { const int e = e0 + u * kRxScanThreads + threadIdx.x; v[u] = e < total ? table[e] : 0u; }
#pragma unroll
    for (int u = 0; u < 8; ++u) { const int e = e0 + u * kRxScanThreads + threadIdx.x; if (e < total) rx_tab[e] = v[u]; }
  }
  __syncthreads();
  const int L = (total + kRxScanThreads - 1) / kRxScanThreads;   // <= 32 consecutive entries per thread
  const int b0 = threadIdx.x * L;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t sum = 0;
  for (int k = 0; k < L; ++k) if (b0 + k < total) sum += rx_tab[b0 + k];
  uint32_t x = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
  if (lane == 31) wsum[warp] = x;
  __syncthreads();
  if (warp == 0) {
    const uint32_t w = wsum[lane];
    uint32_t xi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, xi, o); if (lane >= o) xi += y; }
    wsum[lane] = xi - w;
  }
  __syncthreads();
  uint32_t run = wsum[warp] + x - sum;
  for (int k = 0; k < L; ++k)
    if (b0 + k < total) { const uint32_t v = rx_tab[b0 + k]; rx_tab[b0 + k] = run; run += v; }
  __syncthreads();
  for (int e = threadIdx.x; e < total; e += kRxScanThreads) table[e] = rx_tab[e];
}

template <int ITEMS>
__global__ void __launch_bounds__(kRxThreads) radix_scatter_kernel(const uint64_t* __restrict__ kin, const uint32_t* __restrict__ vin,
                                                                   uint64_t* __restrict__ kout, uint32_t* __restrict__ vout,
                                                                   size_t n, int shift, const uint32_t* __restrict__ table,
                                                                   int n_tiles, RxBatch rb) {
  if (rb.n_arr) {
    n = (size_t)rb.n_arr[blockIdx.y];
    kin += blockIdx.y * rb.key_stride; vin += blockIdx.y * rb.val_stride; kout += blockIdx.y * rb.key_stride; vout += blockIdx.y * rb.val_stride;
    table += blockIdx.y * rb.table_stride;
  }
  __shared__ uint32_t wcnt[kRxWarps][256];
  __shared__ uint32_t gbase[256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int w = 0; w < kRxWarps; ++w) wcnt[w][threadIdx.x] = 0;
  gbase[threadIdx.x] = table[(size_t)threadIdx.x * n_tiles + blockIdx.x];
  __syncthreads();
  // item order inside the tile = memory order: index = base + warp * (32 * ITEMS) + it * 32 + lane
  const size_t base = (size_t)blockIdx.x * (kRxThreads * ITEMS) + (size_t)warp * (32 * ITEMS) + lane;
  uint64_t k[ITEMS];
  uint32_t v[ITEMS];
  uint32_t rk[ITEMS];
#pragma unroll
  for (int it = 0; it < ITEMS; ++it) {
    const size_t i = base + (size_t)it * 32;
    k[it] = i < n ? kin[i] : ~0ull;
    v[it] = i < n ? vin[i] : 0u;
  }
#pragma unroll
  for (int it = 0; it < ITEMS; ++it) {
    const size_t i = base + (size_t)it * 32;
    const bool ok = i < n;
    const unsigned d = ok ? (unsigned)((k[it] >> shift) & 255ull) : (256u + lane);
    const unsigned m = __match_any_sync(0xffffffffu, d);
    const uint32_t prior = ok ? wcnt[warp][d] : 0u;
    __syncwarp();
    if (ok && lane == __ffs(m) - 1) wcnt[warp][d] = prior + (uint32_t)__popc(m);
    __syncwarp();
    rk[it] = prior + (uint32_t)__popc(m & ((1u << lane) - 1u));
  }
  __syncthreads();
  {  // exclusive scan over the warps, one digit per thread
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < kRxWarps; ++w) { const uint32_t t = wcnt[w][threadIdx.x]; wcnt[w][threadIdx.x] = run; run += t; }
  }
  __syncthreads();
#pragma unroll
  for (int it = 0; it < ITEMS; ++it) {
    const size_t i = base + (size_t)it * 32;
    if (i < n) {
      const unsigned d = (unsigned)((k[it] >> shift) & 255ull);
      const size_t pos = (size_t)gbase[d] + wcnt[warp][d] + rk[it];
      kout[pos] = k[it];
      vout[pos] = v[it];
    }
  }
}

}  // namespace

constexpr size_t kRxSmallN = (size_t)1 << 18;
constexpr int kRxSmallItems = 4;
inline int rx_tile_items(size_t n) { return kRxThreads * (n < kRxSmallN ? kRxSmallItems : kRxItems); }

size_t radix_table_words(size_t n) {
  const size_t tiles = (n + rx_tile_items(n) - 1) / rx_tile_items(n);
  return 256 * (tiles ? tiles : 1) + 2 * 256;  // table + 256 u64 totals
}

// Sorts n pairs by bits [begin_bit, end_bit) of the key.  Ping-pongs between (keys, vals) and (keys_alt, vals_alt);
// *in_alt tells where the result is.  Returns the number of kernels enqueued.
// Batched: `batch` independent sorts in the same launches -- segment s holds n_arr[s] (device array) <= n items at offset
// s * key_stride of the key arrays and s * val_stride of the payload arrays, its table at s * table_stride words (>= radix_table_words(n), even).
int radix_sort_pairs_batched(uint64_t* keys, uint32_t* vals, uint64_t* keys_alt, uint32_t* vals_alt, uint32_t* table, size_t n,
                             const unsigned long long* n_arr, size_t key_stride, size_t val_stride, size_t table_stride, int batch, int begin_bit, int end_bit,
                             cudaStream_t st, bool* in_alt) {
  *in_alt = false;
  if (n == 0 || batch <= 0) return 0;
  const bool small = n < kRxSmallN;
  const int n_tiles = (int)((n + rx_tile_items(n) - 1) / rx_tile_items(n));
  unsigned long long* totals = reinterpret_cast<unsigned long long*>(table + (size_t)256 * n_tiles);
  RxBatch rb; rb.n_arr = n_arr; rb.key_stride = key_stride; rb.val_stride = val_stride; rb.table_stride = table_stride;
  const int by = n_arr ? batch : 1;
  int launches = 0;
  uint64_t* ki = keys; uint32_t* vi = vals; uint64_t* ko = keys_alt; uint32_t* vo = vals_alt;
  for (int shift = begin_bit; shift < end_bit; shift += 8) {
    if (small) radix_hist_kernel<kRxSmallItems><<<dim3(n_tiles, by), kRxThreads, 0, st>>>(ki, n, shift, table, n_tiles, rb);
    else radix_hist_kernel<kRxItems><<<dim3(n_tiles, by), kRxThreads, 0, st>>>(ki, n, shift, table, n_tiles, rb);
    if (n_tiles <= kRxScanMaxTiles) {
      static PerDeviceOnce attr;
      attr.run([] { return cudaFuncSetAttribute(radix_tablescan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 256 * kRxScanMaxTiles * 4); });
      radix_tablescan_kernel<<<dim3(1, by), kRxScanThreads, (size_t)256 * n_tiles * 4, st>>>(table, n_tiles, rb);
      launches -= 1;
    } else {
      radix_rowsum_kernel<<<dim3(256, by), 256, 0, st>>>(table, n_tiles, totals, rb);
      radix_rowscan_kernel<<<dim3(256, by), 256, 0, st>>>(table, n_tiles, totals, rb);
    }
    if (small) radix_scatter_kernel<kRxSmallItems><<<dim3(n_tiles, by), kRxThreads, 0, st>>>(ki, vi, ko, vo, n, shift, table, n_tiles, rb);
    else radix_scatter_kernel<kRxItems><<<dim3(n_tiles, by), kRxThreads, 0, st>>>(ki, vi, ko, vo, n, shift, table, n_tiles, rb);
    launches += 4;
    uint64_t* tk = ki; ki = ko; ko = tk;
    uint32_t* tv = vi; vi = vo; vo = tv;
    *in_alt = !*in_alt;
  }
  return launches;
}

int radix_sort_pairs(uint64_t* keys, uint32_t* vals, uint64_t* keys_alt, uint32_t* vals_alt, uint32_t* table, size_t n,
                     int begin_bit, int end_bit, cudaStream_t st, bool* in_alt) {
  return radix_sort_pairs_batched(keys, vals, keys_alt, vals_alt, table, n, nullptr, 0, 0, 0, 1, begin_bit, end_bit, st, in_alt);
}

}  // namespace b2r
