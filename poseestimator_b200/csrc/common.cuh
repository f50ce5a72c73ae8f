// Internal declarations shared by the kernels of libb2r.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "../../include/b2r.h"

namespace b2r {

// cudaFuncSetAttribute (dynamic shared memory opt-in) applies to the CURRENT device only: handles on several devices in one
// process each need it once, and a second thread must not launch before the first one has finished setting it.
struct PerDeviceOnce {
  std::mutex m;
  unsigned long long done = 0;   // bit d: set on device d
  template <class F>
  cudaError_t run(F&& f) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> g(m);
    if ((done >> (dev & 63)) & 1ull) return cudaSuccess;
    e = f();
    if (e == cudaSuccess) done |= 1ull << (dev & 63);
    return e;
  }
};

constexpr int kTile = 128;            // consistency-graph tile edge (rows and columns per CTA)
constexpr int kMaxN = 65535;          // degree / id packing in the clique kernels (16 + 16 bits)
constexpr int kMaxCompactWords = 1024;  // register bitsets in the exact search: n' <= 32768

// error / info bits in ProbState::flags
enum : uint32_t {
  FLAG_ARENA_OVERFLOW = 1u << 0,   // exact search ran out of per-warp stack
  FLAG_TIME_LIMIT = 1u << 1,       // exact search stopped by max_clique_time_limit
  FLAG_COMPACT_TOO_LARGE = 1u << 2,  // pruned graph wider than kMaxCompactWords
  FLAG_CANON_FAILED = 1u << 3,     // canonicalisation could not build a clique of size omega (bug guard)
  FLAG_SHARD_FALLBACK = 1u << 5,   // b2r_solve_sharded: some rank's tensor-core filter refused the problem -> every rank solves it alone
  FLAG_SHARD_TIMEOUT = 1u << 6,    // b2r_solve_sharded: a peer's flag did not arrive within the time limit
};

// Per-problem device state: written by the kernels, copied back once per batch.
struct ProbState {
  int n;
  int W;                 // adjacency words per row in use (multiple of 4)
  int max_deg;
  int lb;                // heuristic lower bound (clique size)
  unsigned long long heu_key;   // (size << 32) | (0xffffffff - seed): deterministic heuristic winner
  unsigned long long seed_key[8];  // per id class (v & 7): (deg << 32) | (0xffffffff - v) of its highest-degree vertex
  unsigned long long deg_sum;   // 2 * edges
  int T;                 // prune threshold used for the compact graph
  int n_alive;           // compact graph size n'
  int Wc;                // compact words per row (multiple of 4)
  int omega;
  int clique_size;
  int n_rot_tims, n_rot_inl, n_trans_inl, gnc_iters;
  double gnc_cost;
  double scale, R[9], t[3];
  int valid;
  uint32_t flags;
  unsigned long long bnb_nodes;
  unsigned long long slow_pairs;
  unsigned int root_counter;
  unsigned int canon_next;
  int canon_found;
  int use32;             // graph filter v3: 1 when the packed-fp32 filter may decide pairs of this problem (graph_prep_kernel)
  double ctr[6];         // centre of the bounding boxes of src (0..2) and dst (3..5): the fp32 copies are taken relative to it
  float f32c[8];         // v3 constants: -2 beta^2, beta^4, beta^2, a3, a2, a1, a0 (band polynomial), k (info)
  // tensor-core filter (graph_tc.cu): band T(L) = tcc[1] + tcc[0] * L, tcc[2] = L above which both lengths may be below beta
  float tcc[4];
  double kappa;          // 1 / beta: the 22-bit copies are (x - ctr) * kappa
  unsigned int amb_count;   // pairs appended to ProbDesc::amb_list (may exceed amb_cap: then tc_overflow is set)
  int tc_overflow;       // 1: the list overflowed -> the whole problem is redone by the fp64 filter
  int deg_ready;         // 1: deg[] was accumulated by the tensor-core filter and its fix-up (degree_kernel skips the popcount)
  float tc_pi;           // tensor-core filter: upper bound of |s'|^2 + |d'|^2 over the problem's points (beta units), for the per-row band
  // fused tail (tail.cu)
  int tail_done;         // 1: the fused tail kernel finished this problem (the multi-kernel clique / rotation / translation path skips it)
  int tail_why;          // why the fused tail handed the problem to the multi-kernel path (0 = it did not)
  unsigned int tail_clk[32]; // SM cycles the fused tail spent up to the end of: fix-up, degrees, heuristic, prune, clique, rotation, translation, start;
                             // [8..31]: finer marks inside the stages (b2r_debug_tail_cycles)
};

// cycle marks of the fused tail (test hook): thread 0 of the CTA stores clock64() - t0 into slot k
struct TailMarks {
  unsigned int* clk = nullptr;
  long long t0 = 0;
  __device__ __forceinline__ void mark(int k) const { if (clk && threadIdx.x == 0) clk[k] = (unsigned int)(clock64() - t0); }
};

// Per-problem pointers (device copies of host-built descriptors).
struct ProbDesc {
  const double* src;     // packed 3 x n, column-major
  const double* dst;
  uint32_t* adj;         // n x pitch words
  int* deg;              // n
  int* ids;              // compact id -> original id (n)
  uint32_t* cadj;        // compact adjacency, n' x pitch words (first Wc used)
  int* clique;           // output: original ids ascending (n)
  double* weights;       // GNC weights (tim_cap)
  double* timx;          // pruned TIMs for rotation, 3 x tim_cap
  double* timy;
  unsigned char* rot_mask;   // tim_cap
  int* rot_inl;          // positions with w >= 0.5 (tim_cap)
  unsigned char* trans_mask; // tim_cap
  double* tls_val;       // adaptive voting scratch: 2 * tim_cap endpoints
  int* tls_id;
  double* raw;           // 3 x tim_cap raw translations
  ProbState* st;
  float4* pts32;         // tensor-core filter: pre-swizzled operand rows (column, row-L, row-D: 3 x n_pad x 128 B), graph_tc.cu
  uint32_t* amb_list;    // tensor-core filter: (i << 16) | j of the pairs the exact kernel must decide
  int amb_cap;
  int pad2_;
  int n;
  int pitch;             // words per adjacency row (multiple of 4)
  int tim_cap;
  int pad_;
};

struct KernelParams {
  double beta;           // 2 * noise_bound * sqrt(cbar2)
  double half_beta2;     // 0.5 * beta^2
  int e_beta2;           // biased fp64 exponent of beta^2
  int filter_shift;      // widens the ambiguity band of the sqrt-free filter by 2^shift (tests)
  double noise_bound, cbar2;
  double rot_noise_bound;
  double gnc_factor, cost_threshold;
  int max_iterations;
  int tim_graph;
  int inlier_mode;
  int estimate_scaling;
  double kcore_threshold;
  unsigned long long time_limit_ns;
  int exact_only;
  int graph_v1;           // 1: the first filter kernel (one column per thread), kept for A/B measurements
  int graph_v2;           // 1: the fp64 filter with two columns per thread (v2) instead of the packed-fp32 filter (v3)
  int graph_v3;           // 1: the packed-fp32 filter on coordinate differences (v3) instead of the norm-expansion filter (v4)
  int graph_v4;           // 1: the packed-fp32 norm-expansion filter (v4) instead of the tensor-core filter (v5)
  double filter_tau_min;  // graph_filter_tau of the params (0 = none): lower limit of v3's band tau_G (tests widen it)
  unsigned band_k1, band_k2;  // high-word constants of the v2 ambiguity band (api.cu fill_kernel_params)
  // One problem tile-sharded over the GPUs of a box (b2r_solve_sharded): this rank evaluates the work items i with
  // i % shard_world == shard_rank and stores every adjacency word it produces into its own matrix AND, over NVLink, into the
  // matrices of the n_peer other ranks (same offsets).  shard_world <= 1: not sharded.
  int shard_rank, shard_world, n_peer;
  uint32_t* peer_adj[7];
};

// launchers (each returns the number of kernels it enqueued)
int launch_tim_graph(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, cudaStream_t st);
// graph_tc.cu (v5): prep (boxes, constants, 22-bit records) + tcgen05 filter; graph.cu: exact fix-up of the listed pairs and the
// fp64 filter for problems the fp32 path may not decide
int launch_tim_graph_tc(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, cudaStream_t st);
// sharded form (kp.shard_world > 1, batch 1): the filter kernel with peer stores, then the exact fix-up of this rank's listed pairs
// on every copy of the matrix (bits only; the degrees are recomputed from the complete matrix)
int launch_tim_graph_tc_sharded(const ProbDesc* d_desc, int n, const KernelParams& kp, cudaStream_t st);
int launch_tim_graph_tc_and_fix(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, cudaStream_t st);
int launch_graph_tc_debug(const ProbDesc* d_desc, int prob, int bi, int v, float* out_dm, float* out_l, cudaStream_t st);
// every unit of one problem checked on the device against fp64 (b2r_debug_tc_sweep); out: 8 doubles, zeroed by the caller
int launch_graph_tc_sweep(const ProbDesc* d_desc, int prob, int n, const KernelParams& kp, double* out, cudaStream_t st);
int launch_graph_fix(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, cudaStream_t st);
// listed pairs per problem: about 1.4e-4 of the pairs are ambiguous at config 3; room for eight times that, at least 32768
inline size_t amb_cap_for(size_t n_cap) { const size_t c = n_cap * n_cap / 1024; return c < ((size_t)1 << 15) ? ((size_t)1 << 15) : c; }
int launch_tim_graph_rows(const double* src, const double* dst, int n, int row_begin, int row_end, uint32_t* adj,
                          int pitch, const KernelParams& kp, unsigned long long* slow_pairs, cudaStream_t st);
int launch_degree(const ProbDesc* d_desc, int batch, int n_max, cudaStream_t st);
// tail.cu: everything after the graph in one launch, one CTA per problem (mode 0: tensor-core graph of this batch, 1: final graph + degree_kernel)
int launch_tail(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, int mode, cudaStream_t st);
int launch_clique(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, uint32_t* arena,
                  size_t arena_words_per_warp, int bnb_blocks_per_problem, cudaStream_t st, cudaEvent_t* ev /*4 or null*/,
                  bool bound_known = false /*degrees, seeds and greedy bound of every open problem are the fused tail's*/);
int launch_rotation_translation(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, cudaStream_t st,
                                cudaEvent_t* ev /*1 or null*/);

// radix.cu: stable LSD sort of (u64 key, u32 payload) pairs on bits [begin_bit, end_bit); result in the alt buffers
// when *in_alt.  table: radix_table_words(n) u32 words.
size_t radix_table_words(size_t n);
int radix_sort_pairs(uint64_t* keys, uint32_t* vals, uint64_t* keys_alt, uint32_t* vals_alt, uint32_t* table, size_t n,
                     int begin_bit, int end_bit, cudaStream_t st, bool* in_alt);
// scale.cu: adaptive-voting scale estimate of one problem into st->scale
int radix_sort_pairs_batched(uint64_t* keys, uint32_t* vals, uint64_t* keys_alt, uint32_t* vals_alt, uint32_t* table, size_t n,
                             const unsigned long long* n_arr, size_t key_stride, size_t val_stride, size_t table_stride, int batch, int begin_bit,
                             int end_bit, cudaStream_t st, bool* in_alt);
size_t scale_workspace_bytes(int n);
// batch <= scale_max_chunk() problems per call; ws: batch * scale_workspace_bytes(n_max) bytes
int scale_max_chunk();
int launch_scale_estimate_batch(const ProbDesc* dd, int batch, int n_max, double beta, void* ws, cudaStream_t stream);

constexpr int kBnbWarpsPerBlock = 8;
constexpr int kCanonWarps = 8;

}  // namespace b2r
