// C ABI of libb2r.so (include/b2r.h): handle, workspace, batching, result getters.
// Host-side orchestration only; every stage is a kernel in graph.cu / clique.cu / rt.cu.
#if defined(__x86_64__) || defined(_M_X64)
#include <emmintrin.h>
#endif
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <string>
#include <thread>
#include <vector>

#include "common.cuh"
#include "../../include/b2r_debug.h"

using namespace b2r;

#include "solver.cuh"

namespace {

__global__ void gather_points_kernel(const double* __restrict__ in, long long rs, long long cs, int n, double* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 3 * n) out[i] = in[(long long)(i % 3) * rs + (long long)(i / 3) * cs];
}

}  // namespace

namespace {

// Staging copy with non-temporal stores.  Measured on the B200 hosts (tools/peaks/h2d_pack.cu, profiles/r01_h2d_pack.jsonl): a
// pinned block freshly written with ordinary stores by 8 threads is DMA-read at 12 GB/s (the lines are dirty in eight cores'
// caches), written with streaming stores at 54 GB/s -- 1.2 ms vs 0.28 ms for the 15 MB of a 64 x 5000 batch.
void nt_memcpy(void* dst_, const void* src_, size_t n) {
#if defined(__x86_64__) || defined(_M_X64)
  char* dst = (char*)dst_;
  const char* src = (const char*)src_;
  size_t i = 0;
  for (; i < n && ((uintptr_t)(dst + i) & 15); ++i) dst[i] = src[i];
  for (; i + 64 <= n; i += 64) {
    const __m128i a = _mm_loadu_si128((const __m128i*)(src + i)), b = _mm_loadu_si128((const __m128i*)(src + i + 16));
    const __m128i c = _mm_loadu_si128((const __m128i*)(src + i + 32)), d = _mm_loadu_si128((const __m128i*)(src + i + 48));
    _mm_stream_si128((__m128i*)(dst + i), a); _mm_stream_si128((__m128i*)(dst + i + 16), b);
    _mm_stream_si128((__m128i*)(dst + i + 32), c); _mm_stream_si128((__m128i*)(dst + i + 48), d);
  }
  for (; i < n; ++i) dst[i] = src[i];
  _mm_sfence();
#else
  memcpy(dst_, src_, n);
#endif
}

void fill_kernel_params(b2r_solver* s) {
  const b2r_params& p = s->params;
  KernelParams& k = s->kp;
  k.beta = 2 * p.noise_bound * sqrt(p.cbar2);  // (2*nb)*sqrt(cbar2), A.4-2
  const double beta2 = k.beta * k.beta;
  k.half_beta2 = 0.5 * beta2;
  {
    unsigned long long bits;
    memcpy(&bits, &beta2, sizeof(bits));
    k.e_beta2 = (int)((bits >> 52) & 0x7ff);
    // default band: |F| < 2^-24..2^-28 of Sigma^2; graph_filter_tau widens it (tests force the exact path with it)
    int shift = 0;
    if (p.reserved1 > 0) { shift = (int)ceil(log2(p.reserved1) + 28.0); if (shift < 0) shift = 0; if (shift > 64) shift = 64; }
    k.filter_shift = shift;
  }
  k.noise_bound = p.noise_bound; k.cbar2 = p.cbar2;
  k.rot_noise_bound = p.rotation_noise_bound;
  k.gnc_factor = p.rotation_gnc_factor; k.cost_threshold = p.rotation_cost_threshold;
  k.max_iterations = p.rotation_max_iterations;
  k.tim_graph = p.rotation_tim_graph;
  k.inlier_mode = p.inlier_selection_mode;
  k.estimate_scaling = p.estimate_scaling;
  k.kcore_threshold = p.kcore_heuristic_threshold;
  const double lim = p.max_clique_time_limit;
  k.time_limit_ns = (lim > 0 && lim < 1e9) ? (unsigned long long)(lim * 1e9) : 0ull;
  k.exact_only = p.reserved0 == 1;
  k.filter_tau_min = p.reserved1 > 0 ? p.reserved1 : 0.0;
  {
    // filter generation: b2r_debug_set_graph_filter(3 / 2 / 4 / 5) or B2R_GRAPH_V = 1 / 2 / 3 / 4 select the older filters v1 / v2 (fp64) / v3 / v4 (A/B runs, tests)
    const char* gv = getenv("B2R_GRAPH_V");
    const int v = p.reserved0 == 3 ? 1 : p.reserved0 == 2 ? 2 : p.reserved0 == 4 ? 3 : p.reserved0 == 5 ? 4
                  : (gv && gv[0] >= '1' && gv[0] <= '5') ? gv[0] - '0' : 5;
    k.graph_v1 = v == 1;
    k.graph_v2 = v == 2;
    k.graph_v3 = v == 3;
    k.graph_v4 = v == 4;
    // v2 band: |F| < 2^(max(2 e(H) - 1023, e(beta^4)) - 22 + shift), compared on the high words of H and F
    const double beta4 = beta2 * beta2;
    unsigned long long bits;
    memcpy(&bits, &beta4, sizeof(bits));
    long long e4 = (long long)((bits >> 52) & 0x7ff);
    long long k1 = 1045 - (long long)k.filter_shift;
    if (k1 < 0) k1 = 0;
    long long k2 = e4 + 1023;
    if (k2 < k1) k2 = k1;
    if (k2 > 4095) k2 = 4095;
    k.band_k1 = (unsigned)(k1 << 20);
    k.band_k2 = (unsigned)(k2 << 20);
  }
}

b2r_status check_params(b2r_solver* s, const b2r_params* p) {
  if (!p) return fail(s, B2R_ERR_INVALID_ARG, "params is NULL");
  if (!(p->noise_bound != 0)) return fail(s, B2R_ERR_INVALID_ARG, "noise_bound must be non-zero");
  if (!(p->rotation_gnc_factor > 1)) return fail(s, B2R_ERR_INVALID_ARG, "rotation_gnc_factor must be > 1");
  if (p->rotation_estimation_algorithm != B2R_ROT_GNC_TLS)
    return fail(s, B2R_ERR_INVALID_ARG, "only GNC_TLS rotation estimation is implemented (FGR is not on the reference's path)");
  if (p->inlier_selection_mode < 0 || p->inlier_selection_mode > 3) return fail(s, B2R_ERR_INVALID_ARG, "bad inlier_selection_mode");
  if (p->rotation_tim_graph < 0 || p->rotation_tim_graph > 1) return fail(s, B2R_ERR_INVALID_ARG, "bad rotation_tim_graph");
  if (p->reserved0 != 0 || p->reserved1 != 0) return fail(s, B2R_ERR_INVALID_ARG, "reserved fields of b2r_params must be 0");
  return B2R_OK;
}

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

b2r_status ensure_workspace(b2r_solver* s, int batch, int n_max) {
  const int pitch = b2r_pitch_words(n_max);
  long long tim_cap = n_max;
  if (s->params.rotation_tim_graph == B2R_TIM_COMPLETE) {
    const long long full = (long long)n_max * (n_max - 1) / 2;
    const long long lim = 1ll << 22;
    tim_cap = full < lim ? (full > n_max ? full : n_max) : lim;
  }
  if (batch <= s->batch_cap && n_max <= s->n_cap && tim_cap <= s->tim_cap) return B2R_OK;
  const int bc = batch > s->batch_cap ? batch : s->batch_cap;
  const int nc = n_max > s->n_cap ? n_max : s->n_cap;
  const int pc = b2r_pitch_words(nc);
  const long long tc = tim_cap > s->tim_cap ? tim_cap : s->tim_cap;
  (void)pitch;
  CK(cudaStreamSynchronize(s->stream));
  const size_t B = (size_t)bc;
  s->bnb_blocks = (int)(296 / B);
  if (s->bnb_blocks < 1) s->bnb_blocks = 1;
  if (s->bnb_blocks > 37) s->bnb_blocks = 37;
  const char* ak = getenv("B2R_ARENA_KB");
  if (ak && atoi(ak) > 0) s->arena_words_per_warp = (size_t)atoi(ak) * 256;
  // Not failure-atomic buffer by buffer (DevBuf::ensure frees before it allocates), so a failure anywhere drops the WHOLE
  // workspace and its recorded capacity: the next call re-allocates from scratch instead of trusting half-grown buffers.
  auto grow = [&]() -> cudaError_t {
    cudaError_t e;
#define B2R_TRY(x) if ((e = (x)) != cudaSuccess) return e
    B2R_TRY(s->d_pts.ensure(B * 6 * nc * sizeof(double)));
    B2R_TRY(s->d_adj.ensure(B * nc * pc * sizeof(uint32_t)));
    B2R_TRY(s->d_cadj.ensure(B * nc * pc * sizeof(uint32_t)));
    B2R_TRY(s->d_pts32.ensure(B * (size_t)(pc * 32) * 128));   // tensor-core filter: one 128-byte operand row per point, pc * 32 = n_cap rounded up to the tile edge
    B2R_TRY(s->d_amb.ensure(B * amb_cap_for(nc) * sizeof(uint32_t)));
    // ints per problem: deg, ids, clique (n each) + rot_inl (tim) + tls_id (4*tim)
    B2R_TRY(s->d_int.ensure(B * (3 * (size_t)nc + 5 * (size_t)tc) * sizeof(int)));
    // doubles per problem: weights (tim), timx, timy, raw (3*tim each), tls_val (4*tim)
    B2R_TRY(s->d_dbl.ensure(B * (14 * (size_t)tc) * sizeof(double)));
    B2R_TRY(s->d_u8.ensure(B * 2 * (size_t)tc));
    B2R_TRY(s->d_state.ensure(B * sizeof(ProbState)));
    B2R_TRY(s->d_desc.ensure(B * sizeof(ProbDesc)));
    B2R_TRY(s->h_state.ensure(B * sizeof(ProbState)));
    B2R_TRY(s->h_desc.ensure(B * sizeof(ProbDesc)));
    B2R_TRY(s->h_pts.ensure(B * 6 * nc * sizeof(double)));
    B2R_TRY(s->d_arena.ensure(B * s->bnb_blocks * kBnbWarpsPerBlock * s->arena_words_per_warp * sizeof(uint32_t)));
#undef B2R_TRY
    return cudaSuccess;
  };
  const cudaError_t ge = grow();
  if (ge != cudaSuccess) {
    cudaGetLastError();   // the failed cudaMalloc must not surface in a later cudaGetLastError()
    s->d_pts.release(); s->d_adj.release(); s->d_cadj.release(); s->d_pts32.release(); s->d_amb.release(); s->d_int.release();
    s->d_dbl.release(); s->d_u8.release(); s->d_state.release(); s->d_desc.release(); s->d_arena.release();
    s->h_state.release(); s->h_desc.release(); s->h_pts.release();
    s->batch_cap = 0; s->n_cap = 0; s->pitch = 0; s->tim_cap = 0;
    s->have_result = false; s->batch = 0;
    s->err = std::string("workspace allocation: ") + cudaGetErrorString(ge);
    return ge == cudaErrorMemoryAllocation ? B2R_ERR_OOM : B2R_ERR_CUDA;
  }
  s->batch_cap = bc; s->n_cap = nc; s->pitch = pc; s->tim_cap = tc;
  return B2R_OK;
}

void build_desc(b2r_solver* s, int b, int n, const double* dsrc, const double* ddst, const uint32_t* ext_adj, int ext_pitch) {
  ProbDesc& d = ((ProbDesc*)s->h_desc.p)[b];
  const size_t nc = (size_t)s->n_cap, tc = (size_t)s->tim_cap, pc = (size_t)s->pitch;
  d.src = dsrc; d.dst = ddst;
  d.adj = ext_adj ? const_cast<uint32_t*>(ext_adj) : (uint32_t*)s->d_adj.p + (size_t)b * nc * pc;
  d.cadj = (uint32_t*)s->d_cadj.p + (size_t)b * nc * pc;
  int* ib = (int*)s->d_int.p + (size_t)b * (3 * nc + 5 * tc);
  d.deg = ib; d.ids = ib + nc; d.clique = ib + 2 * nc; d.rot_inl = ib + 3 * nc; d.tls_id = ib + 3 * nc + tc;
  double* db = (double*)s->d_dbl.p + (size_t)b * 14 * tc;
  d.weights = db; d.timx = db + tc; d.timy = db + 4 * tc; d.raw = db + 7 * tc; d.tls_val = db + 10 * tc;
  unsigned char* ub = (unsigned char*)s->d_u8.p + (size_t)b * 2 * tc;
  d.rot_mask = ub; d.trans_mask = ub + tc;
  d.st = (ProbState*)s->d_state.p + b;
  d.pts32 = (float4*)((unsigned char*)s->d_pts32.p + (size_t)b * pc * 32 * 128);
  d.amb_list = (uint32_t*)s->d_amb.p + (size_t)b * amb_cap_for(nc);
  d.amb_cap = (int)amb_cap_for(nc);
  d.pad2_ = 0;
  d.n = n;
  d.pitch = ext_adj ? ext_pitch : (int)pc;
  d.tim_cap = (int)tc;
  d.pad_ = 0;
}

b2r_status ensure_events(b2r_solver* s) {
  if (s->ev_ready) return B2R_OK;
  for (int i = 0; i < 12; ++i) CK(cudaEventCreate(&s->ev[i]));
  s->ev_ready = true;
  return B2R_OK;
}

// The multi-kernel clique / rotation / translation chain (one launch per stage).  Every kernel skips problems the fused tail
// finished (ProbState::tail_done), so the same chain serves as the whole tail and as the hand-over path.
// bound_known: every open problem was handed over by the fused tail with its degrees, seeds and greedy bound in place (tail_why == 2):
// the chain starts at the core peel instead of computing them again (0.02 ms and two launches of a config-2 chain solve)
b2r_status enqueue_chain(b2r_solver* s, const ProbDesc* dd, int batch, int n_max, cudaStream_t st, bool prof, bool bound_known = false) {
  B2R_NVTX("b2r:clique_rotation_translation_chain");
  if (!bound_known) s->launches += launch_degree(dd, batch, n_max, st);
  if (prof) CK(cudaEventRecord(s->ev[3], st));
  s->launches += launch_clique(dd, batch, n_max, s->kp, (uint32_t*)s->d_arena.p, s->arena_words_per_warp, s->bnb_blocks, st,
                               prof ? &s->ev[4] : nullptr, bound_known);
  s->launches += launch_rotation_translation(dd, batch, n_max, s->kp, st, prof ? &s->ev[8] : nullptr);
  return B2R_OK;
}

bool fused_tail_enabled(const b2r_solver* s) {
  static const bool off = [] { const char* e = getenv("B2R_NO_FUSED_TAIL"); return e && e[0] == '1'; }();
  return !off && s->kp.inlier_mode == B2R_INLIER_PMC_EXACT;
}

// enqueue stages 1.. on descriptors already in h_desc
b2r_status enqueue_pipeline(b2r_solver* s, int batch, int n_max, bool skip_graph, bool prepared = false) {
  B2R_NVTX("b2r:solve_pipeline");
  cudaStream_t st = s->stream;
  const bool prof = s->profiling;
  if (!prepared) {   // (the sharded solve clears the state and uploads the descriptor before its graph kernels, shard.cu)
    CK(cudaMemsetAsync(s->d_state.p, 0, (size_t)batch * sizeof(ProbState), st));
    CK(cudaMemcpyAsync(s->d_desc.p, s->h_desc.p, (size_t)batch * sizeof(ProbDesc), cudaMemcpyHostToDevice, st));
  }
  const ProbDesc* dd = (const ProbDesc*)s->d_desc.p;
  if (s->kp.estimate_scaling) {  // A.4-5: the scale vote comes first; the graph is its inlier mask
    B2R_NVTX("b2r:scale_vote");
    if (prof) CK(cudaEventRecord(s->ev[10], st));
    // all problems of a chunk in one set of launches; the chunk is what fits a 24 GB scratch (N = 5000: 0.8 GB per problem)
    const size_t per = scale_workspace_bytes(n_max);
    int chunk = (int)(((size_t)24 << 30) / per);
    if (chunk < 1) chunk = 1;
    if (chunk > batch) chunk = batch;
    if (chunk > scale_max_chunk()) chunk = scale_max_chunk();
    CK(s->d_scale.ensure((size_t)chunk * per));
    for (int b0 = 0; b0 < batch; b0 += chunk)
      s->launches += launch_scale_estimate_batch(dd + b0, (batch - b0) < chunk ? (batch - b0) : chunk, n_max, s->kp.beta, s->d_scale.p, st);
    if (prof) CK(cudaEventRecord(s->ev[11], st));
  }
  const bool fused = fused_tail_enabled(s);
  // does the graph come from the tensor-core filter (pair list + degrees accumulated on the fly)?
  const bool tc = !skip_graph && !s->kp.estimate_scaling && !s->kp.exact_only && !s->kp.graph_v1 && !s->kp.graph_v2 && !s->kp.graph_v3 &&
                  !s->kp.graph_v4;
  if (prof) CK(cudaEventRecord(s->ev[1], st));
  if (!skip_graph) {
    B2R_NVTX("b2r:consistency_graph");
    if (tc && fused) s->launches += launch_tim_graph_tc(dd, batch, n_max, s->kp, st);   // the tail decides the listed pairs itself
    else s->launches += launch_tim_graph(dd, batch, n_max, s->kp, st);
  }
  if (prof) CK(cudaEventRecord(s->ev[2], st));
  if (s->stream_hi) {  // the tail continues on the high-priority stream
    CK(cudaEventRecord(s->ev_graph, st));
    st = s->stream_hi;
    CK(cudaStreamWaitEvent(st, s->ev_graph, 0));
  }
  s->fused = fused;
  s->handed_over = 0;
  s->n_max = n_max;
  if (fused) {
    B2R_NVTX("b2r:fused_tail(clique+rotation+translation)");
    if (!tc) s->launches += launch_degree(dd, batch, n_max, st);
    s->launches += launch_tail(dd, batch, n_max, s->kp, tc ? 0 : 1, st);
  } else {
    b2r_status r = enqueue_chain(s, dd, batch, n_max, st, prof);
    if (r != B2R_OK) return r;
  }
  if (prof) CK(cudaEventRecord(s->ev[9], st));
  CK(cudaMemcpyAsync(s->h_state.p, s->d_state.p, (size_t)batch * sizeof(ProbState), cudaMemcpyDeviceToHost, st));
  if (s->stream_hi) {  // the handle's stream (what callers synchronise on) waits for the tail
    CK(cudaEventRecord(s->ev_tail, st));
    CK(cudaStreamWaitEvent(s->stream, s->ev_tail, 0));
  }
  CK(cudaGetLastError());
  s->pending = true;
  s->have_result = false;
  return B2R_OK;
}

// Problems the fused tail handed over (ProbState::tail_done == 0 after the batch): the fp64 graph filter where the graph is not
// final, then the multi-kernel chain, which skips every finished problem.  Called from b2r_sync with the stream idle.
b2r_status run_handed_over(b2r_solver* s) {
  const ProbState* hs = (const ProbState*)s->h_state.p;
  int todo = 0, regraph = 0;
  for (int b = 0; b < s->batch; ++b) {
    if (!hs[b].tail_done) { ++todo; if (hs[b].tail_why != 2) ++regraph; }
  }
  s->handed_over = todo;
  if (!todo) return B2R_OK;
  const ProbDesc* dd = (const ProbDesc*)s->d_desc.p;
  // on the high-priority stream: these are small kernels that would otherwise queue behind the graph kernels of other handles
  cudaStream_t st = s->stream_hi ? s->stream_hi : s->stream;
  if (regraph) s->launches += launch_graph_fix(dd, s->batch, s->n_max, s->kp, st);
  const bool prof = s->profiling && s->ev_ready;
  if (prof) CK(cudaEventRecord(s->ev[2], st));
  b2r_status r = enqueue_chain(s, dd, s->batch, s->n_max, st, prof, regraph == 0 && s->kp.inlier_mode == B2R_INLIER_PMC_EXACT);
  if (r != B2R_OK) return r;
  if (prof) CK(cudaEventRecord(s->ev[9], st));
  CK(cudaMemcpyAsync(s->h_state.p, s->d_state.p, (size_t)s->batch * sizeof(ProbState), cudaMemcpyDeviceToHost, st));
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
  // stage times of the handed-over chain: [2] .. [8] (the fused tail's own time stays in [11])
  if (prof) for (int i = 2; i < 9; ++i) cudaEventElapsedTime(&s->timings[i], s->ev[i], s->ev[i + 1]);
  return B2R_OK;
}

// B2R_CANARY=1: guard zones around every workspace buffer (solver.cuh)
const char* broken_canary(const b2r_solver* s) {
  if (!canary_bytes()) return nullptr;
  const struct { const DevBuf* b; const char* name; } all[] = {
      {&s->d_pts, "points"}, {&s->d_adj, "adjacency"}, {&s->d_cadj, "compact adjacency"}, {&s->d_int, "int arrays"}, {&s->d_dbl, "double arrays"},
      {&s->d_u8, "masks"}, {&s->d_state, "problem state"}, {&s->d_desc, "descriptors"}, {&s->d_arena, "search arena"}, {&s->d_scale, "scale vote"},
      {&s->d_pre, "pre-steps"}, {&s->d_nn, "nn gemm"}, {&s->d_grid, "knn grid"}, {&s->d_refine, "refine"}, {&s->d_pts32, "tc operand rows"},
      {&s->d_amb, "ambiguous pairs"}};
  for (const auto& e : all) if (!e.b->intact()) return e.name;
  return nullptr;
}

b2r_status check_index(b2r_solver* s, int idx) {
  if (!s) return B2R_ERR_INVALID_ARG;
  if (s->pending) { b2r_status r = b2r_sync(s); if (r != B2R_OK) return r; }
  if (!s->have_result) return fail(s, B2R_ERR_STATE, "no solve has completed on this handle");
  if (idx < 0 || idx >= s->batch) return fail(s, B2R_ERR_STATE, "problem index out of range");
  return B2R_OK;
}

}  // namespace

// ---- used by shard.cu ------------------------------------------------------------------------------------------------
// workspace for one problem of n correspondences; with operands: state cleared and the descriptor of problem 0 on the device
namespace {
__global__ void __launch_bounds__(32) publish_words_kernel(const unsigned* __restrict__ src, int n, volatile unsigned* mbox, unsigned seq) {
  const int t = threadIdx.x;
  if (t < n) mbox[2 + t] = src[t];
  __threadfence_system();
  __syncwarp();
  if (t == 0) mbox[0] = seq;
}
}  // namespace

b2r_status b2r_internal_read_words(b2r_solver* s, const void* dev_words, int n_words, void* host_out) {
  if (n_words < 1 || n_words > 30) return fail(s, B2R_ERR_INVALID_ARG, "internal: read_words size");
  static const bool off = [] { const char* e = getenv("B2R_NO_MAILBOX"); return e && *e && *e != '0'; }();
  if (off) {   // the plain form, for A/B measurements
    CK(cudaMemcpyAsync(host_out, dev_words, 4 * (size_t)n_words, cudaMemcpyDeviceToHost, s->stream));
    CK(cudaStreamSynchronize(s->stream));
    return B2R_OK;
  }
  if (!s->mbox_h) {
    void* h = nullptr; void* d = nullptr;
    CK(cudaHostAlloc(&h, 256, cudaHostAllocMapped));
    memset(h, 0, 256);
    cudaError_t e = cudaHostGetDevicePointer(&d, h, 0);
    if (e != cudaSuccess) { cudaFreeHost(h); CK(e); }
    s->mbox_h = static_cast<volatile unsigned*>(h); s->mbox_d = static_cast<unsigned*>(d); s->mbox_seq = 0;
  }
  unsigned seq = ++s->mbox_seq;
  if (seq == 0) seq = ++s->mbox_seq;
  publish_words_kernel<<<1, 32, 0, s->stream>>>(static_cast<const unsigned*>(dev_words), n_words, s->mbox_d, seq);
  ++s->launches;
  CK(cudaGetLastError());
  for (unsigned long long it = 1;; ++it) {
    if (s->mbox_h[0] == seq) break;
    if ((it & 0x3fffull) == 0) {   // a faulted kernel never publishes: ask the stream now and then
      const cudaError_t e = cudaStreamQuery(s->stream);
      if (e != cudaSuccess && e != cudaErrorNotReady) CK(e);
    }
    // short waits spin; past ~50 us (the stage before the read is still running) the core is offered to other threads -- several
    // frames in flight per GPU and several ranks per box poll at the same time
    if (it > 1024) std::this_thread::yield();
#if defined(__x86_64__)
    else __builtin_ia32_pause();
#endif
  }
  std::atomic_thread_fence(std::memory_order_acquire);
  unsigned* out = static_cast<unsigned*>(host_out);
  for (int i = 0; i < n_words; ++i) out[i] = s->mbox_h[2 + i];
  return B2R_OK;
}

b2r_status b2r_internal_prepare_single(b2r_solver* s, int n, const double* src_dev, const double* dst_dev) {
  b2r_status r = ensure_workspace(s, 1, n);
  if (r != B2R_OK) return r;
  if (!src_dev) return B2R_OK;
  if (s->profiling) { r = ensure_events(s); if (r != B2R_OK) return r; CK(cudaEventRecord(s->ev[0], s->stream)); }
  s->batch = 1; s->n.assign(1, n);
  s->ext_adj = nullptr;
  build_desc(s, 0, n, src_dev, dst_dev, nullptr, 0);
  CK(cudaMemsetAsync(s->d_state.p, 0, sizeof(ProbState), s->stream));
  CK(cudaMemcpyAsync(s->d_desc.p, s->h_desc.p, sizeof(ProbDesc), cudaMemcpyHostToDevice, s->stream));
  return B2R_OK;
}
// the matrix of problem 0 is complete: degrees, fused tail, state to the host; synchronises
b2r_status b2r_internal_finish_from_graph(b2r_solver* s, int n) {
  b2r_status r = enqueue_pipeline(s, 1, n, true, true);
  if (r != B2R_OK) return r;
  return b2r_sync(s);
}

extern "C" {

void b2r_default_params(b2r_params* p) {
  memset(p, 0, sizeof(*p));
  p->noise_bound = 0.01; p->cbar2 = 1.0; p->estimate_scaling = 1;
  p->rotation_estimation_algorithm = B2R_ROT_GNC_TLS;
  p->rotation_gnc_factor = 1.4; p->rotation_max_iterations = 100; p->rotation_cost_threshold = 1e-6;
  p->rotation_tim_graph = B2R_TIM_CHAIN; p->inlier_selection_mode = B2R_INLIER_PMC_EXACT;
  p->kcore_heuristic_threshold = 0.5; p->max_clique_time_limit = 3000; p->max_clique_num_threads = 0;
}

int b2r_version(void) { return B2R_VERSION; }

int b2r_device_count(void) {
  int c = 0;
  if (cudaGetDeviceCount(&c) != cudaSuccess) { cudaGetLastError(); return 0; }
  return c;
}

int b2r_pitch_words(int n) { return 4 * ((n + kTile - 1) / kTile); }

b2r_status b2r_create(const b2r_params* params, int device, void* stream, b2r_solver** out) {
  if (!out) return B2R_ERR_INVALID_ARG;
  *out = nullptr;
  b2r_solver* s = new b2r_solver();
  b2r_params def;
  if (!params) { b2r_default_params(&def); params = &def; }
  b2r_status r = check_params(s, params);
  if (r != B2R_OK) { delete s; return r; }
  int cnt = 0;
  if (cudaGetDeviceCount(&cnt) != cudaSuccess || cnt <= 0 || device < 0 || device >= cnt) {
    cudaGetLastError();
    delete s;
    return B2R_ERR_CUDA;  // no CPU fallback by design
  }
  s->device = device;
  if (cudaSetDevice(device) != cudaSuccess) { delete s; return B2R_ERR_CUDA; }
  if (stream) { s->stream = (cudaStream_t)stream; s->own_stream = false; }
  else {
    if (cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess) { delete s; return B2R_ERR_CUDA; }
    s->own_stream = true;

  }
  {
  int lo = 0, hi = 0;
  const char* nohi = getenv("B2R_NO_HI_STREAM");
  if (!(nohi && nohi[0] == '1') && cudaDeviceGetStreamPriorityRange(&lo, &hi) == cudaSuccess && hi < lo) {
    if (cudaStreamCreateWithPriority(&s->stream_hi, cudaStreamNonBlocking, hi) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->ev_graph, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s->ev_tail, cudaEventDisableTiming) != cudaSuccess) {
      cudaGetLastError();
      if (s->stream_hi) cudaStreamDestroy(s->stream_hi);
      s->stream_hi = nullptr;
    }
  }
  }

  s->params = *params;
  fill_kernel_params(s);
  *out = s;
  return B2R_OK;
}

b2r_status b2r_destroy(b2r_solver* s) {
  if (!s) return B2R_OK;
  cudaSetDevice(s->device);
  cudaStreamSynchronize(s->stream);
  if (const char* bad = broken_canary(s)) {   // B2R_CANARY=1 only: an out-of-bounds write must not pass silently
    fprintf(stderr, "libb2r: guard zone of the '%s' workspace buffer was overwritten\n", bad);
    abort();
  }
  s->d_pts.release(); s->d_adj.release(); s->d_cadj.release(); s->d_int.release(); s->d_dbl.release(); s->d_u8.release();
  s->d_pts32.release(); s->d_amb.release(); s->d_state.release(); s->d_desc.release(); s->d_arena.release(); s->d_scale.release(); s->d_pre.release(); s->d_nn.release(); s->d_grid.release(); s->d_refine.release(); s->h_refine.release();
  s->h_state.release(); s->h_desc.release(); s->h_pts.release();
  if (s->mbox_h) { cudaFreeHost(const_cast<unsigned*>(s->mbox_h)); s->mbox_h = nullptr; }
  if (s->ev_ready) for (int i = 0; i < 12; ++i) cudaEventDestroy(s->ev[i]);
  if (s->stream_hi) { cudaStreamDestroy(s->stream_hi); if (s->ev_graph) cudaEventDestroy(s->ev_graph); if (s->ev_tail) cudaEventDestroy(s->ev_tail); }
  if (s->own_stream) cudaStreamDestroy(s->stream);
  delete s;
  return B2R_OK;
}

b2r_status b2r_reset(b2r_solver* s, const b2r_params* params) {
  if (!s) return B2R_ERR_INVALID_ARG;
  b2r_status r = check_params(s, params);
  if (r != B2R_OK) return r;
  if (s->pending) { r = b2r_sync(s); if (r != B2R_OK) return r; }
  const int32_t dbg_filter = s->params.reserved0;   // b2r_debug_set_graph_filter survives reset(params)
  const double dbg_tau = s->params.reserved1;
  s->params = *params;
  s->params.reserved0 = dbg_filter; s->params.reserved1 = dbg_tau;
  fill_kernel_params(s);
  s->have_result = false; s->batch = 0;
  return B2R_OK;
}

const char* b2r_last_error(const b2r_solver* s) { return s ? s->err.c_str() : "null handle"; }
int b2r_batch_size(const b2r_solver* s) { return s ? s->batch : 0; }
int64_t b2r_launch_count(const b2r_solver* s) { return s ? s->launches : 0; }

b2r_status b2r_solve_batch_async(b2r_solver* s, int batch, const int* n, const double* const* src,
                                 const double* const* dst, int on_device) {
  if (!s) return B2R_ERR_INVALID_ARG;
  if (batch <= 0 || !n || !src || !dst) return fail(s, B2R_ERR_INVALID_ARG, "batch/n/src/dst");
  B2R_NVTX("b2r:solve_batch_async");
  CK(cudaSetDevice(s->device));
  if (s->pending) { b2r_status r = b2r_sync(s); if (r != B2R_OK) return r; }
  int n_max = 0;
  size_t total = 0;
  for (int b = 0; b < batch; ++b) {
    if (n[b] < 2) return fail(s, B2R_ERR_INVALID_ARG, "every problem needs at least 2 correspondences");
    if (n[b] > kMaxN) return fail(s, B2R_ERR_LIMIT, "N > 65535 correspondences is not supported");
    if (!src[b] || !dst[b]) return fail(s, B2R_ERR_INVALID_ARG, "null point pointer");
    n_max = n[b] > n_max ? n[b] : n_max;
    total += (size_t)3 * n[b];
  }
  if (s->params.estimate_scaling && scale_workspace_bytes(n_max) > (size_t)96 << 30)
    return fail(s, B2R_ERR_LIMIT, "estimate_scaling=True: the 2P-endpoint vote for this N needs more than 96 GB of workspace");
  b2r_status r = ensure_workspace(s, batch, n_max);
  if (r != B2R_OK) return r;
  if (s->profiling) { r = ensure_events(s); if (r != B2R_OK) return r; if (on_device == 1) CK(cudaEventRecord(s->ev[0], s->stream)); }
  s->batch = batch;
  s->n.assign(n, n + batch);
  s->ext_adj = nullptr;
  if (on_device < 0 || on_device > 2) return fail(s, B2R_ERR_INVALID_ARG, "on_device must be 0 (host), 1 (device) or 2 (page-locked host)");
  if (on_device == 1) {
    for (int b = 0; b < batch; ++b) build_desc(s, b, n[b], src[b], dst[b], nullptr, 0);
  } else if (on_device == 2) {
    // page-locked host operands: the copy engine reads the caller's memory directly (no staging pass over the input on the
    // host); operands that follow each other in memory go in one copy
    double* ds = (double*)s->d_pts.p;
    double* dd = ds + total;
    std::vector<size_t> offs((size_t)batch + 1);
    size_t off = 0;
    for (int b = 0; b < batch; ++b) {
      offs[b] = off;
      build_desc(s, b, n[b], ds + off, dd + off, nullptr, 0);
      off += (size_t)3 * n[b];
    }
    offs[batch] = off;
    if (s->profiling) CK(cudaEventRecord(s->ev[0], s->stream));
    for (int op = 0; op < 2; ++op) {
      const double* const* hp = op ? dst : src;
      double* dbase = op ? dd : ds;
      for (int b0 = 0; b0 < batch;) {
        int b1 = b0 + 1;
        while (b1 < batch && hp[b1] == hp[b1 - 1] + (size_t)3 * n[b1 - 1]) ++b1;
        CK(cudaMemcpyAsync(dbase + offs[b0], hp[b0], sizeof(double) * (offs[b1] - offs[b0]), cudaMemcpyHostToDevice, s->stream));
        b0 = b1;
      }
    }
  } else {
    // pack all problems into one pinned staging block -> one H2D copy per operand
    double* hs = (double*)s->h_pts.p;
    double* hd = hs + total;
    double* ds = (double*)s->d_pts.p;
    double* dd = ds + total;
    std::vector<size_t> offs((size_t)batch);
    size_t off = 0;
    for (int b = 0; b < batch; ++b) {
      offs[b] = off;
      build_desc(s, b, n[b], ds + off, dd + off, nullptr, 0);
      off += (size_t)3 * n[b];
    }
    // staging copy: the only host-side work proportional to the input; spread it over a few threads when it is large
    auto pack = [&](int b0, int b1) {
      for (int b = b0; b < b1; ++b) {
        nt_memcpy(hs + offs[b], src[b], sizeof(double) * 3 * n[b]);
        nt_memcpy(hd + offs[b], dst[b], sizeof(double) * 3 * n[b]);
      }
    };
    unsigned nthr = std::thread::hardware_concurrency();
    if (nthr > 8) nthr = 8;
    if (const char* pt = getenv("B2R_PACK_THREADS")) {  // several ranks on one host: cores / ranks, not 8 each
      const int v = atoi(pt);
      if (v >= 1 && v <= 64) nthr = (unsigned)v;
    }
    if ((unsigned)batch < nthr) nthr = (unsigned)batch;
    if (total * sizeof(double) * 2 < (1u << 19) || nthr <= 1) pack(0, batch);
    else {
      const int nt = (int)nthr;
      s->pack_pool.run(nt, [&](int t) { pack((int)((size_t)batch * t / nt), (int)((size_t)batch * (t + 1) / nt)); });   // persistent workers of the handle
    }
    if (s->profiling) CK(cudaEventRecord(s->ev[0], s->stream));   // 'upload' = the H2D copy alone (host packing is host time)
    CK(cudaMemcpyAsync(ds, hs, sizeof(double) * 2 * total, cudaMemcpyHostToDevice, s->stream));
  }
  return enqueue_pipeline(s, batch, n_max, false);
}

b2r_status b2r_alloc_pinned(size_t bytes, void** out) {
  if (!out || bytes == 0) return B2R_ERR_INVALID_ARG;
  *out = nullptr;
  return cudaHostAlloc(out, bytes, cudaHostAllocPortable) == cudaSuccess ? B2R_OK : B2R_ERR_CUDA;
}
b2r_status b2r_copy_to_pinned(void* dst, const void* src, size_t bytes) {
  if ((!dst || !src) && bytes) return B2R_ERR_INVALID_ARG;
  nt_memcpy(dst, src, bytes);
  return B2R_OK;
}
b2r_status b2r_free_pinned(void* p) {
  if (!p) return B2R_OK;
  return cudaFreeHost(p) == cudaSuccess ? B2R_OK : B2R_ERR_CUDA;
}

b2r_status b2r_sync(b2r_solver* s) {
  if (!s) return B2R_ERR_INVALID_ARG;
  B2R_NVTX("b2r:sync");
  CK(cudaSetDevice(s->device));
  CK(cudaStreamSynchronize(s->stream));
  if (!s->pending) return B2R_OK;
  s->pending = false;
  if (s->profiling && s->ev_ready) {
    if (s->fused) {   // one launch after the graph: [2] .. [8] have no events of their own, [11] is the fused tail
      cudaEventElapsedTime(&s->timings[0], s->ev[0], s->ev[1]);
      cudaEventElapsedTime(&s->timings[1], s->ev[1], s->ev[2]);
      for (int i = 2; i < 9; ++i) s->timings[i] = 0.f;
      cudaEventElapsedTime(&s->timings[11], s->ev[2], s->ev[9]);
    } else {
      for (int i = 0; i < 9; ++i) cudaEventElapsedTime(&s->timings[i], s->ev[i], s->ev[i + 1]);
      s->timings[11] = 0.f;
    }
    cudaEventElapsedTime(&s->timings[9], s->ev[0], s->ev[9]);
    s->timings[10] = 0.f;
    if (s->kp.estimate_scaling) {
      cudaEventElapsedTime(&s->timings[10], s->ev[10], s->ev[11]);
      s->timings[0] -= s->timings[10];  // the scale vote sits between the upload and graph events
    }
  }
  if (s->fused) {
    b2r_status r = run_handed_over(s);
    if (r != B2R_OK) return r;
  }
  if (const char* bad = broken_canary(s)) return fail(s, B2R_ERR_STATE, std::string("guard zone overwritten: workspace buffer '") + bad + "'");
  s->have_result = true;
  const ProbState* hs = (const ProbState*)s->h_state.p;
  for (int b = 0; b < s->batch; ++b) {
    const uint32_t f = hs[b].flags;
    if (f & FLAG_COMPACT_TOO_LARGE) return fail(s, B2R_ERR_LIMIT, "pruned consistency graph exceeds 32768 vertices; exact clique search not supported at this size");
    if (f & FLAG_ARENA_OVERFLOW) return fail(s, B2R_ERR_LIMIT, "exact clique search ran out of stack arena (set B2R_ARENA_KB to enlarge)");
    if (f & FLAG_CANON_FAILED) return fail(s, B2R_ERR_STATE, "internal error: canonical clique reconstruction failed");
    if (f & (1u << 4)) return fail(s, B2R_ERR_LIMIT, "COMPLETE TIM graph over the clique exceeds the workspace cap");
  }
  return B2R_OK;
}

b2r_status b2r_solve_batch(b2r_solver* s, int batch, const int* n, const double* const* src, const double* const* dst,
                           int on_device) {
  b2r_status r = b2r_solve_batch_async(s, batch, n, src, dst, on_device);
  if (r != B2R_OK) return r;
  return b2r_sync(s);
}

b2r_status b2r_solve(b2r_solver* s, const double* src, const double* dst, int n, int64_t row_stride, int64_t col_stride,
                     int on_device) {
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!src || !dst) return fail(s, B2R_ERR_INVALID_ARG, "null point pointer");
  if (n < 2) return fail(s, B2R_ERR_INVALID_ARG, "need at least 2 correspondences");
  if (n > kMaxN) return fail(s, B2R_ERR_LIMIT, "N > 65535 correspondences is not supported");
  const bool packed = (row_stride == 1 && col_stride == 3);
  if (packed) return b2r_solve_batch(s, 1, &n, &src, &dst, on_device);
  if (on_device != 1) {
    std::vector<double> a((size_t)3 * n), b((size_t)3 * n);
    for (int c = 0; c < n; ++c)
      for (int r = 0; r < 3; ++r) {
        a[(size_t)3 * c + r] = src[r * row_stride + c * col_stride];
        b[(size_t)3 * c + r] = dst[r * row_stride + c * col_stride];
      }
    const double* pa = a.data(); const double* pb = b.data();
    return b2r_solve_batch(s, 1, &n, &pa, &pb, 0);
  }
  // strided device input: gather into the workspace first
  CK(cudaSetDevice(s->device));
  if (s->pending) { b2r_status r = b2r_sync(s); if (r != B2R_OK) return r; }
  b2r_status r = ensure_workspace(s, 1, n);
  if (r != B2R_OK) return r;
  double* ds = (double*)s->d_pts.p;
  double* dd = ds + (size_t)3 * n;
  gather_points_kernel<<<(3 * n + 255) / 256, 256, 0, s->stream>>>(src, row_stride, col_stride, n, ds);
  gather_points_kernel<<<(3 * n + 255) / 256, 256, 0, s->stream>>>(dst, row_stride, col_stride, n, dd);
  s->launches += 2;
  const double* pa = ds; const double* pb = dd;
  return b2r_solve_batch(s, 1, &n, &pa, &pb, 1);
}

b2r_status b2r_get_solution(b2r_solver* s, int idx, b2r_solution* out) {
  b2r_status r = check_index(s, idx);
  if (r != B2R_OK) return r;
  if (!out) return B2R_ERR_INVALID_ARG;
  const ProbState& st = ((const ProbState*)s->h_state.p)[idx];
  out->valid = st.valid;
  if (st.valid) {
    out->scale = st.scale;
    memcpy(out->rotation, st.R, sizeof(double) * 9);
    memcpy(out->translation, st.t, sizeof(double) * 3);
  } else {  // upstream: default-constructed solution (s=1, R=I, t=0, valid=false)
    out->scale = 1.0;
    for (int i = 0; i < 9; ++i) out->rotation[i] = (i % 4 == 0) ? 1.0 : 0.0;
    out->translation[0] = out->translation[1] = out->translation[2] = 0.0;
  }
  return B2R_OK;
}

b2r_status b2r_get_solutions(b2r_solver* s, b2r_solution* out, int count) {
  b2r_status r = check_index(s, 0);
  if (r != B2R_OK) return r;
  if (!out || count < 0 || count > s->batch) return fail(s, B2R_ERR_INVALID_ARG, "get_solutions: out / count");
  for (int i = 0; i < count; ++i) { r = b2r_get_solution(s, i, out + i); if (r != B2R_OK) return r; }
  return B2R_OK;
}

b2r_status b2r_get_stats(b2r_solver* s, int idx, b2r_stats* out) {
  b2r_status r = check_index(s, idx);
  if (r != B2R_OK) return r;
  if (!out) return B2R_ERR_INVALID_ARG;
  const ProbState& st = ((const ProbState*)s->h_state.p)[idx];
  memset(out, 0, sizeof(*out));
  out->n = s->n[idx];
  out->n_edges = (int64_t)(st.deg_sum / 2);
  out->max_degree = st.max_deg;
  out->heuristic_lb = st.lb;
  out->n_pruned = st.n_alive;
  out->omega = st.omega;
  out->clique_size = st.clique_size;
  out->n_rot_tims = st.n_rot_tims;
  out->n_rot_inliers = st.n_rot_inl;
  out->n_trans_inliers = st.n_trans_inl;
  out->gnc_iterations = st.gnc_iters;
  out->gnc_cost = st.gnc_cost;
  out->bnb_nodes = (int64_t)st.bnb_nodes;
  out->graph_slow_pairs = (int64_t)st.slow_pairs;
  out->exact_proven = (st.flags & FLAG_TIME_LIMIT) ? 0 : 1;
  out->flags = (int32_t)st.flags;
  out->fused_tail = (s->fused && st.tail_why == 0 && st.tail_done) ? 1 : 0;
  return B2R_OK;
}

b2r_status b2r_get_indices(b2r_solver* s, int idx, b2r_index_kind kind, int32_t* out, int32_t* count) {
  b2r_status r = check_index(s, idx);
  if (r != B2R_OK) return r;
  const ProbState& st = ((const ProbState*)s->h_state.p)[idx];
  const ProbDesc& d = ((const ProbDesc*)s->h_desc.p)[idx];
  CK(cudaSetDevice(s->device));
  const int C = st.clique_size > 0 ? st.clique_size : 0;
  const int nr = st.n_rot_inl, M = st.n_rot_inl;
  auto fetch_i = [&](const int* dev, int cnt, std::vector<int>& v) -> cudaError_t {
    v.resize(cnt > 0 ? cnt : 0);
    if (cnt <= 0) return cudaSuccess;
    cudaError_t e = cudaMemcpyAsync(v.data(), dev, sizeof(int) * cnt, cudaMemcpyDeviceToHost, s->stream);
    if (e != cudaSuccess) return e;
    return cudaStreamSynchronize(s->stream);
  };
  std::vector<int> res;
  switch (kind) {
    case B2R_IDX_MAX_CLIQUE: CK(fetch_i(d.clique, C, res)); break;
    case B2R_IDX_DEGREES: CK(fetch_i(d.deg, s->n[idx], res)); break;
    case B2R_IDX_ROT_INLIERS: CK(fetch_i(d.rot_inl, nr, res)); break;
    case B2R_IDX_ROT_INLIERS_MAP: {
      std::vector<int> cl, ri;
      CK(fetch_i(d.clique, C, cl)); CK(fetch_i(d.rot_inl, nr, ri));
      res.resize(ri.size());
      for (size_t i = 0; i < ri.size(); ++i) res[i] = C ? cl[ri[i] % C] : -1;
      break;
    }
    case B2R_IDX_TRANS_INLIERS:
    case B2R_IDX_TRANS_INLIERS_MAP: {
      std::vector<unsigned char> tm((size_t)(M > 0 ? M : 0));
      if (M > 0) {
        CK(cudaMemcpyAsync(tm.data(), d.trans_mask, (size_t)M, cudaMemcpyDeviceToHost, s->stream));
        CK(cudaStreamSynchronize(s->stream));
      }
      std::vector<int> cl, ri;
      if (kind == B2R_IDX_TRANS_INLIERS_MAP) { CK(fetch_i(d.clique, C, cl)); CK(fetch_i(d.rot_inl, nr, ri)); }
      for (int i = 0; i < M; ++i)
        if (tm[i]) res.push_back(kind == B2R_IDX_TRANS_INLIERS ? i : (C ? cl[ri[i] % C] : -1));
      break;
    }
    default: return fail(s, B2R_ERR_INVALID_ARG, "unknown index kind");
  }
  if (count) *count = (int32_t)res.size();
  if (out && !res.empty()) memcpy(out, res.data(), sizeof(int) * res.size());
  return B2R_OK;
}

b2r_status b2r_get_adjacency(b2r_solver* s, int idx, uint32_t* out, int words_per_row) {
  b2r_status r = check_index(s, idx);
  if (r != B2R_OK) return r;
  const int n = s->n[idx];
  const int need = (n + 31) / 32;
  if (!out || words_per_row < need) return fail(s, B2R_ERR_INVALID_ARG, "adjacency buffer too narrow");
  const ProbDesc& d = ((const ProbDesc*)s->h_desc.p)[idx];
  CK(cudaSetDevice(s->device));
  memset(out, 0, sizeof(uint32_t) * (size_t)n * words_per_row);
  CK(cudaMemcpy2DAsync(out, sizeof(uint32_t) * words_per_row, d.adj, sizeof(uint32_t) * d.pitch, sizeof(uint32_t) * need, n,
                       cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return B2R_OK;
}

b2r_status b2r_get_gnc_weights(b2r_solver* s, int idx, double* out) {
  b2r_status r = check_index(s, idx);
  if (r != B2R_OK) return r;
  const ProbState& st = ((const ProbState*)s->h_state.p)[idx];
  const ProbDesc& d = ((const ProbDesc*)s->h_desc.p)[idx];
  if (st.n_rot_tims <= 0) return B2R_OK;
  CK(cudaSetDevice(s->device));
  CK(cudaMemcpyAsync(out, d.weights, sizeof(double) * st.n_rot_tims, cudaMemcpyDeviceToHost, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  return B2R_OK;
}

b2r_status b2r_set_profiling(b2r_solver* s, int on) {
  if (!s) return B2R_ERR_INVALID_ARG;
  s->profiling = on != 0;
  return B2R_OK;
}

b2r_status b2r_get_timings(b2r_solver* s, float* ms12) {
  if (!s || !ms12) return B2R_ERR_INVALID_ARG;
  if (s->pending) { b2r_status r = b2r_sync(s); if (r != B2R_OK) return r; }
  memcpy(ms12, s->timings, sizeof(float) * 12);
  return B2R_OK;
}

b2r_status b2r_graph_rows(b2r_solver* s, const double* src_dev, const double* dst_dev, int n, int row_begin, int row_end,
                          uint32_t* adj_out_dev, int pitch_words) {
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!src_dev || !dst_dev || !adj_out_dev || n < 2 || row_begin < 0 || row_end > n || row_begin > row_end)
    return fail(s, B2R_ERR_INVALID_ARG, "graph_rows arguments");
  if (s->params.estimate_scaling) return fail(s, B2R_ERR_INVALID_ARG, "the row-sharded path needs estimate_scaling=False (the scale vote is global)");
  if (pitch_words < b2r_pitch_words(n) || (pitch_words & 3)) return fail(s, B2R_ERR_INVALID_ARG, "pitch_words must be >= b2r_pitch_words(n) and a multiple of 4");
  CK(cudaSetDevice(s->device));
  s->launches += launch_tim_graph_rows(src_dev, dst_dev, n, row_begin, row_end, adj_out_dev, pitch_words, s->kp, nullptr, s->stream);
  CK(cudaGetLastError());
  return B2R_OK;
}

b2r_status b2r_debug_set_graph_filter(b2r_solver* s, int filter, double tau) {
  if (!s) return B2R_ERR_INVALID_ARG;
  if (filter < 0 || filter > 5) return fail(s, B2R_ERR_INVALID_ARG, "graph filter must be 0 .. 5");
  if (s->pending) { b2r_status r = b2r_sync(s); if (r != B2R_OK) return r; }
  s->params.reserved0 = filter;
  s->params.reserved1 = tau > 0 ? tau : 0.0;
  fill_kernel_params(s);
  return B2R_OK;
}

b2r_status b2r_debug_check_canaries(b2r_solver* s) {
  if (!s) return B2R_ERR_INVALID_ARG;
  CK(cudaSetDevice(s->device));
  CK(cudaStreamSynchronize(s->stream));
  if (const char* bad = broken_canary(s)) return fail(s, B2R_ERR_STATE, std::string("guard zone overwritten: workspace buffer '") + bad + "'");
  return B2R_OK;
}

b2r_status b2r_debug_tail_cycles(b2r_solver* s, int idx, uint32_t* out32) {
  b2r_status r = check_index(s, idx);
  if (r != B2R_OK) return r;
  if (!out32) return B2R_ERR_INVALID_ARG;
  const ProbState& st = ((const ProbState*)s->h_state.p)[idx];
  for (int k = 0; k < 32; ++k) out32[k] = st.tail_clk[k];
  return B2R_OK;
}

b2r_status b2r_debug_tc_sweep(b2r_solver* s, int idx, double* out8) {
  b2r_status r = check_index(s, idx);
  if (r != B2R_OK) return r;
  if (!out8) return B2R_ERR_INVALID_ARG;
  const ProbState& st = ((const ProbState*)s->h_state.p)[idx];
  for (int k = 0; k < 8; ++k) out8[k] = 0.0;
  out8[7] = st.use32;
  if (!st.use32) return B2R_OK;
  CK(cudaSetDevice(s->device));
  double* dbuf = nullptr;
  CK(cudaMalloc(&dbuf, 8 * sizeof(double)));
  cudaError_t e = cudaMemsetAsync(dbuf, 0, 8 * sizeof(double), s->stream);
  if (e == cudaSuccess) {
    s->launches += launch_graph_tc_sweep((const ProbDesc*)s->d_desc.p, idx, s->n[idx], s->kp, dbuf, s->stream);
    e = cudaStreamSynchronize(s->stream);
  }
  if (e == cudaSuccess) e = cudaMemcpy(out8, dbuf, 7 * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(dbuf);
  CK(e);
  out8[7] = 1.0;
  return B2R_OK;
}

b2r_status b2r_debug_tc_values(b2r_solver* s, int idx, int row_block, int unit, float* dm_out, float* l_out, double* consts_out) {
  b2r_status r = check_index(s, idx);
  if (r != B2R_OK) return r;
  if (!dm_out || !l_out || !consts_out) return fail(s, B2R_ERR_INVALID_ARG, "debug_tc_values arguments");
  const ProbState& st = ((const ProbState*)s->h_state.p)[idx];
  const int n = s->n[idx];
  const int T = (n + kTile - 1) / kTile;
  if (row_block < 0 || row_block >= T || unit < 0 || unit >= 2 * (T - row_block)) return fail(s, B2R_ERR_INVALID_ARG, "debug_tc_values: block / unit out of range");
  consts_out[0] = st.kappa;
  for (int k = 0; k < 6; ++k) consts_out[1 + k] = st.ctr[k];
  consts_out[7] = -(double)st.tcc[0]; consts_out[8] = st.tcc[1]; consts_out[9] = st.tcc[2]; consts_out[10] = st.tcc[3];
  consts_out[11] = st.use32;
  if (!st.use32) return fail(s, B2R_ERR_STATE, "the tensor-core filter did not run on this problem (use32 == 0)");
  CK(cudaSetDevice(s->device));
  float* dbuf = nullptr;
  CK(cudaMalloc(&dbuf, 2 * 128 * 64 * sizeof(float)));
  s->launches += launch_graph_tc_debug((const ProbDesc*)s->d_desc.p, idx, row_block, unit, dbuf, dbuf + 128 * 64, s->stream);
  cudaError_t e = cudaStreamSynchronize(s->stream);
  if (e == cudaSuccess) e = cudaMemcpy(dm_out, dbuf, 128 * 64 * sizeof(float), cudaMemcpyDeviceToHost);
  if (e == cudaSuccess) e = cudaMemcpy(l_out, dbuf + 128 * 64, 128 * 64 * sizeof(float), cudaMemcpyDeviceToHost);
  cudaFree(dbuf);
  CK(e);
  return B2R_OK;
}

b2r_status b2r_solve_from_adjacency(b2r_solver* s, const double* src_dev, const double* dst_dev, int n, const uint32_t* adj_dev,
                                    int pitch_words) {
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!src_dev || !dst_dev || !adj_dev || n < 2) return fail(s, B2R_ERR_INVALID_ARG, "solve_from_adjacency arguments");
  if (s->params.estimate_scaling) return fail(s, B2R_ERR_INVALID_ARG, "the row-sharded path needs estimate_scaling=False (the scale vote is global)");
  if (n > kMaxN) return fail(s, B2R_ERR_LIMIT, "N > 65535 correspondences is not supported");
  if (pitch_words < b2r_pitch_words(n) || (pitch_words & 3)) return fail(s, B2R_ERR_INVALID_ARG, "pitch_words");
  CK(cudaSetDevice(s->device));
  if (s->pending) { b2r_status r = b2r_sync(s); if (r != B2R_OK) return r; }
  b2r_status r = ensure_workspace(s, 1, n);
  if (r != B2R_OK) return r;
  if (pitch_words != s->pitch && pitch_words > s->pitch) {
    // compact adjacency rows use the workspace pitch; a wider external pitch would overflow them
    return fail(s, B2R_ERR_INVALID_ARG, "external adjacency pitch wider than the workspace pitch");
  }
  if (s->profiling) { r = ensure_events(s); if (r != B2R_OK) return r; CK(cudaEventRecord(s->ev[0], s->stream)); }
  s->batch = 1; s->n.assign(1, n);
  build_desc(s, 0, n, src_dev, dst_dev, adj_dev, pitch_words);
  r = enqueue_pipeline(s, 1, n, true);
  if (r != B2R_OK) return r;
  return b2r_sync(s);
}

}  // extern "C"
