// One large registration tile-sharded over the GPUs of a box (BASELINE.json config 5).
//
// The reference has no counterpart (one solver.solve call, HelpersRealtime.py:166); the north-star asks for the consistency graph of
// a very large N to be sharded with the adjacency exchanged over NVLink.  Instead of "compute rows, then all-gather" (two steps, the
// NCCL form kept in poseestimator_b200/sharded.py as the baseline), the exchange is FUSED into the graph kernel:
//
//   * every rank holds the replicated 3 x N operands and a full N x N bit matrix of its own, mapped into every other rank's address
//     space through CUDA IPC (b2r_shard_export / b2r_shard_attach);
//   * rank r evaluates the work items r, r + w, ... of the tensor-core filter (tim_graph6_kernel<.., PEERS>): 1 / w of the pair tests;
//     when an item's words leave shared memory they are stored to the local matrix and, with plain NVLink stores, to the same
//     offsets of the w - 1 peer matrices -- the transfer overlaps the arithmetic item by item, no staging buffer, no second pass;
//   * the rank's ambiguous pairs are decided exactly and their bits OR-ed into every copy (graph_fix_peers_kernel);
//   * two flag words per peer pair order the epochs: READY (peers have finished reading their matrix of the previous solve, so it may
//     be overwritten) and DONE (all of a rank's stores have been issued: st.release.sys after the kernels of that rank);
//   * after DONE from everyone each rank owns the complete matrix and continues alone: degrees, fused tail (replicated; the
//     max-clique stage is "replicas only").
// A rank whose filter refuses the problem (ill-conditioned input, pair-list overflow) says so in its DONE word; every rank then
// solves the problem by itself through the ordinary path -- same result, no exchange.
#include <cuda.h>
#include <string.h>

#include "solver.cuh"

using namespace b2r;

// api.cu
b2r_status b2r_internal_prepare_single(b2r_solver* s, int n, const double* src_dev, const double* dst_dev);
b2r_status b2r_internal_finish_from_graph(b2r_solver* s, int n);

namespace {

struct ShardFlags {
  unsigned* local;
  unsigned* peer[7];
  int n_peer, rank, world;
};

// slot layout of a rank's flag words: [which * 8 + sender]
__global__ void shard_signal_kernel(ShardFlags f, int which, unsigned epoch, const ProbState* st) {
  unsigned v = epoch << 1;
  if (st && (!st->use32 || st->tc_overflow)) v |= 1u;   // DONE word: "my part of the matrix is not valid"
  const int p = threadIdx.x;
  __threadfence_system();
  if (p < f.n_peer) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f.peer[p] + which * 8 + f.rank), "r"(v) : "memory");
  if (p == f.n_peer) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f.local + which * 8 + f.rank), "r"(v) : "memory");
}

__global__ void shard_wait_kernel(ShardFlags f, int which, unsigned epoch, ProbState* st, long long timeout_cycles) {
  __shared__ unsigned bad, late;
  if (threadIdx.x == 0) { bad = 0u; late = 0u; }
  __syncthreads();
  const int p = threadIdx.x;
  if (p < f.world) {
    const long long t0 = clock64();
    for (;;) {
      unsigned v;
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f.local + which * 8 + p) : "memory");
      if ((int)((v >> 1) - epoch) >= 0) { if (v & 1u) atomicOr(&bad, 1u); break; }
      if (clock64() - t0 > timeout_cycles) { atomicOr(&late, 1u); break; }
      __nanosleep(200);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0 && st) {
    if (late) st->flags |= FLAG_SHARD_TIMEOUT;
    else if (bad) st->flags |= FLAG_SHARD_FALLBACK;
  }
}

typedef CUresult (*GetRangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);

bool alloc_base(const void* p, const void** base) {
  static GetRangeFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) { cudaGetLastError(); f = nullptr; }
    return (GetRangeFn)f;
  }();
  if (!fn) return false;
  CUdeviceptr b = 0; size_t sz = 0;
  if (fn(&b, &sz, (CUdeviceptr)p) != CUDA_SUCCESS) return false;
  *base = (const void*)b;
  return true;
}

void detach(b2r_solver* s) {
  for (int i = 0; i < s->shard.n_opened; ++i) cudaIpcCloseMemHandle(s->shard.opened[i]);
  s->shard.n_opened = 0;
  s->shard.attached = false;
}

}  // namespace

extern "C" {

b2r_status b2r_shard_export(b2r_solver* s, int n, b2r_peer_handle* out) {
  if (!s || !out) return B2R_ERR_INVALID_ARG;
  if (n < 2 || n > kMaxN) return fail(s, B2R_ERR_INVALID_ARG, "shard_export: n");
  if (s->params.estimate_scaling) return fail(s, B2R_ERR_INVALID_ARG, "the sharded path needs estimate_scaling=False (the scale vote is global)");
  B2R_NVTX("b2r:shard_export");
  CK(cudaSetDevice(s->device));
  if (s->pending) { b2r_status r = b2r_sync(s); if (r != B2R_OK) return r; }
  detach(s);
  b2r_status r = b2r_internal_prepare_single(s, n, nullptr, nullptr);   // sizes the workspace for (1, n)
  if (r != B2R_OK) return r;
  if (!s->shard.flags) {
    CK(cudaMalloc(&s->shard.flags, 4096));
    CK(cudaMemset(s->shard.flags, 0, 4096));
  }
  memset(out, 0, sizeof(*out));
  const void* base = nullptr;
  if (!alloc_base(s->d_adj.p, &base)) return fail(s, B2R_ERR_CUDA, "cuMemGetAddressRange unavailable");
  out->adj_offset = (int64_t)((const char*)s->d_adj.p - (const char*)base);
  CK(cudaIpcGetMemHandle((cudaIpcMemHandle_t*)out->adjacency, const_cast<void*>(base)));
  if (!alloc_base(s->shard.flags, &base)) return fail(s, B2R_ERR_CUDA, "cuMemGetAddressRange unavailable");
  out->flag_offset = (int64_t)((const char*)s->shard.flags - (const char*)base);
  CK(cudaIpcGetMemHandle((cudaIpcMemHandle_t*)out->flags, const_cast<void*>(base)));
  out->n = n;
  out->device = s->device;
  s->shard.exported = true;
  s->shard.n = n;
  return B2R_OK;
}

b2r_status b2r_shard_attach(b2r_solver* s, int rank, int world, const b2r_peer_handle* all) {
  if (!s || !all) return B2R_ERR_INVALID_ARG;
  if (world < 1 || world > 8 || rank < 0 || rank >= world) return fail(s, B2R_ERR_INVALID_ARG, "shard_attach: 1 <= world <= 8, 0 <= rank < world");
  if (!s->shard.exported) return fail(s, B2R_ERR_STATE, "shard_attach before shard_export");
  B2R_NVTX("b2r:shard_attach");
  CK(cudaSetDevice(s->device));
  detach(s);
  int k = 0;
  for (int p = 0; p < world; ++p) {
    if (p == rank) continue;
    if (all[p].n != s->shard.n) return fail(s, B2R_ERR_INVALID_ARG, "shard_attach: the ranks exported different problem sizes");
    void* a = nullptr; void* f = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&a, *(const cudaIpcMemHandle_t*)all[p].adjacency, cudaIpcMemLazyEnablePeerAccess);
    if (e == cudaSuccess) { s->shard.opened[s->shard.n_opened++] = a; e = cudaIpcOpenMemHandle(&f, *(const cudaIpcMemHandle_t*)all[p].flags, cudaIpcMemLazyEnablePeerAccess); }
    if (e == cudaSuccess) s->shard.opened[s->shard.n_opened++] = f;
    if (e != cudaSuccess) {
      cudaGetLastError();
      detach(s);
      s->err = std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e) + " (no peer access between the two GPUs?)";
      return B2R_ERR_CUDA;
    }
    s->shard.peer_adj[k] = (uint32_t*)((char*)a + all[p].adj_offset);
    s->shard.peer_flags[k] = (unsigned*)((char*)f + all[p].flag_offset);
    ++k;
  }
  s->shard.rank = rank; s->shard.world = world;
  s->shard.attached = true;
  s->shard.epoch = 0;
  CK(cudaMemset(s->shard.flags, 0, 4096));
  return B2R_OK;
}

b2r_status b2r_shard_detach(b2r_solver* s) {
  if (!s) return B2R_ERR_INVALID_ARG;
  cudaSetDevice(s->device);
  cudaStreamSynchronize(s->stream);
  detach(s);
  if (s->shard.flags) { cudaFree(s->shard.flags); s->shard.flags = nullptr; }
  s->shard.exported = false;
  return B2R_OK;
}

b2r_status b2r_solve_sharded(b2r_solver* s, const double* src_dev, const double* dst_dev, int n) {
  if (!s) return B2R_ERR_INVALID_ARG;
  if (!src_dev || !dst_dev) return fail(s, B2R_ERR_INVALID_ARG, "solve_sharded: null operand");
  if (!s->shard.attached) return fail(s, B2R_ERR_STATE, "solve_sharded before shard_attach");
  if (n != s->shard.n) return fail(s, B2R_ERR_INVALID_ARG, "solve_sharded: n differs from the exported size");
  B2R_NVTX("b2r:solve_sharded");
  CK(cudaSetDevice(s->device));
  if (s->pending) { b2r_status r = b2r_sync(s); if (r != B2R_OK) return r; }
  const void* adj_before = s->d_adj.p;
  b2r_status r = b2r_internal_prepare_single(s, n, src_dev, dst_dev);   // state cleared, descriptor of problem 0 on the device
  if (r != B2R_OK) return r;
  if (s->d_adj.p != adj_before) return fail(s, B2R_ERR_STATE, "the workspace was re-allocated since shard_export: export and attach again");
  cudaStream_t st = s->stream;

  ShardFlags f;
  f.local = s->shard.flags; f.n_peer = s->shard.world - 1; f.rank = s->shard.rank; f.world = s->shard.world;
  for (int p = 0; p < 7; ++p) f.peer[p] = s->shard.peer_flags[p];
  KernelParams kp = s->kp;
  kp.shard_rank = s->shard.rank; kp.shard_world = s->shard.world; kp.n_peer = s->shard.world - 1;
  for (int p = 0; p < 7; ++p) kp.peer_adj[p] = s->shard.peer_adj[p];
  const unsigned epoch = ++s->shard.epoch;
  const ProbDesc* dd = (const ProbDesc*)s->d_desc.p;
  ProbState* dst0 = (ProbState*)s->d_state.p;
  long long timeout = 4000000000ll;   // ~2 s of SM clock: a peer that never arrives must not hang the GPU
  if (const char* e = getenv("B2R_SHARD_TIMEOUT_MS")) { const long long ms = atoll(e); if (ms > 0) timeout = ms * 2000000ll; }
  // READY: every peer has finished reading its matrix of the previous solve (its tail ran before its READY signal, in stream order)
  shard_signal_kernel<<<1, 32, 0, st>>>(f, 0, epoch, nullptr);
  shard_wait_kernel<<<1, 32, 0, st>>>(f, 0, epoch, dst0, timeout);
  s->launches += 2;
  s->launches += launch_tim_graph_tc_sharded(dd, n, kp, st);
  // DONE: all of this rank's stores (kernels above, stream order) are issued and fenced before the flag becomes visible
  shard_signal_kernel<<<1, 32, 0, st>>>(f, 1, epoch, dst0);
  shard_wait_kernel<<<1, 32, 0, st>>>(f, 1, epoch, dst0, timeout);
  s->launches += 2;
  CK(cudaGetLastError());
  r = b2r_internal_finish_from_graph(s, n);   // degrees from the complete matrix, fused tail, state to the host; synchronises
  if (r != B2R_OK) return r;
  const ProbState* hs = (const ProbState*)s->h_state.p;
  s->shard.last_fell_back = false;
  if (hs[0].flags & FLAG_SHARD_TIMEOUT) { s->have_result = false; return fail(s, B2R_ERR_STATE, "solve_sharded: a peer rank did not arrive (flag time-out)"); }
  if (hs[0].flags & FLAG_SHARD_FALLBACK) {   // decided identically on every rank: solve alone, ordinary path
    s->shard.last_fell_back = true;
    return b2r_solve_batch(s, 1, &n, &src_dev, &dst_dev, 1);
  }
  return B2R_OK;
}

int b2r_shard_last_fell_back(const b2r_solver* s) { return s && s->shard.last_fell_back ? 1 : 0; }

}  // extern "C"
