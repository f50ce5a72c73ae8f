// Internal: the solver handle behind include/b2r.h, shared by api.cu (registration) and presteps.cu / nn_gemm.cu.
#pragma once
#include <stdlib.h>

#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <nvtx3/nvToolsExt.h>   // header-only NVTX v3: a no-op function-pointer test unless a profiler is attached

#include "common.cuh"

namespace b2r {

// Device buffer of the workspace.  With B2R_CANARY=1 in the environment every buffer is allocated with a 4 KB guard zone of 0xA5
// bytes on either side; b2r_sync / b2r_destroy verify the zones (b2r_debug_check_canaries), so a kernel that writes past either end of
// a workspace buffer fails the call instead of corrupting a neighbour silently (compute-sanitizer is not available on the GPU pool).
inline size_t canary_bytes() {
  static const size_t g = [] { const char* e = getenv("B2R_CANARY"); return (e && e[0] == '1') ? (size_t)4096 : (size_t)0; }();
  return g;
}
struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  cudaError_t ensure(size_t need) {
    if (need <= bytes) return cudaSuccess;
    release();
    const size_t g = canary_bytes();
    char* base = nullptr;
    cudaError_t e = cudaMalloc(&base, need + 2 * g);
    if (e != cudaSuccess) return e;
    if (g) {
      e = cudaMemset(base, 0xA5, g);
      if (e == cudaSuccess) e = cudaMemset(base + g + need, 0xA5, g);
      if (e != cudaSuccess) { cudaFree(base); return e; }
    }
    p = base + g; bytes = need;
    return cudaSuccess;
  }
  void release() { if (p) cudaFree((char*)p - canary_bytes()); p = nullptr; bytes = 0; }
  // true when both guard zones still hold the pattern (always true without B2R_CANARY)
  bool intact() const {
    const size_t g = canary_bytes();
    if (!g || !p) return true;
    std::vector<unsigned char> h(2 * g);
    if (cudaMemcpy(h.data(), (char*)p - g, g, cudaMemcpyDeviceToHost) != cudaSuccess) return false;
    if (cudaMemcpy(h.data() + g, (char*)p + bytes, g, cudaMemcpyDeviceToHost) != cudaSuccess) return false;
    for (unsigned char c : h) if (c != 0xA5) return false;
    return true;
  }
};
struct HostBuf {
  void* p = nullptr;
  size_t bytes = 0;
  cudaError_t ensure(size_t need) {
    if (need <= bytes) return cudaSuccess;
    if (p) cudaFreeHost(p);
    p = nullptr; bytes = 0;
    cudaError_t e = cudaMallocHost(&p, need);
    if (e == cudaSuccess) bytes = need;
    return e;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; bytes = 0; }
};

}  // namespace b2r

// Worker threads of a handle for the host-side staging copy (on_device = 0): created once, woken per batch.  Spawning and joining
// std::threads per call cost more than the copy of a small batch (config 4: 3 MB per frame).
class PackPool {
 public:
  ~PackPool() { stop(); }
  // run fn(t) for t in [0, nthreads) -- t = 0 on the calling thread -- and return when all are done
  void run(int nthreads, const std::function<void(int)>& fn) {
    if (nthreads <= 1) { fn(0); return; }
    ensure(nthreads - 1);
    {
      std::lock_guard<std::mutex> lk(m_);
      job_ = &fn; active_ = nthreads - 1; pending_ = nthreads - 1; ++gen_;
    }
    cv_.notify_all();
    fn(0);
    std::unique_lock<std::mutex> lk(m_);
    done_.wait(lk, [&] { return pending_ == 0; });
    job_ = nullptr;
  }
  void stop() {
    { std::lock_guard<std::mutex> lk(m_); quit_ = true; ++gen_; }
    cv_.notify_all();
    for (auto& t : th_) t.join();
    th_.clear();
    quit_ = false;
  }

 private:
  void ensure(int workers) {
    while ((int)th_.size() < workers) {
      const int id = (int)th_.size() + 1;
      unsigned long long seen;
      { std::lock_guard<std::mutex> lk(m_); seen = gen_; }
      th_.emplace_back([this, id, seen]() mutable {
        for (;;) {
          const std::function<void(int)>* job = nullptr;
          {
            std::unique_lock<std::mutex> lk(m_);
            cv_.wait(lk, [&] { return gen_ != seen; });
            seen = gen_;
            if (quit_) return;
            if (id <= active_) job = job_;
          }
          if (job) {
            (*job)(id);
            std::lock_guard<std::mutex> lk(m_);
            if (--pending_ == 0) done_.notify_one();
          }
        }
      });
    }
  }
  std::mutex m_;
  std::condition_variable cv_, done_;
  std::vector<std::thread> th_;
  const std::function<void(int)>* job_ = nullptr;
  unsigned long long gen_ = 0;
  int active_ = 0, pending_ = 0;
  bool quit_ = false;
};

struct b2r_solver {
  PackPool pack_pool;
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  // latency-bound tail of a batch (degrees, clique, rotation, translation: 64 small CTAs per kernel) on a HIGH-priority stream, so that
  // with several handles in flight it is not starved by another handle's graph kernel, which fills every SM
  cudaStream_t stream_hi = nullptr;
  cudaEvent_t ev_graph = nullptr, ev_tail = nullptr;
  b2r_params params;
  b2r::KernelParams kp;
  std::string err;
  long long launches = 0;

  // workspace capacity
  int batch_cap = 0, n_cap = 0, pitch = 0;
  long long tim_cap = 0;
  b2r::DevBuf d_pts, d_adj, d_cadj, d_int, d_dbl, d_u8, d_state, d_desc, d_arena, d_scale, d_pre, d_nn, d_grid, d_refine, d_pts32, d_amb;
  b2r::HostBuf h_state, h_desc, h_pts, h_refine;
  size_t arena_words_per_warp = 64 * 1024;
  int bnb_blocks = 1;

  // last batch
  int batch = 0;
  std::vector<int> n;
  bool pending = false;
  bool have_result = false;
  const uint32_t* ext_adj = nullptr;  // solve_from_adjacency
  bool fused = false;                 // the last batch went through the fused tail kernel (tail.cu)
  int handed_over = 0;                // problems of the last batch the fused tail handed to the multi-kernel path
  int n_max = 0;

  // mailbox for the small host reads that size the next stage (counts, bounds): page-locked, mapped words a one-warp kernel writes and
  // the host polls (b2r_internal_read_words, api.cu) -- a cudaMemcpyAsync into pageable memory plus a stream synchronise cost ~3x more
  volatile unsigned* mbox_h = nullptr;
  unsigned* mbox_d = nullptr;
  unsigned mbox_seq = 0;

  // one problem tile-sharded over the GPUs of a box (shard.cu): peers' matrices and flag words mapped through CUDA IPC
  struct Shard {
    bool exported = false, attached = false;
    int rank = 0, world = 1, n = 0;
    unsigned epoch = 0;
    unsigned* flags = nullptr;          // this rank's flag words: ready[8], done[8] (own cudaMalloc, exported)
    uint32_t* peer_adj[7] = {};
    unsigned* peer_flags[7] = {};
    void* opened[14] = {};
    int n_opened = 0;
    bool last_fell_back = false;
  } shard;

  bool profiling = false;
  cudaEvent_t ev[12] = {};
  bool ev_ready = false;
  float timings[12] = {};
};


#define CK(call)                                                                                  \
  do {                                                                                            \
    cudaError_t e__ = (call);                                                                     \
    if (e__ != cudaSuccess) {                                                                     \
      s->err = std::string(#call) + ": " + cudaGetErrorString(e__);                               \
      return e__ == cudaErrorMemoryAllocation ? B2R_ERR_OOM : B2R_ERR_CUDA;                       \
    }                                                                                             \
  } while (0)


// NVTX range over one stage of the path: shows as "b2r:<stage>" rows in Nsight Systems / ncu --nvtx
struct B2rNvtx {
  explicit B2rNvtx(const char* name) { nvtxRangePushA(name); }
  ~B2rNvtx() { nvtxRangePop(); }
  B2rNvtx(const B2rNvtx&) = delete;
  B2rNvtx& operator=(const B2rNvtx&) = delete;
};
#define B2R_NVTX_CAT2(a, b) a##b
#define B2R_NVTX_CAT(a, b) B2R_NVTX_CAT2(a, b)
#define B2R_NVTX(name) B2rNvtx B2R_NVTX_CAT(nvtx_range_, __LINE__)(name)

// n_words (<= 30) 32-bit words of device memory to the host, ordered after everything queued on s->stream so far; returns when they
// have arrived (api.cu).  The stream itself is NOT synchronised beyond the one-warp kernel that publishes the words.
b2r_status b2r_internal_read_words(b2r_solver* s, const void* dev_words, int n_words, void* host_out);

static inline b2r_status fail(b2r_solver* s, b2r_status code, const std::string& msg) {
  if (s) s->err = msg;
  return code;
}
