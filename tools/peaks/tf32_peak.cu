// Microbenchmark: sustained tcgen05.mma kind::tf32 rate on this GPU (operands resident in shared memory, accumulators in TMEM, no
// loads in the loop) -- the denominator for the "tensor pipe x % busy" figures of tim_graph6_kernel and nn_gemm_kernel
// ("TF32-tcgen05 ... measure on the box").  One CTA per SM (or two with B2R_CTAS=2), one thread issues M = 128,
// N = 256, K = 8 instructions back to back into two alternating accumulators and commits every 64 of them.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tf32_peak tf32_peak.cu ; prints one JSON line.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ constexpr uint32_t umma_idesc_tf32(int m, int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

constexpr int M = 128, N = 256;
struct Ctl { uint64_t bar; uint32_t tmem; };

__global__ void __launch_bounds__(128, 1) k(int rounds, float* sink) {
  extern __shared__ __align__(1024) unsigned char raw[];
  unsigned char* sm = raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
  unsigned char* sA = sm;                    // 128 rows x 128 B (one 32-float k-block, SWIZZLE_128B)
  unsigned char* sB = sm + M * 128;          // 256 rows x 128 B
  Ctl* c = reinterpret_cast<Ctl*>(sm + (M + N) * 128);
  for (int i = threadIdx.x; i < (M + N) * 32; i += 128) reinterpret_cast<float*>(sm)[i] = 1.0f;   // finite operands
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(&c->bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&c->tmem)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = c->tmem;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = umma_idesc_tf32(M, N);
    const uint64_t ad = umma_desc_sw128(smem_u32(sA)), bd = umma_desc_sw128(smem_u32(sB));
    uint32_t ph = 0;
    for (int r = 0; r < rounds; ++r) {
#pragma unroll 1
      for (int i = 0; i < 64; ++i) tc_mma_tf32(tmem + (uint32_t)(i & 1) * N, ad + 2 * (i & 3), bd + 2 * (i & 3), idesc, 1u);
      tc_commit(&c->bar);
      mbar_wait(&c->bar, ph);
      ph ^= 1u;
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
  if (sink && threadIdx.x == 999) sink[0] = 1.f;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int smem = (M + N) * 128 + 2048;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int blocks = p.multiProcessorCount, rounds = 400;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<<<blocks, 128, smem>>>(rounds, nullptr);
  if (cudaDeviceSynchronize() != cudaSuccess) { printf("{\"error\": \"%s\"}\n", cudaGetErrorString(cudaGetLastError())); return 1; }
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0); k<<<blocks, 128, smem>>>(rounds, nullptr); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  const double flops = (double)blocks * rounds * 64.0 * 2.0 * M * N * 8;
  int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"tf32_tcgen05_tflops\": %.1f, \"shape\": \"M128 N256 K8, cta_group::1, one CTA per SM\", "
         "\"mac_per_sm_per_clk_at_max_clock\": %.0f, \"ms\": %.4f, \"max_clock_khz\": %d}\n",
         p.name, p.multiProcessorCount, flops / (best * 1e-3) / 1e12, flops / 2 / (best * 1e-3) / (p.multiProcessorCount * (double)clk * 1e3), best, clk);
  return 0;
}
