// Microbenchmark: sustained FP64-pipe instruction rate on this GPU (DFMA / DADD / DMUL, 8 independent chains per thread).
// The consistency-graph kernel is FP64-pipe bound; MEASURED_PEAKS.json has no FP64 number (SURVEY.md 8d asks for one).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_peak fp64_peak.cu ; prints one JSON line.
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__global__ void __launch_bounds__(256) k(double* out, int iters, double a, double b) {
  double x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = a + threadIdx.x * 1e-9 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (OP == 0) x[i] = fma(x[i], a, b);
        else if (OP == 1) x[i] = __dadd_rn(x[i], b);
        else x[i] = __dmul_rn(x[i], a);
      }
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += x[i];
  if (s == 123.456) out[0] = s;
}

template <int OP>
double run(int blocks, int iters) {
  double* d; cudaMalloc(&d, 8);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<OP><<<blocks, 256>>>(d, iters, 1.0000001, 1e-9);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0); k<OP><<<blocks, 256>>>(d, iters, 1.0000001, 1e-9); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  cudaFree(d);
  return (double)blocks * 256 * iters * 32.0 / (best * 1e-3);  // thread-instructions per second
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int blocks = p.multiProcessorCount * 8, iters = 20000;
  const double fma_ = run<0>(blocks, iters), add_ = run<1>(blocks, iters), mul_ = run<2>(blocks, iters);
  int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"dfma_ginst_s\": %.1f, \"dadd_ginst_s\": %.1f, \"dmul_ginst_s\": %.1f, \"dfma_tflops\": %.2f, "
         "\"lanes_per_sm_per_clk_at_max_clock\": %.1f, \"max_clock_khz\": %d}\n",
         p.name, p.multiProcessorCount, fma_ / 1e9, add_ / 1e9, mul_ / 1e9, 2 * fma_ / 1e12, fma_ / (p.multiProcessorCount * (double)clk * 1e3), clk);
  return 0;
}
