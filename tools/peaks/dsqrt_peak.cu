// Microbenchmark: sustained rate of IEEE fp64 square roots (__dsqrt_rn) and divisions on this GPU -- the two operations of the
// reference's edge verdict |sqrt(A) - sqrt(B)| <= beta that the graph filters avoid ("measure ... DSQRT").
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dsqrt_peak dsqrt_peak.cu ; prints one JSON line.
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__global__ void __launch_bounds__(256) k(double* out, int iters, double a) {
  double x[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = a + threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) x[i] = __dsqrt_rn(x[i]) + 2.0;      // one DSQRT sequence + one DADD (keeps the value in range)
      else x[i] = __ddiv_rn(a, x[i]) + 2.0;
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
  k<OP><<<blocks, 256>>>(d, iters, 3.7);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0); k<OP><<<blocks, 256>>>(d, iters, 3.7); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  cudaFree(d);
  return (double)blocks * 256 * iters * 8.0 / (best * 1e-3);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int blocks = p.multiProcessorCount * 8, iters = 4000;
  const double sq = run<0>(blocks, iters), dv = run<1>(blocks, iters);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"dsqrt_rn_gops_s\": %.1f, \"ddiv_rn_gops_s\": %.1f, "
         "\"note\": \"IEEE-rounded fp64 sqrt / div are instruction sequences (MUFU seed + DFMA Newton steps); each op also carries one DADD\"}\n",
         p.name, p.multiProcessorCount, sq / 1e9, dv / 1e9);
  return 0;
}
