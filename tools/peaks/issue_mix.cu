// Issue-mix microbenchmark for sm_100a: does the integer pipe (LOP3 / SHF / FSETP) or an LDS issue in the shadow of the
// 2-cycle packed-fp32 instructions (FFMA2 / FADD2), and what is the dependent-issue latency of FFMA2?  The graph filter's hot
// loop is a mix of exactly these.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o issue_mix issue_mix.cu
// Output: cycles per iteration per SM sub-partition at the maximum clock, per warp-iteration (8 FFMA2 + extras).
#include <cstdio>
#include <cuda_runtime.h>

#define FMA2(acc, p, q) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc) : "l"(p), "l"(q))
#define LOP(x, y) asm volatile("lop3.b32 %0, %0, %1, %1, 0x96;" : "+r"(x) : "r"(y))
#define SHF(x, y) asm volatile("shf.l.wrap.b32 %0, %1, %0, 1;" : "+r"(x) : "r"(y))
#define SETP(pw, a, b) asm volatile("{.reg .pred q; setp.lt.f32 q, %1, %2; @q or.b32 %0, %0, 1;}" : "+r"(pw) : "f"(a), "f"(b))

template <int MODE>
__global__ void k(float* out, int iters, unsigned long long p, unsigned long long q, unsigned s, float fa) {
  __shared__ float4 sm[256];
  sm[threadIdx.x & 255] = make_float4(threadIdx.x & 255, 1.f, 2.f, 3.f);
  __syncthreads();
  unsigned long long acc[8];
  unsigned x[8];
  float4 ld = make_float4(0, 0, 0, 0);
  for (int i = 0; i < 8; ++i) { acc[i] = p + i * 17 + threadIdx.x; x[i] = s + i + threadIdx.x; }
  for (int it = 0; it < iters; ++it) {
    if (MODE == 1) {  // one dependent chain of 8
#pragma unroll
      for (int i = 0; i < 8; ++i) FMA2(acc[0], p, q);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        FMA2(acc[i], p, q);
        if (MODE == 2) LOP(x[i], s);
        if (MODE == 3) SETP(x[i], __uint_as_float(x[(i + 1) & 7]), fa);
        if (MODE == 4) SHF(x[i], s);
        if (MODE == 5 && (i & 1) == 0) { const float4 v = sm[(it + i) & 255]; ld.x += v.x; }
        if (MODE == 6) { LOP(x[i], s); SHF(x[i], s); }
        if (MODE == 7) { LOP(x[i], s); SHF(x[i], s); LOP(x[(i + 1) & 7], s); }
        if (MODE == 8 && (i & 1) == 0) { const float4 v = sm[(it + i) & 255]; x[i] ^= __float_as_uint(v.x) ^ __float_as_uint(v.y) ^ __float_as_uint(v.z) ^ __float_as_uint(v.w); }
      }
    }
  }
  float r = ld.x;
  for (int i = 0; i < 8; ++i) r += (float)(acc[i] & 0xffff) + x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE>
void run(const char* name, int warps_per_smsp, float* out) {
  int dev = 0, sms = 0, khz = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  const int iters = 20000, threads = 128 * warps_per_smsp, blocks = sms;
  const float one = 1.0001f, two = 0.9999f;
  unsigned long long p, q;
  p = ((unsigned long long)__builtin_bit_cast(unsigned, one) << 32) | __builtin_bit_cast(unsigned, two);
  q = ((unsigned long long)__builtin_bit_cast(unsigned, two) << 32) | __builtin_bit_cast(unsigned, one);
  k<MODE><<<blocks, threads>>>(out, 100, p, q, 12345u, 0.5f);
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  k<MODE><<<blocks, threads>>>(out, iters, p, q, 12345u, 0.5f);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) { printf("{\"mode\": \"%s\", \"error\": \"%s\"}\n", name, cudaGetErrorString(err)); fflush(stdout); return; }
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  const double cyc = ms * 1e-3 * khz * 1e3;                       // cycles at max clock
  const double per_warp_iter = cyc / ((double)iters * warps_per_smsp);
  printf("{\"mode\": \"%s\", \"warps_per_smsp\": %d, \"ms\": %.3f, \"cycles_per_warp_iteration_of_8_FFMA2\": %.2f}\n", name, warps_per_smsp, ms,
         per_warp_iter);
  fflush(stdout);
}

int main() {
  float* out;
  cudaMalloc(&out, 1 << 24);
  run<0>("8 FFMA2 independent", 8, out);
  run<0>("8 FFMA2 independent", 4, out);
  run<0>("8 FFMA2 independent", 1, out);
  run<1>("8 FFMA2 one dependent chain (latency x 8)", 1, out);
  run<1>("8 FFMA2 one dependent chain", 4, out);
  run<2>("8 FFMA2 + 8 LOP3", 8, out);
  run<3>("8 FFMA2 + 8 (FSETP + predicated LOP3)", 8, out);
  run<4>("8 FFMA2 + 8 SHF", 8, out);
  run<5>("8 FFMA2 + 4 LDS.128 (+4 FADD)", 8, out);
  run<6>("8 FFMA2 + 8 LOP3 + 8 SHF", 8, out);
  run<7>("8 FFMA2 + 16 LOP3 + 8 SHF", 8, out);
  run<8>("8 FFMA2 + 4 LDS.128 + 16 LOP3", 8, out);
  run<6>("8 FFMA2 + 8 LOP3 + 8 SHF", 4, out);
  return 0;
}
