// Packed-fp32 (FADD2 / FMUL2 / FFMA2) issue-rate microbenchmark for sm_100a: warp-instructions per clock per SM sub-partition
// for the operand forms the graph filter uses.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp32x2_peak fp32x2_peak.cu
#include <cstdio>
#include <cuda_runtime.h>
#define F2(name, ptx) __device__ __forceinline__ float2 name(float2 a, float2 b) { float2 r; \
  asm volatile("{.reg .b64 ra, rb, rc; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; " ptx " rc, ra, rb; mov.b64 {%0,%1}, rc;}" \
  : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y)); return r; }
F2(add2, "add.rn.f32x2") F2(mul2, "mul.rn.f32x2")
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { float2 r;
  asm volatile("{.reg .b64 ra, rb, rc, rd; mov.b64 ra, {%2,%3}; mov.b64 rb, {%4,%5}; mov.b64 rc, {%6,%7}; fma.rn.f32x2 rd, ra, rb, rc; mov.b64 {%0,%1}, rd;}"
  : "=f"(r.x), "=f"(r.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y)); return r; }
__device__ __forceinline__ float sfma(float a, float b, float c) { float r; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }

template <int MODE>
__global__ void k(float* out, int iters, float s0, float s1) {
  float2 acc[8];
  for (int i = 0; i < 8; ++i) acc[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f);
  float2 p = make_float2(s0, s1), q = make_float2(s1, s0);
  float sa[8]; for (int i = 0; i < 8; ++i) sa[i] = threadIdx.x + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) acc[i] = fma2(acc[i], p, q);                 // FFMA2, three distinct register pairs
      if (MODE == 1) acc[i] = fma2(p, p, acc[i]);                 // FFMA2 d, d, acc
      if (MODE == 2) acc[i] = add2(acc[i], make_float2(s0, s0));  // FADD2 pair + scalar broadcast
      if (MODE == 3) acc[i] = mul2(acc[i], acc[i]);               // FMUL2 d, d
      if (MODE == 4) acc[i] = add2(acc[i], p);                    // FADD2 pair + pair
      if (MODE == 5) sa[i] = sfma(sa[i], s0, s1);                 // scalar FFMA
      if (MODE == 6) { acc[i] = fma2(acc[i], p, q); sa[i] = sfma(sa[i], s0, s1); }  // one FFMA2 + one FFMA
      if (MODE == 7) { acc[i] = fma2(acc[i], p, q); sa[i] = sfma(sa[i], s0, s1); sa[(i + 1) & 7] = sfma(sa[(i + 1) & 7], s1, s0); }
    }
  }
  float r = 0; for (int i = 0; i < 8; ++i) r += acc[i].x + acc[i].y + sa[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}
template <int MODE> void run(const char* name, int instr_per_iter, float* out) {
  int dev = 0, sms = 0, khz = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  const int iters = 20000, threads = 256, blocks = sms * 4;   // 8 warps per SMSP
  k<MODE><<<blocks, threads>>>(out, 100, 1.0001f, 0.9999f);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a); k<MODE><<<blocks, threads>>>(out, iters, 1.0001f, 0.9999f); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double winst = (double)blocks * (threads / 32) * iters * 8.0 * instr_per_iter;
  const double per_clk_smsp = winst / (ms * 1e-3) / (sms * 4.0) / (khz * 1e3);
  printf("{\"mode\": \"%s\", \"ms\": %.3f, \"warp_inst_per_clk_per_smsp_at_max_clock\": %.3f, \"Tinst_s\": %.2f}\n", name, ms, per_clk_smsp, winst * 32 / (ms * 1e-3) / 1e12);
}
int main() {
  float* out; cudaMalloc(&out, 1 << 24);
  run<0>("FFMA2 a,b,c distinct pairs", 1, out);
  run<1>("FFMA2 p,p,acc", 1, out);
  run<2>("FADD2 pair + scalar", 1, out);
  run<3>("FMUL2 d,d", 1, out);
  run<4>("FADD2 pair + pair", 1, out);
  run<5>("FFMA scalar", 1, out);
  run<6>("FFMA2 + FFMA (2 inst)", 2, out);
  run<7>("FFMA2 + 2 FFMA (3 inst)", 3, out);
  return 0;
}
