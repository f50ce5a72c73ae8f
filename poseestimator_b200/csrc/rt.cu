// Rotation (GNC-TLS) and translation (adaptive voting) on the pruned correspondences.
//
// Replaces upstream's chain/complete TIMs over the clique + GNCTLSRotationSolver::solveForRotation (A.6) and
// TLSTranslationSolver / ScalarTLSEstimator::estimate (A.7), reached from `solver.solve`
// (registration_utils.py:255-261 sets GNC_TLS, factor 1.4, 100 iterations, cost threshold 1e-12).
// The stage bodies live in rt.cuh (shared with the fused tail kernel, tail.cu); these are the one-kernel-per-stage launches of
// the multi-kernel path (inlier modes other than PMC_EXACT, problems the fused tail hands over).
#include "rt.cuh"

namespace b2r {

__global__ void __launch_bounds__(kRtThreads) rotation_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  __shared__ double tim_cache[8 * 512];   // TIMs, weights and residuals of up to 512 pruned TIMs stay on chip during the GNC loop
  const ProbDesc d = descs[blockIdx.x];
  if (d.st->tail_done) return;
  rotation_body(d, kp, 0, TailMarks(), tim_cache, 8 * 512);
}

__global__ void __launch_bounds__(3 * kRtThreads) translation_kernel(const ProbDesc* __restrict__ descs, KernelParams kp) {
  extern __shared__ double shd[];
  const ProbDesc d = descs[blockIdx.x];
  if (d.st->tail_done) return;
  translation_body(d, kp, shd);
}

int launch_rotation_translation(const ProbDesc* d_desc, int batch, int n_max, const KernelParams& kp, cudaStream_t st,
                                cudaEvent_t* ev) {
  static PerDeviceOnce attr;
  const size_t tsmem = 3 * kTrGroupDoubles * sizeof(double);
  attr.run([&] { return cudaFuncSetAttribute(translation_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tsmem); });
  rotation_kernel<<<batch, kRtThreads, 0, st>>>(d_desc, kp);
  if (ev) cudaEventRecord(ev[0], st);
  translation_kernel<<<batch, 3 * kRtThreads, tsmem, st>>>(d_desc, kp);
  (void)n_max;
  return 2;
}

}  // namespace b2r
