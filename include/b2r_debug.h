/*
 * b2r_debug.h -- test hooks and A/B knobs of libb2r.so.  NOT part of the drop-in boundary (include/b2r.h): nothing a caller of the
 * reference's teaserpp_python path needs is declared here.  tests/ and tools/ use these to force the fallback paths of the
 * consistency-graph filters and to measure the tensor-core arithmetic the exactness argument relies on.
 */
#ifndef B2R_DEBUG_H_
#define B2R_DEBUG_H_

#include "b2r.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Graph filter generation and band of this handle (takes effect at the next solve):
 *   filter 0: default (tcgen05 filter v6, exact fp64 fix-up of the listed pairs); 1: always evaluate both fp64 square roots;
 *          2 / 3: the fp64 sign-test filters v2 / v1; 4 / 5: the packed-fp32 filters v3 / v4 (A/B measurements, parity tests);
 *   tau:   lower limit of the filter's ambiguity band relative to Sigma^2 (tests widen the band to force the exact path);
 *          <= 0: the band the error analysis needs. */
b2r_status b2r_debug_set_graph_filter(b2r_solver* s, int filter, double tau);

/* After a solve on this handle: recompute the raw fp32 values Dm~ and L~ (DESIGN.md "graph filter v5") of rows
 * [128 row_block, +128) x the 64 columns of unit `unit` of that row block for problem idx, as the tcgen05.mma leaves them in TMEM.
 * dm_out / l_out: host, 128 x 64 floats row-major.  consts_out (host, 12 doubles): kappa, centres (6), t1, t0, lcut, k, use32. */
b2r_status b2r_debug_tc_values(b2r_solver* s, int idx, int row_block, int unit, float* dm_out, float* l_out, double* consts_out);

/* After a solve: EVERY 128 x 64 unit of problem idx is recomputed on the tensor cores and compared on the device with (a) the
 * exact fp64 sum of the same operand products and (b) the fp64 value from the unrounded coordinates.  out8 (host):
 *   [0] max |tc - sum of products| / sum |products|  over Dm~ and L~   (the band assumes <= 2^-20)
 *   [1] max |Dm~ - Dm| / E_D      [2] max |L~ - L| / E_L               (the band assumes <= 1)
 *   [3] pairs checked             [4] Pi   [5] E_D   [6] E_L   [7] 1 if the tensor-core filter decided this problem, else 0 */
b2r_status b2r_debug_tc_sweep(b2r_solver* s, int idx, double* out8);

/* SM cycles the fused tail kernel spent on problem idx of the last batch up to the end of each stage:
 * [0] fix-up of listed pairs [1] degrees [2] greedy bound [3] core pruning [4] clique [5] rotation [6] translation
 * [7] state loaded (start of the fix-up); [8..31] finer marks inside the stages (0 where a stage did not run), see tail.cu */
b2r_status b2r_debug_tail_cycles(b2r_solver* s, int idx, uint32_t* out32);

/* With B2R_CANARY=1 in the environment every workspace buffer of a handle carries a 4 KB guard zone on either side (0xA5 bytes);
 * this call (and b2r_sync, b2r_destroy) verifies them: B2R_ERR_STATE names the buffer a kernel wrote out of bounds.  Without the
 * variable it returns B2R_OK.  tools/canary_tests.sh runs the GPU tests this way (compute-sanitizer is closed on the GPU pool). */
b2r_status b2r_debug_check_canaries(b2r_solver* s);

/* b2r_nn_gemm with the error of the tensor-core distances MEASURED: for every candidate pair (i, j) the kernel kept (8 x splits x 4 per
 * query row) the exact fp64 distance D is computed as well and compared with the approximate one.  out3 (host):
 *   [0] max |D~_ij - D_ij| / (|fq_i - c|^2 + |fr_j - c|^2)   [1] the bound the shortlist logic assumes (kEpsRel)   [2] rows that fell back */
b2r_status b2r_debug_nn_gemm_error(b2r_solver* s, const double* fq_dev, int nq, const double* fr_dev, int nr, int dim, int32_t* nn_dev,
                                   double* out3);

#ifdef __cplusplus
}
#endif
#endif /* B2R_DEBUG_H_ */
