/*
 * b2r.h -- C ABI of the B200-native registration library (libb2r.so).
 *
 * This is the drop-in boundary for the registration stage of patricksko/PoseEstimator.
 * The reference reaches this path through the pybind11 module `teaserpp_python`
 * (TEASER++ C++/OpenMP) and through Open3D; each entry point below names the reference
 * call it replaces.  Plain pointers and sizes only: no torch / numpy / C++ types.
 *
 *   reference call site (file:line under /root/reference/src/pose_estimator/EstimHelpers)
 *   ------------------------------------------------------------------------------------
 *   teaserpp_python.RobustRegistrationSolver.Params()      HelpersRealtime.py:158, registration_utils.py:251
 *   teaserpp_python.RobustRegistrationSolver(params)       HelpersRealtime.py:165, registration_utils.py:260
 *   solver.solve(src_corr, dst_corr)                       HelpersRealtime.py:166, registration_utils.py:261
 *   solver.getSolution()                                   HelpersRealtime.py:167, registration_utils.py:262
 *   solver.getRotationInliers()/getScaleInliers()/getTranslationInliers()   registration_utils.py:264-266
 *   depth * scale, mask, RGBD -> cloud (back-projection)   RealSenseClass.py:56-67, registration_utils.py:54-90
 *   remove_statistical_outlier(20, 1.0)                    RealSenseClass.py:68, registration_utils.py:93
 *   voxel_down_sample(voxel)                               PoseEstimator.py:87,93
 *   estimate_normals + compute_fpfh_feature                HelpersRealtime.py:215-224
 *   get_correspondences (north-star: mutual NN on FPFH)    HelpersRealtime.py:194-213
 *
 * Conventions
 *   - every function returns a b2r_status (0 = ok), never throws, never calls exit();
 *   - a handle is bound to one CUDA device and one stream; thread-compatible per handle,
 *     thread-safe across handles;
 *   - point sets are 3xN float64, column-major (x,y,z of point i at 3*i..3*i+2) unless strides are given;
 *   - there is NO CPU fallback: every entry point fails with B2R_ERR_CUDA when no device is usable;
 *   - on_device = 1 pointers must live on the handle's device (the library cannot check a raw pointer's device cheaply;
 *     the Python host raises TypeError for tensors of another device).
 */
#ifndef B2R_H_
#define B2R_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2R_VERSION 200 /* 0.2.0 */

typedef struct b2r_solver b2r_solver; /* opaque */

typedef enum b2r_status {
  B2R_OK = 0,
  B2R_ERR_INVALID_ARG = 1,
  B2R_ERR_CUDA = 2,     /* CUDA runtime error; text in b2r_last_error() */
  B2R_ERR_OOM = 3,
  B2R_ERR_STATE = 4,    /* e.g. getter before solve, index out of range */
  B2R_ERR_LIMIT = 5     /* problem exceeds a documented limit (N > 65535, search arena overflow) */
} b2r_status;

/* enum values follow TEASER++ */
enum { B2R_ROT_GNC_TLS = 0, B2R_ROT_FGR = 1 };
enum { B2R_TIM_CHAIN = 0, B2R_TIM_COMPLETE = 1 };
enum { B2R_INLIER_PMC_EXACT = 0, B2R_INLIER_PMC_HEU = 1, B2R_INLIER_KCORE_HEU = 2, B2R_INLIER_NONE = 3 };

/* RobustRegistrationSolver::Params -- same fields, same defaults (b2r_default_params). */
typedef struct b2r_params {
  double noise_bound;                /* 0.01 */
  double cbar2;                      /* 1.0 */
  int32_t estimate_scaling;          /* 1 (upstream default); the reference sets 0 */
  int32_t rotation_estimation_algorithm; /* B2R_ROT_GNC_TLS */
  double rotation_gnc_factor;        /* 1.4 */
  int32_t rotation_max_iterations;   /* 100 */
  double rotation_cost_threshold;    /* 1e-6 */
  int32_t rotation_tim_graph;        /* B2R_TIM_CHAIN */
  int32_t inlier_selection_mode;     /* B2R_INLIER_PMC_EXACT */
  double kcore_heuristic_threshold;  /* 0.5 */
  double max_clique_time_limit;      /* 3000 s; exact search stops (best-so-far kept, stats.exact_proven=0) */
  int32_t max_clique_num_threads;    /* accepted for API parity; ignored on the GPU */
  /* --- extensions (0 = default behaviour) --- */
  double rotation_noise_bound;       /* GNC noise bound before the 2/scale factor; <=0 -> noise_bound.
                                        Lets the host mirror upstream's in-place compounding (A.6-2). */
  int32_t reserved0;                 /* must be 0 (test hooks live in b2r_debug.h: b2r_debug_set_graph_filter) */
  double reserved1;                  /* must be 0 */
} b2r_params;

typedef struct b2r_solution {
  double scale;
  double rotation[9];   /* row-major 3x3 */
  double translation[3];
  int32_t valid;
} b2r_solution;

typedef struct b2r_stats {
  int32_t n;                 /* correspondences */
  int64_t n_edges;           /* edges of the consistency graph */
  int32_t max_degree;
  int32_t heuristic_lb;      /* clique size found by the greedy heuristic */
  int32_t n_pruned;          /* vertices left after (lb-1)-core pruning */
  int32_t omega;             /* maximum clique size proven by the exact search */
  int32_t clique_size;
  int32_t n_rot_tims;
  int32_t n_rot_inliers;
  int32_t n_trans_inliers;
  int32_t gnc_iterations;
  double gnc_cost;
  int64_t bnb_nodes;         /* branch-and-bound nodes expanded */
  int64_t graph_slow_pairs;  /* pairs the sqrt-free filter sent to the exact fp64 path */
  int32_t exact_proven;      /* 1 unless the time limit stopped the exact search */
  int32_t flags;             /* internal error bits, 0 when ok */
  int32_t fused_tail;        /* 1: everything after the graph ran in the fused tail kernel; 0: multi-kernel path (other inlier
                                modes, or handed over: exact search needed / graph redone in fp64) */
  int32_t pad_;
} b2r_stats;

typedef enum b2r_index_kind {
  B2R_IDX_MAX_CLIQUE = 0,          /* original correspondence ids, ascending            (getInlierMaxClique) */
  B2R_IDX_ROT_INLIERS = 1,         /* positions in the pruned TIM list with w >= 0.5    (getRotationInliers) */
  B2R_IDX_ROT_INLIERS_MAP = 2,     /* max_clique[k] of those positions                  (getRotationInliersMap) */
  B2R_IDX_TRANS_INLIERS = 3,       /* positions in the rotation-inlier list             (getTranslationInliers) */
  B2R_IDX_TRANS_INLIERS_MAP = 4,   /* original ids of those                             (getTranslationInliersMap) */
  B2R_IDX_DEGREES = 5              /* degree of every vertex, length n */
} b2r_index_kind;

void b2r_default_params(b2r_params* p);
int b2r_version(void);
/* number of usable CUDA devices (0 when none: the product then refuses to run) */
int b2r_device_count(void);

/* RobustRegistrationSolver(params).  `stream` is a cudaStream_t (NULL = a private non-blocking stream). */
b2r_status b2r_create(const b2r_params* params, int device, void* stream, b2r_solver** out);
b2r_status b2r_destroy(b2r_solver* s);
/* solver.reset(params) */
b2r_status b2r_reset(b2r_solver* s, const b2r_params* params);
const char* b2r_last_error(const b2r_solver* s);

/* solver.solve(src, dst): one registration, synchronous.
 * element (r, c) of a 3xN operand is at base[r*row_stride + c*col_stride] (strides in elements);
 * on_device != 0: pointers are device pointers on the handle's device. */
b2r_status b2r_solve(b2r_solver* s, const double* src, const double* dst, int n,
                     int64_t row_stride, int64_t col_stride, int on_device);

/* Batch of independent registrations (detections / templates / frames).  Each operand is packed 3 x n[b]
 * column-major.  _async only enqueues on the handle's stream (inputs must stay valid until b2r_sync).
 * on_device: 0 = ordinary host memory (copied through the handle's page-locked staging block), 1 = device pointers,
 * 2 = page-locked host memory (cudaHostAlloc / cudaHostRegister / b2r_alloc_pinned): the copy engine reads the
 * caller's buffers directly, operands adjacent in memory are merged into one copy. */
b2r_status b2r_solve_batch(b2r_solver* s, int batch, const int* n, const double* const* src,
                           const double* const* dst, int on_device);
b2r_status b2r_solve_batch_async(b2r_solver* s, int batch, const int* n, const double* const* src,
                                 const double* const* dst, int on_device);
b2r_status b2r_sync(b2r_solver* s);
/* page-locked host memory for on_device == 2 callers that do not link the CUDA runtime themselves */
b2r_status b2r_alloc_pinned(size_t bytes, void** out);
b2r_status b2r_free_pinned(void* p);
/* Fill page-locked memory with streaming stores (no CUDA call).  On the B200 hosts the copy engine reads lines that
 * are still dirty in CPU caches 4x slower than lines in DRAM (profiles/r01_h2d_pack.jsonl): producers that write
 * their correspondences into a pinned block should do it through this (or their own non-temporal stores). */
b2r_status b2r_copy_to_pinned(void* dst, const void* src, size_t bytes);
int b2r_batch_size(const b2r_solver* s);
/* number of kernels this handle has launched since creation (bench.py's gpu_launches claim) */
int64_t b2r_launch_count(const b2r_solver* s);

/* getSolution() and friends; `idx` selects the problem of the last batch (0 for b2r_solve). */
b2r_status b2r_get_solution(b2r_solver* s, int idx, b2r_solution* out);
/* the first `count` solutions of the last batch in one call (out: array of count b2r_solution) */
b2r_status b2r_get_solutions(b2r_solver* s, b2r_solution* out, int count);
b2r_status b2r_get_stats(b2r_solver* s, int idx, b2r_stats* out);
/* out == NULL: only *count is written. */
b2r_status b2r_get_indices(b2r_solver* s, int idx, b2r_index_kind kind, int32_t* out, int32_t* count);
/* consistency graph as bit rows: out[i*words_per_row + (j>>5)] bit (j&31); words_per_row >= ceil(n/32) */
b2r_status b2r_get_adjacency(b2r_solver* s, int idx, uint32_t* out, int words_per_row);
/* GNC weights of the pruned TIMs (length n_rot_tims) */
b2r_status b2r_get_gnc_weights(b2r_solver* s, int idx, double* out);
/* per-stage device times of the last batch in ms (needs b2r_set_profiling(s,1)):
 * [0] upload [1] graph [2] degree [3] heuristic [4] prune+compact [5] exact search [6] canonical clique
 * [7] rotation [8] translation [9] total [10] scale vote (estimate_scaling=True only)
 * [11] fused tail: on the default PMC_EXACT path everything after the graph is ONE kernel (exact fix-up of the listed pairs,
 *      degrees, greedy bound, core pruning, rotation, translation); [2] .. [8] are then 0 */
b2r_status b2r_set_profiling(b2r_solver* s, int on);
b2r_status b2r_get_timings(b2r_solver* s, float* ms12);

/* ---- pieces of the path exposed for sharding and tests -------------------------------------------- */
/* Rows [row_begin,row_end) of the consistency graph of one problem into adj_out (device, pitch words per
 * row, pitch % 4 == 0).  src/dst: device, packed 3xN.  Used by the row-sharded large-N path (config 5). */
b2r_status b2r_graph_rows(b2r_solver* s, const double* src_dev, const double* dst_dev, int n,
                          int row_begin, int row_end, uint32_t* adj_out_dev, int pitch_words);
/* Continue a solve from a complete device adjacency (after the all-gather of row shards). */
b2r_status b2r_solve_from_adjacency(b2r_solver* s, const double* src_dev, const double* dst_dev, int n,
                                    const uint32_t* adj_dev, int pitch_words);
/* words per adjacency row the library uses for n vertices (multiple of 4) */
int b2r_pitch_words(int n);

/* ---- one very large registration over the GPUs of one box, adjacency exchanged by NVLink peer stores (config 5) ------------------
 * One process per GPU.  Every rank: b2r_shard_export (sizes its workspace for ONE problem of n correspondences and returns CUDA IPC
 * handles of its bit matrix and flag words), exchanges the b2r_peer_handle structs with the other ranks by any means (the Python
 * host uses torch.distributed.all_gather_object), b2r_shard_attach (maps the peers' buffers; needs peer access between the GPUs),
 * then any number of b2r_solve_sharded calls -- COLLECTIVE: every rank calls it with the same replicated operands.  Rank r runs
 * 1 / world of the tensor-core graph filter and stores its words into every rank's matrix while it computes; each rank then finishes
 * the solve on its own complete matrix (results identical on all ranks and to b2r_solve, getters as usual).  world <= 8.
 * The handle must not be used for larger problems or batches in between (a workspace re-allocation invalidates the export).
 * b2r_shard_attach resets this rank's flag words: the host must put a barrier between the attach of ALL ranks and the first
 * b2r_solve_sharded of any rank (the Python host's all-or-nothing error exchange is that barrier).  A peer that never arrives makes
 * the call fail with B2R_ERR_STATE after ~2 s (B2R_SHARD_TIMEOUT_MS) instead of hanging the GPU. */
typedef struct b2r_peer_handle {
  unsigned char adjacency[64];   /* cudaIpcMemHandle_t of the allocation that holds the bit matrix */
  unsigned char flags[64];       /* cudaIpcMemHandle_t of the allocation that holds the flag words */
  int64_t adj_offset, flag_offset;
  int32_t n, device;
} b2r_peer_handle;
b2r_status b2r_shard_export(b2r_solver* s, int n, b2r_peer_handle* out);
b2r_status b2r_shard_attach(b2r_solver* s, int rank, int world, const b2r_peer_handle* all_ranks /* [world] */);
b2r_status b2r_solve_sharded(b2r_solver* s, const double* src_dev, const double* dst_dev, int n);
/* 1 when the last b2r_solve_sharded fell back to every rank solving alone (a rank's filter refused the problem) */
int b2r_shard_last_fell_back(const b2r_solver* s);
b2r_status b2r_shard_detach(b2r_solver* s);


/* ---- pre-steps: what the reference does with Open3D between the detector's mask and solver.solve -----------------
 * All point / feature buffers are DEVICE pointers on the handle's device unless stated; outputs are written in the
 * canonical orders of oracle/presteps_oracle.cpp (row-major pixel order, ascending voxel key, neighbours by
 * (distance^2, index)).  Calls are synchronous where they return a count to the host. */
typedef struct b2r_camera {
  double fx, fy, cx, cy;  /* pinhole intrinsics (RealSenseClass.py:45-51) */
  double depth_scale;     /* metres per depth unit; applied in float32 like the reference's numpy expression */
  double depth_trunc;     /* Open3D RGBDImage depth_trunc (3.0 on the camera path); <= 0: no truncation */
  double zmin, zmax;      /* registration_utils.py:75 keeps zmin < z < zmax; both <= 0: no range test */
} b2r_camera;

/* depth u16 [H,W] (+ mask u8 [H,W], 0 = outside, may be NULL) -> points f64 [M,3] in row-major scan order inside the
 * bbox [u0,u1) x [v0,v1); pixel_ids (v*W+u, may be NULL).  on_device == 0: depth/mask are host images.
 * Replaces RealSenseClass.get_pcd_from_rgbd (RealSenseClass.py:56-67) / get_pointcloud (registration_utils.py:54-90). */
b2r_status b2r_backproject(b2r_solver* s, const uint16_t* depth, const uint8_t* mask, int H, int W, int u0, int v0, int u1,
                           int v1, const b2r_camera* cam, int on_device, double* points_dev, int32_t* pixel_ids_dev,
                           int capacity, int* count_out);
/* KDTreeFlann hybrid / knn search for every query: the max_nn (<= 128) nearest points, then d2 < radius^2 when
 * radius > 0.  idx/d2: [nq, max_nn] (padding -1 / 0), cnt: [nq]. */
b2r_status b2r_knn(b2r_solver* s, const double* pts_dev, int n, const double* queries_dev, int nq, int max_nn, double radius,
                   int32_t* idx_dev, double* d2_dev, int32_t* cnt_dev);
/* remove_statistical_outlier(nb_neighbors, std_ratio) (RealSenseClass.py:68, registration_utils.py:93).
 * keep_dev [n] / avg_dev [n] / pts_out_dev [kept,3] / idx_out_dev [kept] are optional (NULL). */
b2r_status b2r_statistical_outlier(b2r_solver* s, const double* pts_dev, int n, int nb_neighbors, double std_ratio,
                                   uint8_t* keep_dev, double* avg_dev, double* pts_out_dev, int32_t* idx_out_dev, int* kept_out);
/* voxel_down_sample(voxel) (PoseEstimator.py:87,93): out_dev [V,3] means in ascending voxel key order;
 * voxel_of_point_dev [n], counts_dev [V] optional. */
b2r_status b2r_voxel_down(b2r_solver* s, const double* pts_dev, int n, double voxel, double* out_dev,
                          int32_t* voxel_of_point_dev, int32_t* counts_dev, int* n_voxels_out);
/* estimate_normals(KDTreeSearchParamHybrid(radius, max_nn)) (HelpersRealtime.py:216-219).
 * orient: 0 none, 1 towards view3, 2 away from view3. */
b2r_status b2r_normals(b2r_solver* s, const double* pts_dev, int n, double radius, int max_nn, int orient, const double* view3,
                       double* normals_dev);
/* compute_fpfh_feature(KDTreeSearchParamHybrid(radius, max_nn)) (HelpersRealtime.py:221-224): fpfh_dev [33,n]
 * (Open3D Feature.data layout); spfh_dev [33,n] optional. */
b2r_status b2r_fpfh(b2r_solver* s, const double* pts_dev, const double* normals_dev, int n, double radius, int max_nn,
                    double* fpfh_dev, double* spfh_dev);
/* exact nearest neighbour in feature space (rows = descriptors, [n, dim]): fp64 direct differences, lowest index
 * wins ties.  dist_dev optional. */
b2r_status b2r_nn_exact(b2r_solver* s, const double* fq_dev, int nq, const double* fr_dev, int nr, int dim, int32_t* nn_dev,
                        double* dist_dev);
/* the same result as b2r_nn_exact, found on the tensor cores: 3xTF32 tcgen05 GEMM fed by TMA keeps a per-row top-4
 * shortlist of |fr_j|^2 - 2 fq_i.fr_j, then an exact fp64 re-rank in the oracle's order; rows whose shortlist cannot be
 * proven complete fall back to exact brute force (*n_fallback_out, optional, counts them).  dim <= 40, else the exact
 * kernel runs.  Replaces the feature matching of get_correspondences (HelpersRealtime.py:194-213). */
b2r_status b2r_nn_gemm(b2r_solver* s, const double* fq_dev, int nq, const double* fr_dev, int nr, int dim, int32_t* nn_dev,
                       double* dist_dev, int* n_fallback_out);
/* both directions + mutual filter in one call: nn01 [na], nn10 [nb], pairs [c,2] ascending i */
b2r_status b2r_mutual_nn(b2r_solver* s, const double* fa_dev, int na, const double* fb_dev, int nb, int dim, int32_t* nn01_dev,
                         int32_t* nn10_dev, int32_t* pairs_dev, int* n_pairs_out, int* n_fallback_out);
/* pairs (i, nn01[i]) with nn10[nn01[i]] == i, ascending i -> pairs_dev [c,2] */
b2r_status b2r_mutual_pairs(b2r_solver* s, const int32_t* nn01_dev, int na, const int32_t* nn10_dev, int nb, int32_t* pairs_dev,
                            int* n_pairs_out);
/* the 3xN operands of solver.solve from matched pairs (HelpersRealtime.py:142-143): src = a_pts[pairs[:,0]], dst = b_pts[pairs[:,1]] */
b2r_status b2r_gather_correspondences(b2r_solver* s, const double* a_pts_dev, const double* b_pts_dev, const int32_t* pairs_dev,
                                      int n_pairs, double* src_dev, double* dst_dev);

/* ---- refinement and scoring right after solver.solve (what the reference does per template, PoseEstimator.py:97-108) ------
 * Point buffers are DEVICE pointers ([n,3] float64); 4x4 matrices are HOST, row-major.  All calls are synchronous. */
typedef struct b2r_icp_result {
  double transformation[16];  /* icp_result.transformation */
  double fitness;             /* correspondences / source points */
  double inlier_rmse;
  int32_t n_corr;
  int32_t iterations;         /* loop iterations executed */
  int32_t converged;          /* 1: relative fitness / rmse criterion met; 0: stopped by max_iteration */
  int32_t pad_;
} b2r_icp_result;

/* pcd.transform(T): p' = T * (x, y, z, 1) */
b2r_status b2r_transform_points(b2r_solver* s, const double* T16, const double* in_dev, int n, double* out_dev);
/* nearest reference point of every query (exact fp64, lowest index on ties): squared distance and / or index.
 * cKDTree(ref).query(q, k=1) (HelpersRealtime.py:103-104), pcd.compute_point_cloud_distance (registration_utils.py:277-278). */
b2r_status b2r_nn_distance(b2r_solver* s, const double* q_dev, int nq, const double* ref_dev, int nr, double* dist2_dev, int32_t* idx_dev);
/* o3d.pipelines.registration.registration_icp(src, dst, max_dist, init, TransformationEstimationPointToPoint(),
 * ICPConvergenceCriteria(relative_fitness, relative_rmse, max_iteration)) (PoseEstimator.py:97-101, main_realsense.py:112-115).
 * init16 NULL = identity; corr_dev [ns] optional: nearest target index of every source point, -1 outside max_dist. */
b2r_status b2r_icp(b2r_solver* s, const double* src_dev, int ns, const double* dst_dev, int nd, double max_dist, const double* init16,
                   int max_iter, double rel_fitness, double rel_rmse, b2r_icp_result* out, int32_t* corr_dev);
/* voxel_coverage (HelpersRealtime.py:108-110): distinct floor(p / voxel) int32 triples */
b2r_status b2r_voxel_coverage(b2r_solver* s, const double* pts_dev, int n, double voxel, int* count_out);
/* b2r_voxel_coverage without any allocation.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same count as
 * b2r_voxel_coverage, but the hash table is a file-scope static device array (refine.cu, 2 * B2R_VOXEL_NOALLOC_MAX_POINTS int32 =
 * 4 MiB, held by every GPU that loads the library until exit) instead of the solver's growable workspace, and the result is written
 * to DEVICE memory: count_dev is one int32 on s's device, valid once the work queued on s's stream has finished (the call only
 * enqueues: two memsets and one kernel; no cudaMalloc / cudaFree / allocating library routine, no synchronisation).
 * n > B2R_VOXEL_NOALLOC_MAX_POINTS fails with B2R_ERR_LIMIT.  All calls on one device share the table: calls that overlap on the
 * GPU (different streams or handles of the same device) race; serialise them. */
#define B2R_VOXEL_NOALLOC_MAX_POINTS 524288
b2r_status b2r_voxel_coverage_noalloc(b2r_solver* s, const double* pts_dev, int n, double voxel, int* count_dev);
/* alignment_score (HelpersRealtime.py:112-130): out5 = {median, p90, cov_aligned, cov_full, score} (numpy median / percentile) */
b2r_status b2r_alignment_score(b2r_solver* s, const double* src_aligned_dev, int ns, const double* src_down_dev, int ns2,
                               const double* dst_dev, int nd, double voxel, double* out5);
/* chamfer_distance (registration_utils.py:276-280) */
b2r_status b2r_chamfer(b2r_solver* s, const double* a_dev, int na, const double* b_dev, int nb, double* out);
/* cloud_resolution(pcd, k) (registration_utils.py:229-240) */
b2r_status b2r_cloud_resolution(b2r_solver* s, const double* pts_dev, int n, int k, double* out);
/* pca_axes (registration_utils.py:290-310): R9 columns = principal axes by decreasing variance (det +1), s3 = sqrt(eigenvalues),
 * c3 = centroid; s3 / c3 optional */
b2r_status b2r_pca_axes(b2r_solver* s, const double* pts_dev, int n, double* R9, double* s3, double* c3);
/* initial_align_centroid_pca (registration_utils.py:314-343) -> T0 (src -> dst) */
b2r_status b2r_pca_align(b2r_solver* s, const double* src_dev, int ns, const double* dst_dev, int nd, double* T16);

/* ---- CAD depth render (replaces the Filament OffscreenRenderer of PoseEstimator.py:118-141 / template_creation.py:188-233) ---
 * verts_dev [nv,3] f64 model frame, tris_dev [nt,3] i32, T16 = model -> camera (HOST, row-major; OpenCV camera: x right, y down,
 * z forward).  depth_dev [H,W] f32: view-space z of the nearest surface on the ray ((u-cx)/fx, (v-cy)/fy, 1) of every pixel with
 * z_near < z < z_far, 0 where there is none (render_to_depth_image(z_in_view_space=True) after depth_trunc). */
b2r_status b2r_render_depth(b2r_solver* s, const double* verts_dev, int nv, const int32_t* tris_dev, int nt, const double* T16, double fx,
                            double fy, double cx, double cy, int W, int H, double z_near, double z_far, float* depth_dev);
/* RGBDImage.create_from_color_and_depth(depth_scale, depth_trunc) + PointCloud.create_from_rgbd_image on a FLOAT depth image
 * (PoseEstimator.py:130-135, template_creation.py:230-233): points [M,3] f64 in row-major pixel order, pixel ids optional. */
b2r_status b2r_backproject_depth32(b2r_solver* s, const float* depth_dev, int H, int W, double fx, double fy, double cx, double cy,
                                   double depth_scale, double depth_trunc, double* points_dev, int32_t* pixel_ids_dev, int capacity,
                                   int* count_out);

/* LiDAR cloud -> image selection (main_seibersdorf.py:70-81,149-162): Xc = R p + t, keep z > z_min, project with K (row-major 3x3)
 * and the OpenCV distortion D (4, 5 or 8 coefficients, else ignored) like cv2.projectPoints on float32 input, truncate to int32,
 * keep pixels inside the W x H image whose mask byte is non-zero (mask_dev NULL: no mask).  idx_dev [<= n]: surviving point ids in
 * ascending order, uv_dev [<= n, 2] their pixels (optional).  R9, t3, K9, D are HOST arrays. */
b2r_status b2r_project_select(b2r_solver* s, const double* pts_dev, int n, const double* R9, const double* t3, const double* K9, const double* D,
                              int nD, int W, int H, const uint8_t* mask_dev, double z_min, int32_t* idx_dev, int32_t* uv_dev, int* n_front_out,
                              int* n_in_image_out, int* n_selected_out);

/* uniform_downsample_farthest_point (registration_utils.py:98-128) started at first_index (the reference draws it with
 * np.random.randint): idx_dev [target] point ids in selection order. */
b2r_status b2r_farthest_point_sample(b2r_solver* s, const double* pts_dev, int n, int target, int first_index, int32_t* idx_dev);

#ifdef __cplusplus
}
#endif
#endif /* B2R_H_ */
