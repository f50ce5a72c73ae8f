"""ctypes binding of libb2r.so (include/b2r.h).  No CPU fallback: a missing library or device is an error."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2R_LIB") or os.path.join(_HERE, "csrc", "libb2r.so")  # B2R_LIB: A/B builds of the same library (tools/)


class B2RParams(C.Structure):
    _fields_ = [
        ("noise_bound", C.c_double),
        ("cbar2", C.c_double),
        ("estimate_scaling", C.c_int32),
        ("rotation_estimation_algorithm", C.c_int32),
        ("rotation_gnc_factor", C.c_double),
        ("rotation_max_iterations", C.c_int32),
        ("rotation_cost_threshold", C.c_double),
        ("rotation_tim_graph", C.c_int32),
        ("inlier_selection_mode", C.c_int32),
        ("kcore_heuristic_threshold", C.c_double),
        ("max_clique_time_limit", C.c_double),
        ("max_clique_num_threads", C.c_int32),
        ("rotation_noise_bound", C.c_double),
        ("reserved0", C.c_int32),
        ("reserved1", C.c_double),
    ]


class B2RCamera(C.Structure):
    _fields_ = [("fx", C.c_double), ("fy", C.c_double), ("cx", C.c_double), ("cy", C.c_double), ("depth_scale", C.c_double),
                ("depth_trunc", C.c_double), ("zmin", C.c_double), ("zmax", C.c_double)]


class B2RSolution(C.Structure):
    _fields_ = [("scale", C.c_double), ("rotation", C.c_double * 9), ("translation", C.c_double * 3), ("valid", C.c_int32)]


class B2RIcpResult(C.Structure):
    _fields_ = [("transformation", C.c_double * 16), ("fitness", C.c_double), ("inlier_rmse", C.c_double), ("n_corr", C.c_int32),
                ("iterations", C.c_int32), ("converged", C.c_int32), ("pad_", C.c_int32)]


class B2RStats(C.Structure):
    _fields_ = [
        ("n", C.c_int32), ("n_edges", C.c_int64), ("max_degree", C.c_int32), ("heuristic_lb", C.c_int32),
        ("n_pruned", C.c_int32), ("omega", C.c_int32), ("clique_size", C.c_int32), ("n_rot_tims", C.c_int32),
        ("n_rot_inliers", C.c_int32), ("n_trans_inliers", C.c_int32), ("gnc_iterations", C.c_int32),
        ("gnc_cost", C.c_double), ("bnb_nodes", C.c_int64), ("graph_slow_pairs", C.c_int64),
        ("exact_proven", C.c_int32), ("flags", C.c_int32), ("fused_tail", C.c_int32), ("pad_", C.c_int32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


# every symbol include/b2r.h declares (tests check the library exports all of them)
EXPORTS = [
    "b2r_default_params", "b2r_version", "b2r_device_count", "b2r_create", "b2r_destroy", "b2r_reset", "b2r_last_error",
    "b2r_solve", "b2r_solve_batch", "b2r_solve_batch_async", "b2r_sync", "b2r_alloc_pinned", "b2r_free_pinned", "b2r_copy_to_pinned", "b2r_batch_size", "b2r_launch_count",
    "b2r_get_solution", "b2r_get_solutions", "b2r_get_stats", "b2r_get_indices", "b2r_get_adjacency", "b2r_get_gnc_weights",
    "b2r_set_profiling", "b2r_get_timings", "b2r_graph_rows", "b2r_solve_from_adjacency", "b2r_pitch_words",
    "b2r_backproject", "b2r_knn", "b2r_statistical_outlier", "b2r_voxel_down", "b2r_normals", "b2r_fpfh", "b2r_nn_exact",
    "b2r_mutual_pairs", "b2r_gather_correspondences", "b2r_nn_gemm", "b2r_mutual_nn",
    "b2r_transform_points", "b2r_nn_distance", "b2r_icp", "b2r_voxel_coverage", "b2r_voxel_coverage_noalloc", "b2r_alignment_score", "b2r_chamfer",
    "b2r_cloud_resolution", "b2r_pca_axes", "b2r_pca_align", "b2r_render_depth", "b2r_backproject_depth32", "b2r_project_select", "b2r_farthest_point_sample",
    "b2r_shard_export", "b2r_shard_attach", "b2r_solve_sharded", "b2r_shard_last_fell_back", "b2r_shard_detach",
]

# include/b2r_debug.h: test hooks, not part of the drop-in boundary
DEBUG_EXPORTS = ["b2r_debug_set_graph_filter", "b2r_debug_tc_values", "b2r_debug_tc_sweep", "b2r_debug_tail_cycles",
                 "b2r_debug_nn_gemm_error", "b2r_debug_check_canaries"]

PEER_HANDLE_BYTES = 64 + 64 + 8 + 8 + 4 + 4   # sizeof(b2r_peer_handle)

# numpy view of an array of b2r_solution (112 bytes each)
import numpy as _np  # noqa: E402
SOLUTION_DTYPE = _np.dtype({"names": ["scale", "rotation", "translation", "valid"], "formats": ["<f8", ("<f8", (3, 3)), ("<f8", (3,)), "<i4"],
                            "offsets": [0, 8, 80, 104], "itemsize": 112})

IDX_MAX_CLIQUE, IDX_ROT_INLIERS, IDX_ROT_INLIERS_MAP, IDX_TRANS_INLIERS, IDX_TRANS_INLIERS_MAP, IDX_DEGREES = range(6)

_lib = None


def lib():
    """Load libb2r.so; raises (never falls back) when the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C poseestimator_b200/csrc`). There is no CPU fallback."
        )
    L = C.CDLL(LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
    L.b2r_default_params.argtypes = [C.POINTER(B2RParams)]
    L.b2r_default_params.restype = None
    L.b2r_version.restype = i32
    L.b2r_device_count.restype = i32
    L.b2r_create.argtypes = [C.POINTER(B2RParams), i32, vp, C.POINTER(vp)]
    L.b2r_destroy.argtypes = [vp]
    L.b2r_reset.argtypes = [vp, C.POINTER(B2RParams)]
    L.b2r_last_error.argtypes = [vp]
    L.b2r_last_error.restype = C.c_char_p
    L.b2r_solve.argtypes = [vp, vp, vp, i32, i64, i64, i32]
    L.b2r_solve_batch.argtypes = [vp, i32, vp, vp, vp, i32]
    L.b2r_solve_batch_async.argtypes = [vp, i32, vp, vp, vp, i32]
    L.b2r_sync.argtypes = [vp]
    L.b2r_alloc_pinned.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
    L.b2r_free_pinned.argtypes = [C.c_void_p]
    L.b2r_copy_to_pinned.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    L.b2r_batch_size.argtypes = [vp]
    L.b2r_launch_count.argtypes = [vp]
    L.b2r_launch_count.restype = i64
    L.b2r_get_solution.argtypes = [vp, i32, C.POINTER(B2RSolution)]
    L.b2r_get_solutions.argtypes = [vp, vp, i32]
    L.b2r_get_stats.argtypes = [vp, i32, C.POINTER(B2RStats)]
    L.b2r_get_indices.argtypes = [vp, i32, i32, vp, C.POINTER(C.c_int32)]
    L.b2r_get_adjacency.argtypes = [vp, i32, vp, i32]
    L.b2r_get_gnc_weights.argtypes = [vp, i32, vp]
    L.b2r_set_profiling.argtypes = [vp, i32]
    L.b2r_get_timings.argtypes = [vp, vp]
    L.b2r_graph_rows.argtypes = [vp, vp, vp, i32, i32, i32, vp, i32]
    L.b2r_solve_from_adjacency.argtypes = [vp, vp, vp, i32, vp, i32]
    L.b2r_debug_tc_values.argtypes = [vp, i32, i32, i32, vp, vp, vp]
    L.b2r_debug_set_graph_filter.argtypes = [vp, i32, C.c_double]
    L.b2r_debug_tc_sweep.argtypes = [vp, i32, vp]
    L.b2r_debug_tail_cycles.argtypes = [vp, i32, vp]
    L.b2r_debug_nn_gemm_error.argtypes = [vp, vp, i32, vp, i32, i32, vp, vp]
    L.b2r_debug_check_canaries.argtypes = [vp]
    L.b2r_shard_export.argtypes = [vp, i32, vp]
    L.b2r_shard_attach.argtypes = [vp, i32, i32, vp]
    L.b2r_solve_sharded.argtypes = [vp, vp, vp, i32]
    L.b2r_shard_last_fell_back.argtypes = [vp]
    L.b2r_shard_last_fell_back.restype = i32
    L.b2r_shard_detach.argtypes = [vp]
    L.b2r_pitch_words.argtypes = [i32]
    L.b2r_pitch_words.restype = i32
    d, ip = C.c_double, C.POINTER(C.c_int)
    L.b2r_backproject.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i32, C.POINTER(B2RCamera), i32, vp, vp, i32, ip]
    L.b2r_knn.argtypes = [vp, vp, i32, vp, i32, i32, d, vp, vp, vp]
    L.b2r_statistical_outlier.argtypes = [vp, vp, i32, i32, d, vp, vp, vp, vp, ip]
    L.b2r_voxel_down.argtypes = [vp, vp, i32, d, vp, vp, vp, ip]
    L.b2r_normals.argtypes = [vp, vp, i32, d, i32, i32, vp, vp]
    L.b2r_fpfh.argtypes = [vp, vp, vp, i32, d, i32, vp, vp]
    L.b2r_nn_exact.argtypes = [vp, vp, i32, vp, i32, i32, vp, vp]
    L.b2r_mutual_pairs.argtypes = [vp, vp, i32, vp, i32, vp, ip]
    L.b2r_gather_correspondences.argtypes = [vp, vp, vp, vp, i32, vp, vp]
    L.b2r_nn_gemm.argtypes = [vp, vp, i32, vp, i32, i32, vp, vp, ip]
    L.b2r_mutual_nn.argtypes = [vp, vp, i32, vp, i32, i32, vp, vp, vp, ip, ip]
    L.b2r_transform_points.argtypes = [vp, vp, vp, i32, vp]
    L.b2r_nn_distance.argtypes = [vp, vp, i32, vp, i32, vp, vp]
    L.b2r_icp.argtypes = [vp, vp, i32, vp, i32, d, vp, i32, d, d, C.POINTER(B2RIcpResult), vp]
    L.b2r_voxel_coverage.argtypes = [vp, vp, i32, d, ip]
    L.b2r_voxel_coverage_noalloc.argtypes = [vp, vp, i32, d, vp]
    L.b2r_alignment_score.argtypes = [vp, vp, i32, vp, i32, vp, i32, d, vp]
    L.b2r_chamfer.argtypes = [vp, vp, i32, vp, i32, C.POINTER(C.c_double)]
    L.b2r_cloud_resolution.argtypes = [vp, vp, i32, i32, C.POINTER(C.c_double)]
    L.b2r_pca_axes.argtypes = [vp, vp, i32, vp, vp, vp]
    L.b2r_pca_align.argtypes = [vp, vp, i32, vp, i32, vp]
    L.b2r_render_depth.argtypes = [vp, vp, i32, vp, i32, vp, d, d, d, d, i32, i32, d, d, vp]
    L.b2r_project_select.argtypes = [vp, vp, i32, vp, vp, vp, vp, i32, i32, i32, vp, d, vp, vp, ip, ip, ip]
    L.b2r_farthest_point_sample.argtypes = [vp, vp, i32, i32, i32, vp]
    L.b2r_backproject_depth32.argtypes = [vp, vp, i32, i32, d, d, d, d, d, d, vp, vp, i32, ip]
    _lib = L
    return L


STATUS_NAMES = {0: "B2R_OK", 1: "B2R_ERR_INVALID_ARG", 2: "B2R_ERR_CUDA", 3: "B2R_ERR_OOM", 4: "B2R_ERR_STATE", 5: "B2R_ERR_LIMIT"}


class B2RError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"{STATUS_NAMES.get(status, status)}: {msg}")
        self.status = status


_CANARY = os.environ.get("B2R_CANARY") == "1"


def check(handle, status: int):
    if status == 0 and _CANARY and handle:
        # debug mode (include/b2r_debug.h): every call is followed by a check of the guard zones around the workspace buffers
        status = lib().b2r_debug_check_canaries(handle)
    if status != 0:
        msg = lib().b2r_last_error(handle).decode() if handle else "no handle"
        if status == 3:
            raise MemoryError(msg)
        raise B2RError(status, msg)
