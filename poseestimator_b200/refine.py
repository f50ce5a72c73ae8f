"""Refinement and scoring after TEASER++ on the GPU: the per-template tail of
`PoseEstimator.find_best_template_teaser` and the helpers around it, under the reference's own names.

    registration_icp            o3d.pipelines.registration.registration_icp            PoseEstimator.py:97-101, main_realsense.py:112-115
    nn_residuals / voxel_coverage / alignment_score                                    HelpersRealtime.py:100-130
    chamfer_distance                                                                   registration_utils.py:276-280
    cloud_resolution                                                                   registration_utils.py:229-240
    pca_axes / initial_align_centroid_pca                                              registration_utils.py:290-343
    get_angular_error                                                                  registration_utils.py:21-23
(paths under /root/reference/src/pose_estimator/EstimHelpers/).  Clouds are [n,3] float64 numpy arrays or CUDA tensors; results
that the reference returns as numpy come back as numpy.  Thin host layer over the C ABI (include/b2r.h); no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib
from .presteps import PreSteps, _torch


@dataclass
class RegistrationResult:
    """Fields of open3d.pipelines.registration.RegistrationResult that the reference reads (`.transformation`) plus the rest."""
    transformation: np.ndarray
    fitness: float
    inlier_rmse: float
    correspondence_set: "np.ndarray | None"   # [m,2] (source index, target index), filled when requested
    iterations: int = 0
    converged: bool = False


class Refiner(PreSteps):
    """One handle = one device + one stream, like PreSteps (whose kNN / voxel kernels it shares)."""

    def transform(self, points, T):
        torch = _torch()
        p = self._pts(points)
        out = self._empty(tuple(p.shape), torch.float64)
        T = np.ascontiguousarray(T, np.float64)
        assert T.shape == (4, 4)
        _lib.check(self._h, _lib.lib().b2r_transform_points(self._h, T.ctypes.data, p.data_ptr(), p.shape[0], out.data_ptr()))
        return out

    def nn_distance(self, queries, points, want_index=False):
        """-> Euclidean distance to the nearest point [nq] (CUDA tensor) (and its index)."""
        torch = _torch()
        q, p = self._pts(queries), self._pts(points)
        d2 = self._empty((max(q.shape[0], 1),), torch.float64)
        idx = self._empty((max(q.shape[0], 1),), torch.int32)
        _lib.check(self._h, _lib.lib().b2r_nn_distance(self._h, q.data_ptr(), q.shape[0], p.data_ptr(), p.shape[0], d2.data_ptr(), idx.data_ptr()))
        d = torch.sqrt(d2[: q.shape[0]])
        return (d, idx[: q.shape[0]]) if want_index else d

    def registration_icp(self, source, target, max_correspondence_distance, init=None, max_iteration=30, relative_fitness=1e-6,
                         relative_rmse=1e-6, want_correspondences=False) -> RegistrationResult:
        torch = _torch()
        s, t = self._pts(source), self._pts(target)
        T0 = None if init is None else np.ascontiguousarray(init, np.float64)
        res = _lib.B2RIcpResult()
        corr = self._empty((max(s.shape[0], 1),), torch.int32) if want_correspondences else None
        _lib.check(self._h, _lib.lib().b2r_icp(self._h, s.data_ptr(), s.shape[0], t.data_ptr(), t.shape[0], float(max_correspondence_distance),
                                               T0.ctypes.data if T0 is not None else None, int(max_iteration), float(relative_fitness),
                                               float(relative_rmse), C.byref(res), corr.data_ptr() if corr is not None else None))
        cs = None
        if corr is not None:
            torch.cuda.synchronize(self._dev)
            c = corr[: s.shape[0]].cpu().numpy()
            i = np.nonzero(c >= 0)[0]
            cs = np.stack([i, c[i]], 1).astype(np.int32)
        return RegistrationResult(np.array(res.transformation, dtype=np.float64).reshape(4, 4), res.fitness, res.inlier_rmse, cs,
                                  res.iterations, bool(res.converged))

    def voxel_coverage(self, points, voxel_size) -> int:
        p = self._pts(points)
        c = C.c_int(0)
        _lib.check(self._h, _lib.lib().b2r_voxel_coverage(self._h, p.data_ptr(), p.shape[0], float(voxel_size), C.byref(c)))
        return c.value

    VOXEL_NOALLOC_MAX_POINTS = 524288   # include/b2r.h: B2R_VOXEL_NOALLOC_MAX_POINTS

    def voxel_coverage_noalloc(self, points, voxel_size, count_out):
        """b2r_voxel_coverage_noalloc (no counterpart in the reference): the same count, hash table in static device memory, nothing
        allocated by the call.  points: [n,3] float64 CUDA tensor; count_out: int32 CUDA tensor of >= 1 element, written on the
        handle's stream (valid after a synchronisation).  Calls that overlap on one GPU race (one shared table)."""
        torch = _torch()
        assert points.is_cuda and points.dtype == torch.float64 and points.is_contiguous() and points.shape[-1] == 3
        assert count_out.is_cuda and count_out.dtype == torch.int32 and count_out.numel() >= 1
        _lib.check(self._h, _lib.lib().b2r_voxel_coverage_noalloc(self._h, points.data_ptr(), points.shape[0], float(voxel_size),
                                                                  count_out.data_ptr()))

    def alignment_score(self, src_aligned, src_down, dst_down, voxel_size, details=False):
        a, b, d = self._pts(src_aligned), self._pts(src_down), self._pts(dst_down)
        out = (C.c_double * 5)()
        _lib.check(self._h, _lib.lib().b2r_alignment_score(self._h, a.data_ptr(), a.shape[0], b.data_ptr(), b.shape[0], d.data_ptr(), d.shape[0],
                                                           float(voxel_size), out))
        if details:
            return dict(median=out[0], p90=out[1], cov_aligned=int(out[2]), cov_full=int(out[3]), score=out[4])
        return out[4]

    def chamfer_distance(self, src, dst) -> float:
        a, b = self._pts(src), self._pts(dst)
        out = C.c_double(0)
        _lib.check(self._h, _lib.lib().b2r_chamfer(self._h, a.data_ptr(), a.shape[0], b.data_ptr(), b.shape[0], C.byref(out)))
        return out.value

    def cloud_resolution(self, points, k=8) -> float:
        p = self._pts(points)
        out = C.c_double(0)
        _lib.check(self._h, _lib.lib().b2r_cloud_resolution(self._h, p.data_ptr(), p.shape[0], int(k), C.byref(out)))
        return out.value

    def pca_axes(self, points):
        """-> (R [3,3] columns = principal directions by decreasing variance, det +1; s [3])."""
        p = self._pts(points)
        R = np.zeros((3, 3)); s = np.zeros(3); c = np.zeros(3)
        _lib.check(self._h, _lib.lib().b2r_pca_axes(self._h, p.data_ptr(), p.shape[0], R.ctypes.data, s.ctypes.data, c.ctypes.data))
        return R, s

    def initial_align_centroid_pca(self, src, dst) -> np.ndarray:
        a, b = self._pts(src), self._pts(dst)
        T = np.zeros((4, 4))
        _lib.check(self._h, _lib.lib().b2r_pca_align(self._h, a.data_ptr(), a.shape[0], b.data_ptr(), b.shape[0], T.ctypes.data))
        return T


    # ---- the alternative down-sampling paths of row a-1b --------------------------------------------------------------------
    def uniform_downsample_farthest_point(self, points, target_points=2000, first_index=None, rng=None, want_index=False):
        """True farthest-point sampling (registration_utils.py:98-128).  The reference draws the first index with
        np.random.randint; pass `first_index` or a seeded `rng` for a reproducible draw.  Clouds with <= target_points come back
        unchanged.  -> points [target,3] CUDA tensor in selection order (and the indices)."""
        torch = _torch()
        p = self._pts(points)
        n, target = p.shape[0], int(target_points)
        if n <= target:
            return (p, torch.arange(n, dtype=torch.int32, device=self._dev)) if want_index else p
        if first_index is None:
            first_index = int((rng or np.random).integers(n)) if hasattr(rng or np.random, "integers") else int(np.random.randint(n))
        idx = self._empty((target,), torch.int32)
        _lib.check(self._h, _lib.lib().b2r_farthest_point_sample(self._h, p.data_ptr(), n, target, int(first_index), idx.data_ptr()))
        out = p[idx.long()]
        return (out, idx) if want_index else out

    def uniform_downsample_random(self, points, target_points=500, rng=None, want_index=False):
        """HelpersRealtime.uniform_downsample_farthest_point (despite its name: np.random.choice without replacement, :47-54)."""
        torch = _torch()
        p = self._pts(points)
        idx = (rng or np.random).choice(p.shape[0], size=int(target_points), replace=False)
        it = torch.from_numpy(np.asarray(idx, np.int64)).to(self._dev)
        return (p[it], it.to(torch.int32)) if want_index else p[it]

    def preprocess_point_cloud_uniform(self, points, target_points=500, method="farthest_point", calc_fpfh=True, first_index=None, rng=None,
                                       radius=0.05):
        """registration_utils.preprocess_point_cloud_uniform (:133-174; HelpersRealtime's variant with method="random", :59-98):
        down-sample, hybrid normals (radius, max_nn = min(30, n // 2)), FPFH (2.5 * radius, max_nn 100).
        -> (points_down, fpfh [33,n] or None); (None, None) for empty / too small clouds, as the reference."""
        p = self._pts(points)
        if p.shape[0] == 0:
            return None, None
        if method == "random":
            down = self.uniform_downsample_random(p, target_points, rng) if p.shape[0] >= target_points else p
        else:
            down = self.uniform_downsample_farthest_point(p, target_points, first_index, rng)
        if down.shape[0] < 10:
            return None, None
        if not calc_fpfh:
            return down, None
        nrm = self.normals(down, radius, min(30, down.shape[0] // 2))
        return down, self.fpfh(down, nrm, radius * 2.5, 100)


def get_angular_error(R_exp, R_est) -> float:
    """registration_utils.py:21-23 (host arithmetic on two 3x3 matrices)."""
    return abs(np.arccos(min(max(((np.matmul(np.asarray(R_exp).T, np.asarray(R_est))).trace() - 1) / 2, -1.0), 1.0)))
