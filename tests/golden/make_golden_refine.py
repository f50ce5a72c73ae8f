#!/usr/bin/env python
"""Regenerates tests/golden/refine_*.npz by EXECUTING THE REFERENCE'S OWN FUNCTION BODIES (this container only).

The scoring / Chamfer / resolution / PCA helpers of the reference are plain numpy + scipy:
    HelpersRealtime.py:100-130     nn_residuals, voxel_coverage, alignment_score
    registration_utils.py:21-23    get_angular_error
    registration_utils.py:98-128   uniform_downsample_farthest_point (true FPS)
    registration_utils.py:229-240  cloud_resolution
    registration_utils.py:276-343  chamfer_distance, centroid_of, pca_axes, initial_align_centroid_pca
(paths under /root/reference/src/pose_estimator/EstimHelpers/).  The modules themselves cannot be imported (they import open3d
and teaserpp_python at the top), so the function definitions are cut out of the files with `ast` and executed in a namespace
that holds numpy, scipy's cKDTree and a three-method stand-in for the two Open3D calls they make:
    pcd.points                               -> the ndarray
    pcd.compute_point_cloud_distance(other)  -> cKDTree(other).query(points)[0]         (nearest-neighbour distance, Open3D docs)
    KDTreeFlann(pcd).search_knn_vector_3d    -> brute-force k nearest, ascending squared distance
Everything else -- medians, percentiles, voxel counting, eigh, sign fixes, score formula -- is the reference's code running on
numpy/scipy, so these vectors PIN the oracle (oracle/refine_oracle.cpp) and the CUDA path for those functions.

    python tests/golden/make_golden_refine.py          (needs /root/reference; the .npz files are committed)
"""
import ast
import io
import os
import sys
import contextlib
import types

import numpy as np
from scipy.spatial import cKDTree

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/src/pose_estimator/EstimHelpers"
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))


class Cloud:
    """Stand-in for o3d.geometry.PointCloud: only what the extracted functions touch."""

    def __init__(self, pts):
        self.points = np.ascontiguousarray(pts, np.float64)

    def compute_point_cloud_distance(self, other):
        return cKDTree(other.points).query(self.points, k=1)[0]

    def select_by_index(self, idx):
        c = Cloud(self.points[np.asarray(idx)])
        c.idx = np.asarray(idx)
        return c


class KDTreeFlann:
    def __init__(self, pcd):
        self.p = pcd.points

    def search_knn_vector_3d(self, q, k):
        d2 = ((self.p - np.asarray(q)) ** 2).sum(1)
        idx = np.lexsort((np.arange(len(d2)), d2))[:k]
        return len(idx), idx, d2[idx]


def load_functions(path, names):
    tree = ast.parse(open(path).read())
    o3d = types.SimpleNamespace(geometry=types.SimpleNamespace(PointCloud=Cloud, KDTreeFlann=KDTreeFlann))
    ns = {"np": np, "cKDTree": cKDTree, "o3d": o3d}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            exec(compile(ast.Module([node], []), path, "exec"), ns)
    missing = [n for n in names if n not in ns]
    assert not missing, missing
    return ns


def clouds(seed, n_src, n_dst, noise):
    """A bumpy ellipsoid patch seen twice: dst = rigid motion of a resampling of the same surface + noise."""
    rng = np.random.default_rng(seed)

    def surf(n):
        u = rng.uniform(0, np.pi, n); v = rng.uniform(0, 1.3 * np.pi, n)
        r = 1 + 0.15 * np.sin(3 * u) * np.cos(2 * v)
        return np.stack([0.30 * r * np.sin(u) * np.cos(v), 0.18 * r * np.sin(u) * np.sin(v), 0.10 * r * np.cos(u)], 1)

    src = surf(n_src)
    a = rng.normal(size=3); a /= np.linalg.norm(a); th = 0.2
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    R = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * K @ K
    t = rng.uniform(-0.03, 0.03, 3)
    dst = surf(n_dst) @ R.T + t + rng.normal(0, noise, (n_dst, 3))
    return src, dst, R, t


def main():
    if not os.path.isdir(REF):
        raise SystemExit("the reference is not mounted here; the committed .npz files are the product of this script")
    hr = load_functions(os.path.join(REF, "HelpersRealtime.py"), ["nn_residuals", "voxel_coverage", "alignment_score"])
    ru = load_functions(os.path.join(REF, "registration_utils.py"),
                        ["get_angular_error", "cloud_resolution", "chamfer_distance", "centroid_of", "pca_axes", "initial_align_centroid_pca",
                         "uniform_downsample_farthest_point"])
    for name, seed, ns, nd, noise, voxel in [("a", 1, 400, 500, 0.002, 0.02), ("b", 2, 257, 129, 0.0, 0.05), ("c", 3, 1200, 900, 0.004, 0.01)]:
        src, dst, R, t = clouds(seed, ns, nd, noise)
        src_aligned = src @ R.T + t
        with contextlib.redirect_stdout(io.StringIO()):     # nn_residuals prints a debugging line
            d = hr["nn_residuals"](Cloud(src_aligned), Cloud(dst))
            score = hr["alignment_score"](Cloud(src_aligned), Cloud(src), Cloud(dst), voxel)
        np.random.seed(seed)                                   # the function draws its first index with np.random.randint
        with contextlib.redirect_stdout(io.StringIO()):
            fps = ru["uniform_downsample_farthest_point"](Cloud(src), target_points=min(100, ns // 2)).idx
        T0 = ru["initial_align_centroid_pca"](Cloud(src), Cloud(dst))
        Rs, ss = ru["pca_axes"](Cloud(src))
        out = dict(
            src=src, dst=dst, src_aligned=src_aligned, voxel=voxel,
            nn_dist=d, median=np.median(d), p90=np.percentile(d, 90),
            cov_aligned=hr["voxel_coverage"](src_aligned, voxel), cov_full=hr["voxel_coverage"](src, voxel), score=score,
            chamfer=ru["chamfer_distance"](Cloud(src_aligned), Cloud(dst)),
            resolution_src=ru["cloud_resolution"](Cloud(src)), resolution_dst=ru["cloud_resolution"](Cloud(dst), k=5),
            pca_T0=T0, pca_R_abs=np.abs(Rs), pca_s=ss, centroid_src=ru["centroid_of"](Cloud(src)),
            angular_error=ru["get_angular_error"](R, T0[:3, :3]), fps_idx=np.asarray(fps, np.int32),
        )
        np.savez_compressed(os.path.join(HERE, f"refine_{name}.npz"), **out)
        print(name, {k: (float(v) if np.ndim(v) == 0 else np.shape(v)) for k, v in out.items() if k not in ("src", "dst", "src_aligned")})


if __name__ == "__main__":
    main()
