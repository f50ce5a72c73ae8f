"""Synthetic registration problems (BASELINE.json configs 1, 3, 4, 5).

The reference has no data in-tree (``data/`` is git-ignored, /root/reference/.gitignore:31), so the
benchmark and the parity tests run on seeded synthetic correspondence sets shaped like the ones
``run_teaser`` builds (3xN float64, non-contiguous F-order views --
/root/reference/src/pose_estimator/EstimHelpers/registration_utils.py:248-249).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

# sigma * sqrt(chi2inv_3(0.9999)) -- SURVEY.md Appendix D
NOISE_BOUND_FACTOR = 4.594297390401966


@dataclass
class Problem:
    src: np.ndarray          # (3, N) float64, F-order view like the reference builds
    dst: np.ndarray          # (3, N)
    R: np.ndarray            # ground truth (3,3)
    t: np.ndarray            # ground truth (3,)
    inliers: np.ndarray      # sorted ground-truth inlier ids
    noise_bound: float
    scene_scale: float       # bbox diagonal of src (sqrt(3) for the unit cube)
    seed: int


def random_rotation(rng: np.random.Generator) -> np.ndarray:
    """Uniform random rotation from a unit quaternion (no scipy dependency in the product)."""
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    w, x, y, z = q
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
        [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
        [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)],
    ])


def make_problem(n: int, outlier_ratio: float, seed: int, sigma: float = 0.01, scale: float = 1.0) -> Problem:
    """Config-1/3 style problem: src ~ U[-0.5,0.5]^3, random SE(3), Gaussian noise, uniform outliers."""
    rng = np.random.default_rng(seed)
    src = rng.uniform(-0.5, 0.5, size=(n, 3))
    R = random_rotation(rng)
    t = rng.uniform(-1.0, 1.0, size=3)
    dst = scale * src @ R.T + t + rng.normal(0.0, sigma, size=(n, 3))
    n_out = int(np.floor(outlier_ratio * n))
    perm = rng.permutation(n)
    out_ids = perm[:n_out]
    dst[out_ids] = rng.uniform(-2.5, 2.5, size=(n_out, 3))
    inl = np.sort(perm[n_out:]).astype(np.int32)
    # (N,3) C-order -> .T gives the 3xN F-order view the reference passes to solve()
    return Problem(src=src.T, dst=dst.T, R=R, t=t, inliers=inl, noise_bound=sigma * NOISE_BOUND_FACTOR,
                   scene_scale=float(np.sqrt(3.0)), seed=seed)


def config1(seed: int = 1000) -> Problem:
    return make_problem(1000, 0.5, seed)


def config3(seed: int = 5000) -> Problem:
    return make_problem(5000, 0.95, seed)


def config4_detection(frame: int, det: int) -> Problem:
    """One detection of config 4: N ~ U{200..1000}, outlier ratio ~ U[0.3,0.9], seed = frame*64+det."""
    seed = frame * 64 + det
    rng = np.random.default_rng(10_000_000 + seed)
    n = int(rng.integers(200, 1001))
    ratio = float(rng.uniform(0.3, 0.9))
    return make_problem(n, ratio, seed)


def config5(outlier_ratio: float = 0.95, seed: int = 20000) -> Problem:
    return make_problem(20000, outlier_ratio, seed)


def rotation_error_rad(Ra: np.ndarray, Rb: np.ndarray) -> float:
    """Geodesic distance (same formula as registration_utils.get_angular_error, :21-23, in radians)."""
    c = (np.trace(Ra.T @ Rb) - 1.0) / 2.0
    return float(np.arccos(np.clip(c, -1.0, 1.0)))
