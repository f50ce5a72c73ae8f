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


# ---- config 2: RealSense-shaped depth frame of a CAD-like box ---------------------------------------------------
@dataclass
class Frame:
    depth: np.ndarray        # (480, 640) uint16, depth_scale metres per unit
    mask: np.ndarray         # (480, 640) uint8, 255 inside the object's silhouette (Detector.py:51-52)
    bbox: tuple              # (u0, v0, u1, v1), half-open
    fx: float
    fy: float
    cx: float
    cy: float
    depth_scale: float
    R: np.ndarray            # model -> camera ground truth
    t: np.ndarray
    cad: np.ndarray          # (n_cad, 3) CAD surface samples in the model frame
    cad_normals: np.ndarray  # (n_cad, 3) outward face normals
    half: np.ndarray         # box half extents


def _box_surface_samples(half, n, rng):
    """Area-weighted uniform samples on the 6 faces of an axis-aligned box (model frame), with outward normals."""
    hx, hy, hz = half
    areas = np.array([hy * hz, hy * hz, hx * hz, hx * hz, hx * hy, hx * hy])
    face = rng.choice(6, size=n, p=areas / areas.sum())
    a = rng.uniform(-1, 1, n)
    b = rng.uniform(-1, 1, n)
    pts = np.zeros((n, 3))
    nrm = np.zeros((n, 3))
    for f in range(6):
        m = face == f
        ax = f // 2
        sgn = 1.0 if f % 2 == 0 else -1.0
        o = [i for i in range(3) if i != ax]
        pts[m, ax] = sgn * half[ax]
        pts[m, o[0]] = a[m] * half[o[0]]
        pts[m, o[1]] = b[m] * half[o[1]]
        nrm[m, ax] = sgn
    return pts, nrm


def make_frame(seed: int = 2, n_cad: int = 5000, dist: float = 0.6, size=(0.30, 0.15, 0.10), noise: bool = True,
               H: int = 480, W: int = 640) -> Frame:
    """BASELINE config 2 input: the reference reads 640x480 z16 depth (RealSenseClass.py:14-15) with a YOLO mask; here a
    box is ray-cast analytically at `dist` metres, with depth-dependent noise shaped like template_creation.py:24-30."""
    rng = np.random.default_rng(seed)
    fx = fy = 615.0
    cx, cy = W / 2.0, H / 2.0
    scale = 0.001
    half = np.asarray(size, float) / 2
    R = random_rotation(rng)
    t = np.array([rng.uniform(-0.05, 0.05), rng.uniform(-0.05, 0.05), dist])
    u, v = np.meshgrid(np.arange(W), np.arange(H))
    d = np.stack([(u - cx) / fx, (v - cy) / fy, np.ones_like(u, float)], -1)  # ray directions, z = 1
    o_m = -R.T @ t                       # camera centre in the model frame
    d_m = d @ R                          # R^T d for every pixel
    with np.errstate(divide="ignore", invalid="ignore"):
        t1 = (-half - o_m) / d_m
        t2 = (half - o_m) / d_m
    tn = np.nanmax(np.minimum(t1, t2), -1)
    tf = np.nanmin(np.maximum(t1, t2), -1)
    hit = (tn <= tf) & (tn > 0)
    z = np.where(hit, tn, 0.0)
    if noise:
        z = z + rng.normal(0.0, 1.0, z.shape) * 0.001 * z * z
    depth = np.where(hit, np.clip(np.rint(z / scale), 1, 65535), 0).astype(np.uint16)
    # background plane 1.5 m away so that the bbox (not only the mask) matters
    depth = np.where(hit, depth, np.uint16(1500)).astype(np.uint16)
    mask = np.where(hit, 255, 0).astype(np.uint8)
    vs, us = np.nonzero(hit)
    bbox = (int(us.min()), int(vs.min()), int(us.max()) + 1, int(vs.max()) + 1)
    cad, cad_n = _box_surface_samples(half, n_cad, rng)
    return Frame(depth=depth, mask=mask, bbox=bbox, fx=fx, fy=fy, cx=cx, cy=cy, depth_scale=scale, R=R, t=t, cad=cad,
                 cad_normals=cad_n, half=half)
