"""CPU: the refinement / scoring oracle (oracle/refine_oracle.cpp) against
  * the golden vectors produced by the reference's own numpy / scipy function bodies (tests/golden/refine_*.npz) -- pinned,
  * scipy / numpy restatements and ground-truth recovery for registration_icp (Open3D's; unpinned).
"""
import glob
import os

import numpy as np
import pytest
from scipy.spatial import cKDTree

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "refine_*.npz")))


@pytest.fixture(scope="module")
def orc():
    import oracle.refine as r
    r.lib()
    return r


def test_golden_files_present():
    assert len(GOLD) == 3


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p) for p in GOLD])
def test_scoring_matches_reference_functions(orc, path):
    g = np.load(path)
    d, idx = orc.nn(g["src_aligned"], g["dst"])
    assert np.array_equal(d, g["nn_dist"])                       # cKDTree distances: same sqrt of the same fp64 sum
    sc = orc.alignment_score(g["src_aligned"], g["src"], g["dst"], float(g["voxel"]))
    assert sc["median"] == float(g["median"]) and sc["p90"] == float(g["p90"])
    assert sc["cov_aligned"] == int(g["cov_aligned"]) and sc["cov_full"] == int(g["cov_full"])
    assert sc["score"] == float(g["score"])
    assert orc.voxel_coverage(g["src"], float(g["voxel"])) == int(g["cov_full"])
    assert abs(orc.chamfer(g["src_aligned"], g["dst"]) - float(g["chamfer"])) <= 1e-15 * 8   # np.mean sums pairwise
    assert orc.cloud_resolution(g["src"]) == float(g["resolution_src"])
    assert orc.cloud_resolution(g["dst"], k=5) == float(g["resolution_dst"])


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p) for p in GOLD])
def test_pca_prealignment_matches_reference_functions(orc, path):
    g = np.load(path)
    R, s, c = orc.pca_axes(g["src"])
    assert np.allclose(np.abs(R), g["pca_R_abs"], atol=1e-9)      # eigenvector signs are the solver's; |R| is not
    assert np.allclose(s, g["pca_s"], rtol=1e-12) and np.allclose(c, g["centroid_src"], atol=1e-15)
    assert abs(np.linalg.det(R) - 1) < 1e-12
    T0 = orc.pca_align(g["src"], g["dst"])
    assert np.allclose(T0, g["pca_T0"], atol=1e-9)                # sign fixes make T0 independent of the eigen-solver


def _rigid(rng, ang):
    a = rng.normal(size=3); a /= np.linalg.norm(a)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K


def _icp_numpy(src, dst, max_dist, init, max_it, rf=1e-6, rr=1e-6):
    """Independent restatement with cKDTree + numpy SVD."""
    tree = cKDTree(dst)
    T = init.copy()
    cur = src @ T[:3, :3].T + T[:3, 3]

    def ev(cur):
        d, j = tree.query(cur)
        m = d < max_dist
        return j, m, (m.sum() / len(cur), np.sqrt((d[m] ** 2).sum() / max(m.sum(), 1)))
    j, m, (fit, rm) = ev(cur)
    for _ in range(max_it):
        a, b = cur[m], dst[j[m]]
        ma, mb = a.mean(0), b.mean(0)
        H = (b - mb).T @ (a - ma) / len(a)
        U, S, Vt = np.linalg.svd(H)
        D = np.diag([1, 1, np.sign(np.linalg.det(U) * np.linalg.det(Vt))])
        R = U @ D @ Vt
        Ti = np.eye(4); Ti[:3, :3] = R; Ti[:3, 3] = mb - R @ ma
        T = Ti @ T
        cur = cur @ R.T + Ti[:3, 3]
        pf, pr = fit, rm
        j, m, (fit, rm) = ev(cur)
        if abs(pf - fit) < rf and abs(pr - rm) < rr:
            break
    return T, fit, rm


@pytest.mark.parametrize("seed,n,ang,noise", [(1, 300, 0.15, 0.0), (2, 500, 0.25, 0.002), (3, 64, 0.05, 0.001)])
def test_icp_recovers_motion_and_matches_numpy_restatement(orc, seed, n, ang, noise):
    rng = np.random.default_rng(seed)
    src = rng.uniform(-0.3, 0.3, (n, 3)) * np.array([1.0, 0.6, 0.3])
    R = _rigid(rng, ang); t = rng.uniform(-0.02, 0.02, 3)
    dst = src @ R.T + t + rng.normal(0, noise, (n, 3))
    dst = dst[rng.permutation(n)]
    init = np.eye(4)
    res = orc.icp(src, dst, 0.1, init, 200)
    Tn, fit, rm = _icp_numpy(src, dst, 0.1, init, 200)
    assert np.allclose(res.transformation, Tn, atol=1e-9)
    assert abs(res.fitness - fit) < 1e-12 and abs(res.inlier_rmse - rm) < 1e-9
    if noise == 0.0:
        assert np.allclose(res.transformation[:3, :3], R, atol=1e-6) and np.allclose(res.transformation[:3, 3], t, atol=1e-6)
        assert res.fitness == 1.0 and res.inlier_rmse < 1e-6


def test_icp_edge_cases(orc):
    rng = np.random.default_rng(9)
    src = rng.uniform(-1, 1, (50, 3)); dst = src + 100.0
    r = orc.icp(src, dst, 0.05, np.eye(4), 30)                  # nothing within reach: identity updates, stops after 1 iteration
    assert r.n_corr == 0 and r.fitness == 0.0 and r.inlier_rmse == 0.0 and np.array_equal(r.transformation, np.eye(4))
    assert r.iterations == 1 and (r.correspondence == -1).all()
    r0 = orc.icp(src, src.copy(), 0.05, np.eye(4), 0)           # max_iteration = 0: evaluation only
    assert r0.iterations == 0 and r0.fitness == 1.0 and r0.inlier_rmse == 0.0
    T = np.eye(4); T[:3, 3] = [0.01, -0.02, 0.005]
    r1 = orc.icp(src, src @ np.eye(3) + T[:3, 3], 0.5, T, 30)   # exact initial guess
    assert np.allclose(r1.transformation, T, atol=1e-12) and r1.inlier_rmse < 1e-12


def test_umeyama_reflection_case(orc):
    """Coplanar correspondences whose least-squares orthogonal fit is a reflection: det fix must give a rotation."""
    src = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0.0]])
    dst = src.copy(); dst[:, 2] = [0.01, -0.01, -0.01, 0.01]
    T = orc.umeyama(src, dst, np.arange(4, dtype=np.int32))
    assert abs(np.linalg.det(T[:3, :3]) - 1) < 1e-12 and np.allclose(T[:3, :3] @ T[:3, :3].T, np.eye(3), atol=1e-12)


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p) for p in GOLD])
def test_farthest_point_sampling_matches_reference_function(orc, path):
    g = np.load(path)
    want = g["fps_idx"]
    assert np.array_equal(orc.farthest_point_sample(g["src"], len(want), int(want[0])), want)
    assert len(set(want.tolist())) == len(want)
