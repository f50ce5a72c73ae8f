"""GPU parity: refinement / scoring kernels (csrc/refine.cu through the C ABI) against
  * the golden vectors made by the reference's own numpy / scipy functions (tests/golden/refine_*.npz), and
  * the CPU oracle (oracle/refine_oracle.cpp) on seeded inputs, including the edge cases.
Bars: indices / counts exact; distances, medians, percentiles, scores bit-exact (same fp64 expression, selection not summation);
sums (Chamfer, PCA, ICP) within 1e-12 relative; ICP pose within 1e-5 rad / 1e-6 of the scene scale of the oracle's.
"""
import glob
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "refine_*.npz")))


@pytest.fixture(scope="module")
def rf():
    from poseestimator_b200.refine import Refiner
    return Refiner(0)


@pytest.fixture(scope="module")
def orc():
    import oracle.refine as r
    r.lib()
    return r


def _rot_err(Ra, Rb):
    return abs(np.arccos(min(max((np.trace(Ra.T @ Rb) - 1) / 2, -1.0), 1.0)))


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p) for p in GOLD])
def test_scoring_matches_reference_golden(rf, path):
    g = np.load(path)
    d, idx = rf.nn_distance(g["src_aligned"], g["dst"], want_index=True)
    assert np.array_equal(d.cpu().numpy(), g["nn_dist"])
    sc = rf.alignment_score(g["src_aligned"], g["src"], g["dst"], float(g["voxel"]), details=True)
    assert sc["median"] == float(g["median"]) and sc["p90"] == float(g["p90"])
    assert sc["cov_aligned"] == int(g["cov_aligned"]) and sc["cov_full"] == int(g["cov_full"])
    assert sc["score"] == float(g["score"])
    assert rf.alignment_score(g["src_aligned"], g["src"], g["dst"], float(g["voxel"])) == float(g["score"])
    assert rf.voxel_coverage(g["src"], float(g["voxel"])) == int(g["cov_full"])
    assert abs(rf.chamfer_distance(g["src_aligned"], g["dst"]) - float(g["chamfer"])) <= 1e-12 * float(g["chamfer"])
    assert rf.cloud_resolution(g["src"]) == float(g["resolution_src"])
    assert rf.cloud_resolution(g["dst"], k=5) == float(g["resolution_dst"])


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p) for p in GOLD])
def test_pca_matches_reference_golden(rf, path):
    g = np.load(path)
    R, s = rf.pca_axes(g["src"])
    assert np.allclose(np.abs(R), g["pca_R_abs"], atol=1e-9) and np.allclose(s, g["pca_s"], rtol=1e-12)
    assert abs(np.linalg.det(R) - 1) < 1e-12
    T0 = rf.initial_align_centroid_pca(g["src"], g["dst"])
    assert np.allclose(T0, g["pca_T0"], atol=1e-9)


def _pair(seed, n, ang, noise, nd=None):
    rng = np.random.default_rng(seed)
    src = rng.uniform(-0.3, 0.3, (n, 3)) * np.array([1.0, 0.6, 0.3])
    a = rng.normal(size=3); a /= np.linalg.norm(a)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    R = np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K
    t = rng.uniform(-0.02, 0.02, 3)
    dst = src @ R.T + t + rng.normal(0, noise, (n, 3))
    dst = dst[rng.permutation(n)]
    if nd is not None:
        dst = dst[:nd]
    return src, dst, R, t


@pytest.mark.parametrize("seed,n,ang,noise,nd,max_it,dist", [
    (1, 300, 0.15, 0.0, None, 200, 0.1), (2, 500, 0.25, 0.002, None, 1000, 0.05), (3, 64, 0.05, 0.001, None, 30, 0.1),
    (4, 2000, 0.1, 0.003, 1500, 30, 0.05), (5, 129, 0.3, 0.0, 100, 5, 0.2)])
def test_icp_matches_oracle(rf, orc, seed, n, ang, noise, nd, max_it, dist):
    src, dst, R, t = _pair(seed, n, ang, noise, nd)
    init = np.eye(4)
    if seed == 2:
        init[:3, 3] = [0.004, -0.003, 0.002]
    ref = orc.icp(src, dst, dist, init, max_it)
    res = rf.registration_icp(src, dst, dist, init, max_iteration=max_it, want_correspondences=True)
    scale = np.linalg.norm(src.max(0) - src.min(0))
    assert res.iterations == ref.iterations
    assert _rot_err(res.transformation[:3, :3], ref.transformation[:3, :3]) < 1e-5
    assert np.linalg.norm(res.transformation[:3, 3] - ref.transformation[:3, 3]) < 1e-6 * scale
    assert np.allclose(res.transformation, ref.transformation, atol=1e-10)
    assert abs(res.fitness - ref.fitness) < 1e-15 and abs(res.inlier_rmse - ref.inlier_rmse) <= 1e-10 * max(ref.inlier_rmse, 1e-3)
    i = np.nonzero(ref.correspondence >= 0)[0]
    assert np.array_equal(res.correspondence_set, np.stack([i, ref.correspondence[i]], 1))


def test_icp_edge_cases(rf, orc):
    rng = np.random.default_rng(9)
    src = rng.uniform(-1, 1, (50, 3)); dst = src + 100.0
    r = rf.registration_icp(src, dst, 0.05, np.eye(4), 30, want_correspondences=True)
    assert r.fitness == 0.0 and r.inlier_rmse == 0.0 and np.array_equal(r.transformation, np.eye(4)) and r.iterations == 1
    assert r.correspondence_set.shape == (0, 2)
    r0 = rf.registration_icp(src, src.copy(), 0.05, None, 0)
    assert r0.iterations == 0 and r0.fitness == 1.0 and r0.inlier_rmse == 0.0 and not r0.converged
    one = rf.registration_icp(src[:1], src[:1] + 0.01, 0.5, None, 30)       # a single point: pure translation
    ref = orc.icp(src[:1], src[:1] + 0.01, 0.5, None, 30)
    assert np.allclose(one.transformation, ref.transformation, atol=1e-12) and one.iterations == ref.iterations


def test_large_cloud_uses_grid_search(rf, orc):
    """More than 16384 reference points: the uniform-grid kNN of the pre-steps answers the same indices."""
    rng = np.random.default_rng(3)
    ref_pts = rng.uniform(-1, 1, (20000, 3)); q = rng.uniform(-1, 1, (700, 3))
    d, idx = rf.nn_distance(q, ref_pts, want_index=True)
    od, oi = orc.nn(q, ref_pts)
    assert np.array_equal(idx.cpu().numpy(), oi) and np.array_equal(d.cpu().numpy(), od)


def test_transform_and_scores_against_oracle(rf, orc):
    rng = np.random.default_rng(11)
    src = rng.uniform(-0.5, 0.5, (777, 3)); dst = rng.uniform(-0.5, 0.5, (901, 3))
    T = np.eye(4); T[:3, :3] = np.linalg.qr(rng.normal(size=(3, 3)))[0]; T[:3, 3] = [0.1, -0.2, 0.05]
    moved = rf.transform(src, T)
    assert np.array_equal(moved.cpu().numpy(), orc.transform(T, src))
    for voxel in (0.05, 0.013):
        a = rf.alignment_score(moved, src, dst, voxel, details=True)
        b = orc.alignment_score(orc.transform(T, src), src, dst, voxel)
        assert a == b
    assert abs(rf.chamfer_distance(src, dst) - orc.chamfer(src, dst)) < 1e-13
    assert rf.cloud_resolution(src[:1]) == 0.005 and rf.cloud_resolution(src[:2]) == orc.cloud_resolution(src[:2])
    big = rng.uniform(-1, 1, (12000, 3))                       # sub-sampling stride 2
    assert rf.cloud_resolution(big) == orc.cloud_resolution(big)
    neg = rng.uniform(-3, -1, (300, 3))                        # negative coordinates: floor, not truncation
    assert rf.voxel_coverage(neg, 0.1) == orc.voxel_coverage(neg, 0.1)
    assert rf.voxel_coverage(np.zeros((5, 3)), 0.1) == 1


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p) for p in GOLD])
def test_farthest_point_sampling_matches_reference_golden(rf, orc, path):
    g = np.load(path)
    want = g["fps_idx"]
    pts, idx = rf.uniform_downsample_farthest_point(g["src"], len(want), first_index=int(want[0]), want_index=True)
    assert np.array_equal(idx.cpu().numpy(), want) and np.array_equal(pts.cpu().numpy(), g["src"][want])
    rng = np.random.default_rng(5)
    big = rng.uniform(-1, 1, (5000, 3)); big[100] = big[7]                 # a duplicate point: ties go to the lowest index
    _, idx = rf.uniform_downsample_farthest_point(big, 300, first_index=4321, want_index=True)
    assert np.array_equal(idx.cpu().numpy(), orc.farthest_point_sample(big, 300, 4321))
    same = rf.uniform_downsample_farthest_point(big[:50], 80)              # fewer points than the target: unchanged
    assert same.shape == (50, 3)


def test_preprocess_point_cloud_uniform(rf):
    from oracle import presteps as op
    rng = np.random.default_rng(6)
    u = rng.uniform(0, np.pi, 3000); v = rng.uniform(0, 2 * np.pi, 3000)
    cloud = np.stack([0.3 * np.sin(u) * np.cos(v), 0.2 * np.sin(u) * np.sin(v), 0.1 * np.cos(u)], 1)
    down, f = rf.preprocess_point_cloud_uniform(cloud, 400, first_index=11)
    d = down.cpu().numpy()
    assert d.shape == (400, 3) and f.shape == (33, 400)
    nrm = op.normals(d, 0.05, 30)
    ref = op.fpfh(d, nrm, 0.125, 100)
    assert np.allclose(f.cpu().numpy(), ref, rtol=1e-9, atol=1e-9)
    assert rf.preprocess_point_cloud_uniform(np.zeros((0, 3)), 100) == (None, None)
    assert rf.preprocess_point_cloud_uniform(cloud[:5], 100) == (None, None)
    r, none = rf.preprocess_point_cloud_uniform(cloud, 100, method="random", calc_fpfh=False, rng=np.random.default_rng(0))
    assert r.shape == (100, 3) and none is None
