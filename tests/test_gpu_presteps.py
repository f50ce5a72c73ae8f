"""GPU parity tests of the pre-steps (SURVEY.md rows a-1a .. a-2): CUDA kernels through the C ABI vs oracle/presteps.

Bars: point order, kept indices, voxel membership, neighbour indices and correspondence indices bit-exact; coordinates,
voxel means, neighbour distances bit-exact (same IEEE operations in the same order); normals within 1e-10 for 99 % of the
points (1e-6 worst case) and FPFH within 1e-9 away from histogram bin edges (acos / cos / atan2 come from CUDA's libm
instead of glibc's -- DESIGN.md "pre-steps parity").
"""
import numpy as np
import pytest

from poseestimator_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def op(oracle_mod):
    from oracle import presteps
    presteps.lib()
    return presteps


@pytest.fixture(scope="module")
def ps():
    from poseestimator_b200.presteps import PreSteps
    return PreSteps(0)


@pytest.fixture(scope="module")
def frame():
    return synth.make_frame(seed=2)


def _np(t):
    return t.detach().cpu().numpy()


# ---- a-1a back-projection ----------------------------------------------------------------------------------------
@pytest.mark.parametrize("use_mask,use_bbox,zr", [(True, True, (0.0, 0.0)), (True, False, (0.0, 0.0)), (False, True, (0.0, 0.0)),
                                                    (True, True, (0.01, 10.0)), (False, False, (0.55, 0.62))])
def test_backproject_bit_exact(op, ps, frame, use_mask, use_bbox, zr):
    f = frame
    mask = f.mask if use_mask else None
    bbox = f.bbox if use_bbox else None
    ref_p, ref_pix = op.backproject(f.depth, mask, bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale, 3.0, zr[0], zr[1])
    pts, pix = ps.backproject(f.depth, mask, bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale, 3.0, zr[0], zr[1], want_pixels=True)
    assert ref_p.shape[0] > 1000
    assert np.array_equal(_np(pix), ref_pix)
    assert np.array_equal(_np(pts), ref_p)


def test_backproject_odd_shapes_and_device_input(op, ps):
    """Unaligned bbox edges, a width that is not a multiple of 8 (scalar path), empty results, CUDA-resident image."""
    import torch
    rng = np.random.default_rng(3)
    for (H, W) in [(37, 53), (64, 64), (5, 8)]:
        depth = rng.integers(0, 4000, size=(H, W)).astype(np.uint16)
        mask = (rng.uniform(size=(H, W)) < 0.6).astype(np.uint8) * 255
        for bbox in [(3, 2, W - 5, H - 1), (0, 0, W, H), (7, 1, 9, 3), (W - 1, H - 1, W, H), (-4, -4, W + 9, H + 9)]:
            ref_p, ref_pix = op.backproject(depth, mask, bbox, 300.0, 310.0, W / 2 - 0.5, H / 2, 0.001, 3.0)
            pts, pix = ps.backproject(depth, mask, bbox, 300.0, 310.0, W / 2 - 0.5, H / 2, 0.001, 3.0, want_pixels=True)
            assert np.array_equal(_np(pix), ref_pix) and np.array_equal(_np(pts), ref_p)
            d_t = torch.from_numpy(depth.view(np.int16)).cuda().view(torch.uint16)
            m_t = torch.from_numpy(mask).cuda()
            pts2 = ps.backproject(d_t, m_t, bbox, 300.0, 310.0, W / 2 - 0.5, H / 2, 0.001, 3.0)
            assert np.array_equal(_np(pts2), ref_p)
    empty = ps.backproject(np.zeros((16, 16), np.uint16), None, None, 300.0, 300.0, 8.0, 8.0, 0.001)
    assert empty.shape[0] == 0


@pytest.mark.parametrize("H,W,bbox", [(1536, 2048, None), (1100, 1999, (5, 3, 1990, 1097))])
def test_backproject_large_image_two_pass(op, ps, H, W, bbox):
    """More than 512 tiles of 2048 pixels: the two-pass form (tile counts, scan kernel, write with known bases) and the warp-staged
    output must give the oracle's points in the oracle's order; the second case takes the scalar load path (W % 8 != 0)."""
    rng = np.random.default_rng(H + W)
    depth = rng.integers(300, 3500, size=(H, W)).astype(np.uint16)
    depth[rng.uniform(size=(H, W)) < 0.15] = 0
    mask = (rng.uniform(size=(H, W)) < 0.7).astype(np.uint8) * 255
    mask[:, : W // 5] = 0                                # long runs of tiles without a single valid pixel
    ref_p, ref_pix = op.backproject(depth, mask, bbox, 900.0, 910.0, W / 2 - 0.5, H / 2, 0.001, 3.0)
    pts, pix = ps.backproject(depth, mask, bbox, 900.0, 910.0, W / 2 - 0.5, H / 2, 0.001, 3.0, want_pixels=True)
    assert ref_p.shape[0] > 500000
    assert np.array_equal(_np(pix), ref_pix)
    assert np.array_equal(_np(pts), ref_p)


# ---- kNN ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n,nq,K,radius", [(1, 1, 1, 0.0), (40, 40, 128, 0.0), (700, 300, 20, 0.0), (2000, 2000, 30, 0.08),
                                            (3000, 500, 100, 0.2), (513, 77, 7, 0.05)])
def test_knn_bit_exact(op, ps, n, nq, K, radius):
    rng = np.random.default_rng(n + K)
    pts = rng.uniform(-0.5, 0.5, size=(n, 3))
    pts[n // 2:] = np.round(pts[n // 2:], 2)        # duplicates and exact distance ties
    q = pts[:nq] if nq <= n else rng.uniform(-0.5, 0.5, size=(nq, 3))
    ri, rd, rc = op.knn(pts, q, K, radius)
    gi, gd, gc = ps.knn(pts, q, K, radius)
    assert np.array_equal(_np(gc), rc)
    assert np.array_equal(_np(gi), ri)
    assert np.array_equal(_np(gd), rd)


def test_knn_descending_order_stress(op, ps):
    """Points sorted by decreasing distance from the query: every candidate qualifies, every batch merges."""
    rng = np.random.default_rng(5)
    pts = rng.uniform(-1, 1, size=(1500, 3))
    q = np.zeros((1, 3))
    order = np.argsort(-np.linalg.norm(pts, axis=1))
    pts = pts[order]
    ri, rd, rc = op.knn(pts, q, 100, 0.0)
    gi, gd, gc = ps.knn(pts, q, 100, 0.0)
    assert np.array_equal(_np(gi), ri) and np.array_equal(_np(gd), rd)


@pytest.mark.parametrize("n,nq,K,radius,shape", [(6000, 6000, 20, 0.0, "surface"), (20000, 3000, 20, 0.0, "frame"), (9000, 9000, 30, 0.01, "surface"),
                                                  (8000, 2000, 100, 0.025, "surface"), (6000, 500, 128, 0.0, "line"), (5700, 300, 16, 0.0, "dups"),
                                                  (7000, 400, 12, 0.3, "outside")])
def test_knn_grid_equals_oracle(op, ps, frame, n, nq, K, radius, shape):
    """Clouds of 5632 points and more take the uniform-grid path; it must return exactly what brute force returns."""
    rng = np.random.default_rng(n + K)
    if shape == "frame":
        f = frame
        pts, _ = op.backproject(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
        pts = pts[:: max(1, pts.shape[0] // n)][:n]
    elif shape == "surface":
        uv = rng.uniform(-0.15, 0.15, size=(n, 2))
        pts = np.stack([uv[:, 0], uv[:, 1], 0.6 + 0.2 * uv[:, 0] * uv[:, 1] + rng.normal(0, 5e-4, n)], 1)
    elif shape == "line":
        pts = np.stack([np.linspace(0, 1, n), np.zeros(n), np.zeros(n)], 1)
    elif shape == "dups":
        pts = np.repeat(rng.uniform(-1, 1, size=(n // 50, 3)), 50, axis=0)
        pts = np.vstack([pts, rng.uniform(-1, 1, size=(n - pts.shape[0], 3))])
    else:
        pts = rng.uniform(-0.5, 0.5, size=(n, 3))
    n = pts.shape[0]
    q = pts[rng.permutation(n)[:nq]] if shape != "outside" else rng.uniform(-3, 3, size=(nq, 3))
    ri, rd, rc = op.knn(pts, q, K, radius)
    gi, gd, gc = ps.knn(pts, q, K, radius)
    assert np.array_equal(_np(gc), rc)
    assert np.array_equal(_np(gi), ri)
    assert np.array_equal(_np(gd), rd)


def test_statistical_outlier_full_bbox_cloud(op, ps, frame):
    """The whole back-projected bbox cloud (tens of thousands of points): grid kNN inside remove_statistical_outlier."""
    f = frame
    p, _ = op.backproject(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    keep_ref, avg_ref = op.statistical_outlier(p, 20, 1.0)
    out, idx, keep, avg = ps.statistical_outlier(p, 20, 1.0)
    assert np.array_equal(_np(avg), avg_ref)
    assert np.array_equal(_np(keep).astype(bool), keep_ref)
    assert np.array_equal(_np(out), p[keep_ref])


# ---- statistical outlier removal ---------------------------------------------------------------------------------
def test_statistical_outlier_bit_exact(op, ps, frame):
    f = frame
    p, _ = op.backproject(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    rng = np.random.default_rng(0)
    p = np.vstack([p[::3], rng.uniform(-0.3, 0.3, size=(40, 3)) + [0, 0, 0.6]])   # a few fliers
    keep_ref, avg_ref = op.statistical_outlier(p, 20, 1.0)
    out, idx, keep, avg = ps.statistical_outlier(p, 20, 1.0)
    assert np.array_equal(_np(avg), avg_ref)
    assert np.array_equal(_np(keep).astype(bool), keep_ref)
    assert np.array_equal(_np(idx), np.nonzero(keep_ref)[0])
    assert np.array_equal(_np(out), p[keep_ref])
    assert 0 < keep_ref.sum() < p.shape[0]


def test_statistical_outlier_small_and_duplicates(op, ps):
    rng = np.random.default_rng(9)
    for n in (1, 2, 5, 19, 20, 21):
        p = rng.normal(size=(n, 3))
        keep_ref, avg_ref = op.statistical_outlier(p, 20, 1.0)
        _, _, keep, avg = ps.statistical_outlier(p, 20, 1.0)
        assert np.array_equal(_np(avg), avg_ref) and np.array_equal(_np(keep).astype(bool), keep_ref)
    p = np.repeat(rng.normal(size=(6, 3)), 30, axis=0)   # 30 copies of 6 points: avg distance 0 -> dropped
    keep_ref, avg_ref = op.statistical_outlier(p, 20, 1.0)
    _, _, keep, avg = ps.statistical_outlier(p, 20, 1.0)
    assert np.array_equal(_np(keep).astype(bool), keep_ref)


# ---- a-1b voxel ----------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("voxel", [0.005, 0.02, 0.05])
def test_voxel_down_bit_exact(op, ps, frame, voxel):
    f = frame
    p, _ = op.backproject(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    ref_out, ref_vox, ref_cnt = op.voxel_down(p, voxel)
    out, vox, cnt = ps.voxel_down(p, voxel)
    assert out.shape[0] == ref_out.shape[0]
    assert np.array_equal(_np(vox), ref_vox)
    assert np.array_equal(_np(cnt), ref_cnt)
    assert np.array_equal(_np(out), ref_out)


def test_voxel_down_edge_cases(op, ps):
    rng = np.random.default_rng(4)
    one = rng.normal(size=(1, 3))
    out, vox, cnt = ps.voxel_down(one, 0.05)
    assert np.array_equal(_np(out), one) and _np(cnt).tolist() == [1]
    big = rng.uniform(-50, 50, size=(5000, 3))        # wide extent, tiny voxel: 40-bit keys
    ref_out, ref_vox, ref_cnt = op.voxel_down(big, 0.001)
    out, vox, cnt = ps.voxel_down(big, 0.001)
    assert np.array_equal(_np(out), ref_out) and np.array_equal(_np(vox), ref_vox)
    same = np.tile(rng.normal(size=(1, 3)), (300, 1))
    ref_out, _, _ = op.voxel_down(same, 0.01)
    out, _, cnt = ps.voxel_down(same, 0.01)
    assert np.array_equal(_np(out), ref_out) and _np(cnt).tolist() == [300]


# ---- a-1c normals, a-1d FPFH ---------------------------------------------------------------------------------------
def _cloud(op, frame, voxel=0.005):
    f = frame
    p, _ = op.backproject(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    v, _, _ = op.voxel_down(p, voxel)
    return v


def test_normals_match_oracle(op, ps, frame):
    v = _cloud(op, frame)
    for (r, k, orient) in [(0.005, 30, 0), (0.01, 30, 1), (0.02, 10, 2)]:
        ref = op.normals(v, r, k, orient, (0, 0, 0))
        got = _np(ps.normals(v, r, k, orient, (0, 0, 0)))
        # identical IEEE operations in the same order, acos / cos from the shared include/b2r_exactmath.h: identical bits
        assert np.array_equal(got, ref), np.abs(got - ref).max()
        assert np.allclose(np.linalg.norm(got, axis=1), 1.0, atol=1e-12)


def test_fpfh_matches_oracle(op, ps, frame):
    v = _cloud(op, frame)
    nrm = op.normals(v, 0.01, 30, 1, (0, 0, 0))
    ref, ref_sp = op.fpfh(v, nrm, 0.025, 100, want_spfh=True)
    got, got_sp = ps.fpfh(v, nrm, 0.025, 100, want_spfh=True)
    got, got_sp = _np(got), _np(got_sp)
    # every pair feature lands in the same bin (the exact acos / atan2 decide near a bin edge), so SPFH and FPFH are identical bits
    assert np.array_equal(got_sp, ref_sp), f"{(np.abs(got_sp - ref_sp).max(axis=0) > 0).sum()} SPFH columns differ"
    assert np.array_equal(got, ref), np.abs(got - ref).max()


def test_fpfh_on_exact_planes(op, ps, frame):
    """CAD samples lie exactly on axis-aligned planes with axis-aligned normals: signed zeros, exact ties and pair features that
    sit EXACTLY on bin edges (atan2 = 0, +-pi/2, pi) everywhere -- the guard band of the libm fast path is exercised on most pairs."""
    f = frame
    ref = op.fpfh(f.cad[:2000], f.cad_normals[:2000], 0.025, 100)
    got = _np(ps.fpfh(f.cad[:2000], f.cad_normals[:2000], 0.025, 100))
    assert np.array_equal(got, ref), f"{(np.abs(got - ref).max(axis=0) > 0).sum()} columns differ"


# ---- a-2 nearest neighbour in feature space (exact form) -----------------------------------------------------------
@pytest.mark.parametrize("na,nb,dim", [(1, 1, 33), (300, 1000, 33), (1000, 257, 33), (200, 150, 8)])
def test_nn_exact_and_mutual_bit_exact(op, ps, na, nb, dim):
    rng = np.random.default_rng(na + nb)
    fa = np.round(rng.uniform(0, 100, size=(na, dim)), 1)
    fb = np.round(rng.uniform(0, 100, size=(nb, dim)), 1)
    fb[: min(na, nb) // 2] = fa[: min(na, nb) // 2]        # exact matches
    if nb > 20:
        fb[10] = fb[3]                                      # duplicate rows: lowest index must win
    pairs_ref, nn01_ref, nn10_ref = op.mutual_nn(fa, fb)
    pairs, nn01, nn10 = ps.mutual_nn(fa, fb, method="exact")
    assert np.array_equal(_np(nn01), nn01_ref) and np.array_equal(_np(nn10), nn10_ref)
    assert np.array_equal(_np(pairs), pairs_ref)


def test_gather_correspondences(ps):
    import torch
    rng = np.random.default_rng(1)
    a, b = rng.normal(size=(50, 3)), rng.normal(size=(70, 3))
    pairs = np.stack([rng.permutation(50)[:20], rng.permutation(70)[:20]], 1).astype(np.int32)
    src, dst = ps.gather_correspondences(a, b, torch.from_numpy(pairs).cuda())
    assert np.array_equal(_np(src), a[pairs[:, 0]]) and np.array_equal(_np(dst), b[pairs[:, 1]])


# ---- a-2 on the tensor cores: tcgen05 3xTF32 GEMM shortlist + exact re-rank must return the oracle's indices ---------
@pytest.mark.parametrize("na,nb,kind", [(1, 1, "uniform"), (100, 37, "uniform"), (128, 128, "uniform"), (300, 1000, "uniform"),
                                        (1000, 257, "hist"), (2500, 5000, "hist"), (5000, 4808, "hist"), (640, 700, "dups")])
def test_nn_gemm_indices_equal_oracle(op, ps, na, nb, kind):
    rng = np.random.default_rng(na * 7 + nb)
    if kind == "uniform":
        fa = rng.uniform(0, 100, size=(na, 33)); fb = rng.uniform(0, 100, size=(nb, 33))
    else:   # FPFH-like: three 11-bin histograms summing to 200 each, many near-duplicates
        base = rng.dirichlet(np.full(11, 0.3), size=(64, 3)).reshape(64, 33) * 200.0
        fa = base[rng.integers(0, 64, na)] + rng.normal(0, 0.5 if kind == "hist" else 0.0, size=(na, 33))
        fb = base[rng.integers(0, 64, nb)] + rng.normal(0, 0.5 if kind == "hist" else 0.0, size=(nb, 33))
    pairs_ref, nn01_ref, nn10_ref = op.mutual_nn(fa, fb)
    pairs, nn01, nn10 = ps.mutual_nn(fa, fb, method="gemm")
    assert np.array_equal(_np(nn01), nn01_ref) and np.array_equal(_np(nn10), nn10_ref)
    assert np.array_equal(_np(pairs), pairs_ref)
    if kind == "dups":
        assert ps.last_fallback_rows > 0       # exact duplicates cannot be separated by the shortlist: brute-force rows
    elif kind == "uniform":
        assert ps.last_fallback_rows == 0
    nn, d2 = ps.nn_gemm(fa, fb)
    nn_e, d2_e = ps.nn_exact(fa, fb)
    assert np.array_equal(_np(nn), _np(nn_e)) and np.array_equal(_np(d2), _np(d2_e))


def test_nn_gemm_on_real_fpfh(op, ps, frame):
    """Config-2 shaped: FPFH of the voxelised scene cloud vs FPFH of the 5000-point CAD cloud."""
    v = _cloud(op, frame)
    nrm = op.normals(v, 0.01, 30, 1, (0, 0, 0))
    fs = op.fpfh(v, nrm, 0.025, 100).T.copy()
    fc = op.fpfh(frame.cad, frame.cad_normals, 0.025, 100).T.copy()
    pairs_ref, nn01_ref, nn10_ref = op.mutual_nn(fs, fc)
    pairs, nn01, nn10 = ps.mutual_nn(fs, fc, method="gemm")
    assert np.array_equal(_np(nn01), nn01_ref) and np.array_equal(_np(nn10), nn10_ref)
    assert np.array_equal(_np(pairs), pairs_ref)
    assert pairs.shape[0] > 10


@pytest.mark.parametrize("seed,size", [(2, (0.30, 0.15, 0.10)), (6, (0.30, 0.15, 0.10)), (11, (0.22, 0.22, 0.06)), (12, (0.15, 0.075, 0.05)),
                                       (13, (0.40, 0.10, 0.10))])
def test_nn_gemm_real_fpfh_shapes_and_fallback_rate(op, ps, seed, size):
    """More seeds / box shapes of synth.make_frame: indices == oracle in both directions, fewer than 5 % of the rows need the
    brute-force fallback, and the error of the tensor-core distances stays inside the bound the shortlist logic assumes."""
    f = synth.make_frame(seed=seed, size=size)
    raw, _ = op.backproject(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    keep, _ = op.statistical_outlier(raw, 20, 1.0)      # the config-2 flow; without it isolated noise points give all-zero
    v, _, _ = op.voxel_down(raw[keep], 0.005)           # descriptors, i.e. exact ties, which no shortlist can separate
    tm, _, _ = op.voxel_down(f.cad, 0.005)
    fs = op.fpfh(v, op.normals(v, 0.01, 30, 1, (0, 0, 0)), 0.025, 100).T.copy()
    ft = op.fpfh(tm, op.normals(tm, 0.01, 30, 2, (0, 0, 0)), 0.025, 100).T.copy()
    pairs_ref, nn01_ref, nn10_ref = op.mutual_nn(ft, fs)
    pairs, nn01, nn10 = ps.mutual_nn(ft, fs, method="gemm")
    assert np.array_equal(_np(nn01), nn01_ref) and np.array_equal(_np(nn10), nn10_ref) and np.array_equal(_np(pairs), pairs_ref)
    # config-2 box: < 5 % of the rows; boxes with a square cross-section have whole faces of near-identical descriptors (real near-ties)
    limit = 0.05 if size == (0.30, 0.15, 0.10) else 0.25
    assert ps.last_fallback_rows < limit * (ft.shape[0] + fs.shape[0]), (ps.last_fallback_rows, ft.shape[0] + fs.shape[0])
    for a, b, ref in ((ft, fs, nn01_ref), (fs, ft, nn10_ref)):
        nn, e = ps.nn_gemm_error(a, b)
        assert np.array_equal(_np(nn), ref)
        assert 0 < e["ratio"] < 0.5 * e["bound"], e      # measured error at most half the assumed bound


def test_nn_gemm_error_bound_large(ps):
    """The 3xTF32 / K = 128 error on a GPU-filling case (8192 x 8192 FPFH-like histograms + uniform rows): measured, not assumed."""
    rng = np.random.default_rng(5)
    base = rng.dirichlet(np.full(11, 0.3), size=(256, 3)).reshape(256, 33) * 200.0
    fa = base[rng.integers(0, 256, 8192)] + rng.normal(0, 0.5, size=(8192, 33))
    fb = np.concatenate([base[rng.integers(0, 256, 6144)] + rng.normal(0, 0.5, size=(6144, 33)), rng.uniform(0, 200, size=(2048, 33))])
    nn, e = ps.nn_gemm_error(fa, fb)
    nn_e, _ = ps.nn_exact(fa, fb)
    assert np.array_equal(_np(nn), _np(nn_e))
    assert 0 < e["ratio"] < 0.5 * e["bound"], e
    print("nn_gemm measured error ratio", e)
