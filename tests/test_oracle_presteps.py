"""Self-validation of the pre-step oracle (oracle/presteps_oracle.cpp).  Open3D is not installable here (parity
unpinned), so every function is checked against an independent restatement: numpy written from the reference's own
lines, scipy.cKDTree (the reference's NN tool, HelpersRealtime.py:5,103-104), numpy.linalg.eigh."""
import numpy as np
import pytest
from scipy.spatial import cKDTree

from poseestimator_b200 import synth


@pytest.fixture(scope="module")
def op(oracle_mod):
    from oracle import presteps
    presteps.lib()
    return presteps


@pytest.fixture(scope="module")
def frame():
    return synth.make_frame(seed=2, dist=1.2)


def test_backproject_follows_reference_lines(op, frame):
    f = frame
    # RealSenseClass.py:57-58 + Open3D RGBD defaults (depth_trunc 3.0), written with numpy
    depth_m = f.depth.astype(np.float32) * f.depth_scale
    depth_m[f.mask == 0] = 0.0
    depth_m[depth_m >= np.float32(3.0)] = 0.0
    v, u = np.nonzero(depth_m > 0)
    z = depth_m[v, u].astype(np.float64)
    ref = np.stack([(u - f.cx) * z / f.fx, (v - f.cy) * z / f.fy, z], 1)
    pts, pix = op.backproject(f.depth, f.mask, None, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    assert np.array_equal(pts, ref)
    assert np.array_equal(pix, v * f.depth.shape[1] + u)
    # bbox restriction gives the same points when the mask lies inside the bbox
    pts_b, _ = op.backproject(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    assert np.array_equal(pts_b, ref)
    # bbox only (north-star wording): background plane at 1.5 m is inside the bbox and below the truncation
    pts_nb, _ = op.backproject(f.depth, None, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    u0, v0, u1, v1 = f.bbox
    assert pts_nb.shape[0] == (u1 - u0) * (v1 - v0)
    # offline variant: registration_utils.py:75 keeps 0.01 < z < 10
    pts_z, _ = op.backproject(f.depth, None, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale, depth_trunc=0.0, zmin=0.01, zmax=1.4)
    assert pts_z.shape[0] == ref.shape[0] and np.array_equal(pts_z, ref)
    # empty mask
    e, _ = op.backproject(f.depth, np.zeros_like(f.mask), f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    assert e.shape == (0, 3)


def test_knn_against_ckdtree(op):
    rng = np.random.default_rng(0)
    p = rng.uniform(0, 1, (3000, 3))
    idx, d2, cnt = op.knn(p, p, 20)
    dd, ii = cKDTree(p).query(p, k=20)
    assert np.array_equal(idx, ii) and (cnt == 20).all()
    assert np.allclose(np.sqrt(d2), dd, rtol=0, atol=1e-15)
    r = 0.08
    idx, d2, cnt = op.knn(p, p, 30, radius=r)
    tree = cKDTree(p)
    for i in range(0, 3000, 37):
        ball = np.array(sorted(tree.query_ball_point(p[i], r * (1 - 1e-12)), key=lambda j: (np.sum((p[j] - p[i]) ** 2), j)))[:30]
        assert cnt[i] == len(ball) and np.array_equal(idx[i, :cnt[i]], ball)
        assert (idx[i, cnt[i]:] == -1).all()
    # ties: a regular lattice -- order is (d2, index)
    g = np.stack(np.meshgrid(*[np.arange(5.0)] * 3, indexing="ij"), -1).reshape(-1, 3)
    idx, d2, cnt = op.knn(g, g, 7)
    for i in range(len(g)):
        key = sorted(range(len(g)), key=lambda j: (np.sum((g[j] - g[i]) ** 2), j))[:7]
        assert idx[i].tolist() == key


def test_statistical_outlier_against_numpy(op, frame):
    f = frame
    pts, _ = op.backproject(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    rng = np.random.default_rng(1)
    pts = np.concatenate([pts[::3], rng.uniform(-0.5, 0.5, (40, 3)) + [0, 0, 1.2]])
    keep, avg = op.statistical_outlier(pts, 20, 1.0)
    dd, _ = cKDTree(pts).query(pts, k=20)
    a = dd.mean(1)
    assert np.allclose(avg, a, rtol=1e-12)
    thr = a.mean() + 1.0 * a.std(ddof=1)
    ref = (a > 0) & (a < thr)
    assert (keep != ref).sum() == 0
    assert keep[-40:].sum() <= 2 and keep[:-40].mean() > 0.8
    k1, _ = op.statistical_outlier(pts[:5], 20, 1.0)  # fewer points than neighbours
    assert k1.shape == (5,)


def test_voxel_down_against_numpy(op, frame):
    f = frame
    pts, _ = op.backproject(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    for vs in (0.005, 0.05):
        out, vox, cnt = op.voxel_down(pts, vs)
        org = pts.min(0) - vs * 0.5
        c = np.floor((pts - org) / vs).astype(np.int64)
        dim = np.floor((pts.max(0) - org) / vs).astype(np.int64) + 1
        key = c[:, 0] + dim[0] * (c[:, 1] + dim[1] * c[:, 2])
        uk, inv, counts = np.unique(key, return_inverse=True, return_counts=True)
        assert out.shape[0] == len(uk) and np.array_equal(vox, inv) and np.array_equal(cnt, counts)
        ref = np.zeros((len(uk), 3))
        np.add.at(ref, inv, pts)  # sequential in point order, like Open3D's AccumulatedPoint
        ref /= counts[:, None]
        assert np.array_equal(out, ref)
    one, vox, cnt = op.voxel_down(pts[:1], 0.01)
    assert np.array_equal(one, pts[:1]) and cnt.tolist() == [1]


def test_smallest_eigvec_against_eigh(op):
    rng = np.random.default_rng(3)
    for _ in range(200):
        A = rng.normal(size=(rng.integers(4, 30), 3)) * rng.uniform(0.01, 1, 3)
        Cm = np.cov(A.T, bias=True)
        w, V = np.linalg.eigh(Cm)
        e = op.smallest_eigvec(Cm)
        assert abs(np.linalg.norm(e) - 1) < 1e-9
        if w[1] - w[0] > 1e-6 * w[2]:
            assert abs(abs(e @ V[:, 0]) - 1) < 1e-7
    assert op.smallest_eigvec(np.diag([3.0, 1.0, 2.0])).tolist() == [0, 1, 0]
    assert op.smallest_eigvec(np.zeros((3, 3))).tolist() == [0, 0, 0]
    # exact plane z = 0 -> +-z
    P = np.c_[rng.uniform(-1, 1, (20, 2)), np.zeros(20)]
    assert abs(abs(op.smallest_eigvec(np.cov(P.T, bias=True))[2]) - 1) < 1e-12


def test_normals_on_box_faces(op, frame):
    f = frame
    cad, _, _ = op.voxel_down(f.cad, 0.005)
    n = op.normals(cad, 0.0125, 30, orient=2, view=(0, 0, 0))  # away from the box centre = outward
    # nearest true face normal
    face = np.argmax(np.abs(cad) / f.half, 1)
    interior = np.sort(np.abs(cad) / f.half, 1)[:, 1] < 0.8  # away from edges
    truth = np.zeros_like(cad)
    truth[np.arange(len(cad)), face] = np.sign(cad[np.arange(len(cad)), face])
    dots = np.sum(n * truth, 1)[interior]
    assert np.mean(dots > 0.999) > 0.97
    # fewer than 3 neighbours -> (0,0,1)
    far = np.array([[0, 0, 0], [10.0, 0, 0], [0, 10.0, 0], [0.001, 0, 0]])
    assert op.normals(far, 0.01, 30).tolist() == [[0, 0, 1]] * 4
    # against numpy eigh on a curved surface
    rng = np.random.default_rng(5)
    s = rng.normal(size=(1500, 3)); s /= np.linalg.norm(s, axis=1)[:, None]
    ns = op.normals(s, 0.2, 30, orient=2, view=(0, 0, 0))
    assert np.mean(np.sum(ns * s, 1) > 0.99) > 0.99
    idx, d2, cnt = op.knn(s, s, 30, 0.2)
    for i in range(0, 1500, 50):
        nb = s[idx[i, :cnt[i]]]
        w, V = np.linalg.eigh(np.cov(nb.T, bias=True))
        assert abs(abs(ns[i] @ V[:, 0]) - 1) < 1e-6


def _fpfh_numpy(p, n, r, max_nn):
    """From-the-definition restatement of SURVEY.md B.5 (pure numpy/python; small clouds only)."""
    N = len(p)
    tree = cKDTree(p)
    nbrs = []
    for i in range(N):
        cand = tree.query_ball_point(p[i], r * (1 + 1e-9))
        cand = [j for j in cand if np.sum((p[j] - p[i]) ** 2) < r * r]
        cand.sort(key=lambda j: (((p[j, 0] - p[i, 0]) ** 2 + (p[j, 1] - p[i, 1]) ** 2) + (p[j, 2] - p[i, 2]) ** 2, j))
        nbrs.append(cand[:max_nn])
    spfh = np.zeros((33, N))
    for i in range(N):
        nb = nbrs[i]
        if len(nb) <= 1:
            continue
        inc = 100.0 / (len(nb) - 1)
        for j in nb[1:]:
            dp = p[j] - p[i]
            d = np.sqrt((dp[0] * dp[0] + dp[1] * dp[1]) + dp[2] * dp[2])
            if d == 0:
                f = (0, 0, 0)
            else:
                n1, n2 = n[i], n[j]
                a1, a2 = n1 @ dp / d, n2 @ dp / d
                if np.arccos(abs(a1)) > np.arccos(abs(a2)):
                    n1, n2, dp, f2 = n[j], n[i], -dp, -a2
                else:
                    f2 = a1
                v = np.cross(dp, n1)
                vn = np.linalg.norm(v)
                if vn == 0:
                    f = (0, 0, 0)
                else:
                    v = v / vn
                    w = np.cross(n1, v)
                    f = (np.arctan2(w @ n2, n1 @ n2), v @ n2, f2)
            b0 = min(10, max(0, int(np.floor(11 * (f[0] + np.pi) / (2 * np.pi)))))
            b1 = min(10, max(0, int(np.floor(11 * (f[1] + 1) * 0.5))))
            b2 = min(10, max(0, int(np.floor(11 * (f[2] + 1) * 0.5))))
            spfh[b0, i] += inc; spfh[11 + b1, i] += inc; spfh[22 + b2, i] += inc
    out = np.zeros((33, N))
    for i in range(N):
        nb = nbrs[i]
        if len(nb) <= 1:
            continue
        s = np.zeros(3)
        for j in nb[1:]:
            d2 = np.sum((p[j] - p[i]) ** 2)
            if d2 == 0:
                continue
            val = spfh[:, j] / d2
            out[:, i] += val
            s += val.reshape(3, 11).sum(1)
        s = np.where(s != 0, 100.0 / np.where(s != 0, s, 1), 0)
        out[:, i] = out[:, i] * np.repeat(s, 11) + spfh[:, i]
    return out, spfh


def test_fpfh_against_numpy_definition(op):
    rng = np.random.default_rng(7)
    s = rng.normal(size=(400, 3)); s /= np.linalg.norm(s, axis=1)[:, None]
    s *= [1.0, 0.7, 0.5]
    nr = op.normals(s, 0.25, 30, orient=2, view=(0, 0, 0))
    F, S = op.fpfh(s, nr, 0.4, 40, want_spfh=True)
    Fr, Sr = _fpfh_numpy(s, nr, 0.4, 40)
    # identical up to pairs sitting on a bin edge (different libm / summation order in the numpy restatement)
    assert np.mean(np.abs(S - Sr) > 1e-9) < 0.002
    assert np.mean(np.abs(F - Fr) > 1e-6) < 0.01
    has = S.sum(0) > 0
    assert np.allclose(S[:, has].reshape(3, 11, -1).sum(1), 100.0)
    assert np.allclose(F[:, has].reshape(3, 11, -1).sum(1), 200.0)


def test_fpfh_is_rigid_motion_invariant(op):
    rng = np.random.default_rng(9)
    s = rng.normal(size=(500, 3)); s /= np.linalg.norm(s, axis=1)[:, None]
    s *= [1.0, 0.6, 0.8]
    R = synth.random_rotation(rng)
    nr = op.normals(s, 0.25, 30, orient=2, view=(0, 0, 0))
    F1 = op.fpfh(s, nr, 0.4, 100)
    F2 = op.fpfh(s @ R.T + [1, 2, 3], nr @ R.T, 0.4, 100)
    assert np.mean(np.abs(F1 - F2) > 1e-6) < 0.02


def test_mutual_nn_against_ckdtree(op):
    rng = np.random.default_rng(11)
    fa = rng.uniform(0, 100, (700, 33))
    fb = np.concatenate([fa[rng.permutation(700)[:400]] + rng.normal(0, 0.5, (400, 33)), rng.uniform(0, 100, (300, 33))])
    pairs, nn01, nn10 = op.mutual_nn(fa, fb)
    _, j01 = cKDTree(fb).query(fa)
    _, j10 = cKDTree(fa).query(fb)
    assert np.array_equal(nn01, j01) and np.array_equal(nn10, j10)
    ref = [(i, j01[i]) for i in range(700) if j10[j01[i]] == i]
    assert pairs.tolist() == [list(x) for x in ref] and len(ref) >= 350
    # ties: duplicated descriptors -> lowest index wins
    fb2 = np.concatenate([fb[:5], fb])
    _, nn01b, _ = op.mutual_nn(fa, fb2)
    hit = nn01 < 5
    assert (nn01b[hit] == nn01[hit]).all() and (nn01b[~hit] == nn01[~hit] + 5).all()
