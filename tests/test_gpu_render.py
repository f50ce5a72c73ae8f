"""GPU parity: CUDA depth rasteriser + float-depth back-projection (csrc/raster.cu through the C ABI) vs the oracle's brute force.
Bar: the depth image and the point cloud are bit-identical."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _icosphere(sub=3, r=0.1):
    t = (1 + 5 ** 0.5) / 2
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t), (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6), (7, 1, 8), (3, 9, 4), (3, 4, 2),
         (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    v = [np.array(p, float) / np.linalg.norm(p) for p in v]
    for _ in range(sub):
        cache, nf = {}, []

        def mid(a, b):
            k = (min(a, b), max(a, b))
            if k not in cache:
                m = v[a] + v[b]; v.append(m / np.linalg.norm(m)); cache[k] = len(v) - 1
            return cache[k]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    return np.array(v) * r * np.array([1.0, 0.6, 0.4]), np.array(f, np.int32)


def _pose(seed, dist):
    rng = np.random.default_rng(seed)
    Q = np.linalg.qr(rng.normal(size=(3, 3)))[0]; Q *= np.sign(np.linalg.det(Q))
    T = np.eye(4); T[:3, :3] = Q; T[:3, 3] = [rng.uniform(-0.03, 0.03), rng.uniform(-0.03, 0.03), dist]
    return T


@pytest.mark.parametrize("seed,sub,W,H,dist", [(1, 2, 160, 120, 0.4), (2, 3, 320, 240, 0.5), (3, 1, 64, 48, 0.12)])
def test_render_matches_oracle_bit_for_bit(seed, sub, W, H, dist):
    import oracle.raster as orr
    from poseestimator_b200.render import DepthRenderer
    v, f = _icosphere(sub)
    T = _pose(seed, dist)                     # seed 3: the object fills the image and crosses its border
    r = DepthRenderer(v, f, W, H, fov_deg=60.0)
    d = r.render_depth(T, 0.01, 5.0).cpu().numpy()
    ref = orr.render_depth(v, f, T, r.fx, r.fy, r.cx, r.cy, W, H, 0.01, 5.0)
    assert (ref > 0).sum() > 100 and np.array_equal(d.view(np.uint32), ref.view(np.uint32))
    pts, pix = r.backproject_depth(d, 1.0, 5.0, want_pixels=True)
    op, opix = orr.backproject_depth32(ref, r.fx, r.fy, r.cx, r.cy, 1.0, 5.0)
    assert np.array_equal(pts.cpu().numpy(), op) and np.array_equal(pix.cpu().numpy(), opix)


def test_render_near_far_and_empty():
    import oracle.raster as orr
    from poseestimator_b200.render import DepthRenderer
    v, f = _icosphere(1)
    r = DepthRenderer(v, f, 96, 72)
    T = _pose(5, 0.3)
    for near, far in [(0.29, 0.31), (0.01, 0.25), (0.4, 5.0)]:       # slab through the object, in front of it, behind it
        d = r.render_depth(T, near, far).cpu().numpy()
        ref = orr.render_depth(v, f, T, r.fx, r.fy, r.cx, r.cy, 96, 72, near, far)
        assert np.array_equal(d.view(np.uint32), ref.view(np.uint32))
    behind = T.copy(); behind[2, 3] = -0.3                             # object behind the camera
    assert not r.render_depth(behind, 0.01, 5.0).any()
    inside = T.copy(); inside[:3, 3] = 0                               # camera inside the mesh: vertices on both sides of z = 0
    d = r.render_depth(inside, 0.001, 5.0).cpu().numpy()
    ref = orr.render_depth(v, f, inside, r.fx, r.fy, r.cx, r.cy, 96, 72, 0.001, 5.0)
    assert (ref > 0).all() and np.array_equal(d.view(np.uint32), ref.view(np.uint32))
    pts = r.create_template_from_H(T, 200, rng=np.random.default_rng(0))
    assert pts.shape == (200, 3)


def test_render_templates_views_cover_the_model():
    """template_creation.render_templates: five views, points back in the model frame lie on the (millimetre-scaled) mesh."""
    from poseestimator_b200.render import render_templates
    v, f = _icosphere(3)
    out = render_templates(v * 1000.0 + np.array([5.0, -3.0, 2.0]), f, n_points=3000, rng=np.random.default_rng(1))
    assert [o["type"] for o in out] == ["6", "7", "8", "11", "12"]
    c = (v + np.array([5.0, -3.0, 2.0]) / 1000.0)
    for o in out:
        p = o["points"]
        assert p.shape == (3000, 3)
        q = (p - c.mean(0)) / (0.1 * np.array([1.0, 0.6, 0.4]))       # unit sphere coordinates of the ellipsoid
        assert np.all(np.abs(np.linalg.norm(q, axis=1) - 1) < 0.02)   # on the surface up to the faceting of the icosphere
