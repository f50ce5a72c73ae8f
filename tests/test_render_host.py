"""CPU: host-side camera geometry of poseestimator_b200/render.py against vectors produced by the reference's own functions
(tests/golden/render_helpers.npz), and the depth-render oracle against analytic scenes."""
import os

import numpy as np
import pytest

G = os.path.join(os.path.dirname(__file__), "golden", "render_helpers.npz")


def _helpers():
    # render.py imports the CUDA binding lazily; the geometry helpers need neither a GPU nor the library
    from poseestimator_b200 import render
    return render


def test_camera_helpers_match_reference_functions():
    r = _helpers()
    g = np.load(G)
    assert np.array_equal(np.array([r.fx_from_fov(f, w) for f, w in zip(g["fov"], g["width"])]), g["fx"])
    cams = r.get_reduced_camera_positions(0.73)
    assert [c["type"] for c in cams] == list(g["cam_type"])
    assert np.array_equal(np.array([c["eye"] for c in cams]), g["cam_eye"])
    assert np.array_equal(np.array([c["up"] for c in cams], float), g["cam_up"])
    assert np.array_equal(np.array([r.o3d_lookat(c["eye"], c["target"], c["up"]) for c in cams]), g["cam_lookat"])
    T = np.array([r.o3d_lookat(e, t, u) for e, t, u in zip(g["la_eye"], g["la_target"], g["la_up"])])
    assert np.array_equal(T, g["la_T"])
    for H, e, t, u in zip(g["H"], g["H_eye"], g["H_target"], g["H_up"]):
        ee, tt, uu = r.camera_eye_lookat_up_from_H(H)
        assert np.array_equal(ee, e) and np.array_equal(tt, t) and np.array_equal(uu, u)


def test_lookat_roundtrip_gives_back_the_pose():
    """look_at(target, eye, up) of camera_eye_lookat_up_from_H(H) is the view whose OpenCV model->camera pose is H again."""
    r = _helpers()
    g = np.load(G)
    for H in g["H"]:
        eye, target, up = r.camera_eye_lookat_up_from_H(H)
        assert np.allclose(r.model_to_camera_from_lookat(eye, target, up), H, atol=1e-9)


def test_depth_oracle_on_analytic_scenes():
    import oracle.raster as orr
    # a plane z = 2 facing the camera: every covered pixel reads exactly 2
    v = np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0.0]]); t = np.array([[0, 1, 2], [0, 2, 3]])
    T = np.eye(4); T[2, 3] = 2.0
    d = orr.render_depth(v, t, T, 100, 100, 32, 24, 64, 48, 0.01, 10)
    assert d.shape == (48, 64) and np.all(d == 2.0)
    # a closed box: the render equals the analytic slab intersection, no cracks along the shared edges
    from poseestimator_b200 import synth
    f = synth.make_frame(seed=2, noise=False, H=120, W=160)
    half = f.half
    c = np.array([[sx, sy, sz] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)], float) * half
    quads = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    tris = np.array([[q[0], q[1], q[2]] for q in quads] + [[q[0], q[2], q[3]] for q in quads])
    T = np.eye(4); T[:3, :3] = f.R; T[:3, 3] = f.t
    fx = f.fx * 160 / 640
    d = orr.render_depth(c, tris, T, fx, fx, 80.0, 60.0, 160, 120, 0.01, 5.0)
    u, vv = np.meshgrid(np.arange(160), np.arange(120))
    dirs = np.stack([(u - 80.0) / fx, (vv - 60.0) / fx, np.ones_like(u, float)], -1)
    o_m = -f.R.T @ f.t; d_m = dirs @ f.R
    with np.errstate(divide="ignore", invalid="ignore"):
        t1 = (-half - o_m) / d_m; t2 = (half - o_m) / d_m
    tn = np.nanmax(np.minimum(t1, t2), -1); tf = np.nanmin(np.maximum(t1, t2), -1)
    hit = (tn <= tf) & (tn > 0)
    inner = hit & (tf - tn > 1e-6)
    assert np.all(d[inner] > 0) and np.allclose(d[inner], tn[inner], rtol=1e-6)
    assert np.all(d[~hit & ~np.isclose(tn, tf, atol=1e-6)] == 0)
    pts, pix = orr.backproject_depth32(d, fx, fx, 80.0, 60.0, 1.0, 5.0)
    assert pts.shape[0] == int((d > 0).sum()) and np.array_equal(pix, np.flatnonzero(d.ravel() > 0))
    on_box = np.abs((pts - f.t) @ f.R)             # back in the model frame: every point lies on the box surface
    assert np.all(np.max(on_box - half, axis=1) < 1e-6) and np.all(np.min(np.abs(on_box - half), axis=1) < 1e-6)
