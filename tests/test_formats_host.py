"""CPU: file formats of the reference's callers (poseestimator_b200/formats.py): PLY round trips and hand-written files,
BOP json, calib.yaml against the reference's own load_calib (tests/golden/formats_calib.npz)."""
import json
import os

import numpy as np
import pytest

from poseestimator_b200 import formats

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_ply_round_trips(tmp_path):
    rng = np.random.default_rng(0)
    p = rng.normal(size=(257, 3)); nrm = rng.normal(size=(257, 3)); col = rng.integers(0, 256, (257, 3)) / 255.0
    tri = rng.integers(0, 257, (40, 3)).astype(np.int32)
    for asc in (False, True):
        f = tmp_path / f"c{int(asc)}.ply"
        formats.write_ply(f, p, nrm, col, tri, ascii=asc)
        r = formats.read_ply(f)
        assert np.array_equal(r["points"], p) and np.array_equal(r["normals"], nrm)      # doubles survive both encodings exactly
        assert np.allclose(r["colors"], col, atol=1e-12) and np.array_equal(r["triangles"], tri)
    formats.write_ply(tmp_path / "bare.ply", p)
    r = formats.read_ply(tmp_path / "bare.ply")
    assert np.array_equal(r["points"], p) and r["normals"] is None and r["colors"] is None and r["triangles"] is None
    formats.write_ply(tmp_path / "empty.ply", np.zeros((0, 3)))
    assert formats.read_ply(tmp_path / "empty.ply")["points"].shape == (0, 3)          # PoseEstimator.py:73 checks for empty clouds


def test_ply_foreign_layouts(tmp_path):
    """float32 vertices with extra properties, a quad face (fan-triangulated), big-endian binary."""
    txt = ("ply\nformat ascii 1.0\ncomment made by hand\nelement vertex 4\nproperty float x\nproperty float y\nproperty float z\nproperty uchar red\n"
           "property uchar green\nproperty uchar blue\nproperty float quality\nelement face 2\nproperty list uchar int vertex_indices\nend_header\n"
           "0 0 0 255 0 0 0.5\n1 0 0 0 255 0 0.5\n1 1 0 0 0 255 0.5\n0 1 0 10 20 30 0.5\n4 0 1 2 3\n3 0 2 3\n")
    (tmp_path / "a.ply").write_text(txt)
    r = formats.read_ply(tmp_path / "a.ply")
    assert r["points"].shape == (4, 3) and np.array_equal(r["triangles"], [[0, 1, 2], [0, 2, 3], [0, 2, 3]])
    assert np.allclose(r["colors"][3], [10 / 255, 20 / 255, 30 / 255])
    v = np.array([[0.5, 1.5, -2.0], [3.0, 4.0, 5.0]], ">f4")
    hdr = b"ply\nformat binary_big_endian 1.0\nelement vertex 2\nproperty float x\nproperty float y\nproperty float z\nelement face 1\nproperty list uchar uint vertex_index\nend_header\n"
    (tmp_path / "b.ply").write_bytes(hdr + v.tobytes() + b"\x03" + np.array([0, 1, 1], ">u4").tobytes())
    r = formats.read_ply(tmp_path / "b.ply")
    assert np.array_equal(r["points"], v.astype(np.float64)) and np.array_equal(r["triangles"], [[0, 1, 1]])
    (tmp_path / "bad.ply").write_text("not a ply")
    with pytest.raises(ValueError):
        formats.read_ply(tmp_path / "bad.ply")


def test_bop_json(tmp_path):
    cam = {"0": {"cam_K": [572.4, 0.0, 325.3, 0.0, 573.6, 242.0, 0.0, 0.0, 1.0], "depth_scale": 0.1}, "7": {"cam_K": [1, 0, 2, 0, 3, 4, 0, 0, 1], "depth_scale": 1.0}}
    (tmp_path / "scene_camera.json").write_text(json.dumps(cam))
    fx, fy, cx, cy, ds, K = formats.load_camera_intrinsics(tmp_path / "scene_camera.json", 0)
    assert (fx, fy, cx, cy, ds) == (572.4, 573.6, 325.3, 242.0, 0.1) and K == cam["0"]["cam_K"]
    assert formats.load_camera_intrinsics(tmp_path / "scene_camera.json", "7")[:4] == (1, 3, 2, 4)
    with pytest.raises(ValueError):
        formats.load_camera_intrinsics(tmp_path / "scene_camera.json", 3)
    gt = {"5": [{"cam_R_m2c": [0, -1, 0, 1, 0, 0, 0, 0, 1], "cam_t_m2c": [10.0, 20.0, 300.0], "obj_id": 1}], "12": [{"cam_R_m2c": list(np.eye(3).ravel()), "cam_t_m2c": [0, 0, 1], "obj_id": 2}]}
    (tmp_path / "scene_gt.json").write_text(json.dumps(gt))
    T = formats.load_scene_gt(tmp_path / "scene_gt.json")            # sorted(keys)[0] is "12" (string order, as in main_image.py:103)
    assert np.array_equal(T[:3, 3], [0, 0, 1])
    T = formats.load_scene_gt(tmp_path / "scene_gt.json", 5, unit_scale=0.001)
    assert np.allclose(T[:3, 3], [0.01, 0.02, 0.3]) and np.array_equal(T[:3, :3], [[0, -1, 0], [1, 0, 0], [0, 0, 1]])


def test_calib_yaml_matches_reference_load_calib(tmp_path):
    g = np.load(os.path.join(GOLD, "formats_calib.npz"))
    for tag in ("a", "b"):
        p = tmp_path / f"calib_{tag}.yaml"
        p.write_text(str(g[f"calib_{tag}_text"]))
        K, D, T = formats.load_calib(p)
        assert np.array_equal(K, g[f"calib_{tag}_K"]) and np.array_equal(D, g[f"calib_{tag}_D"])
        assert np.allclose(T, g[f"calib_{tag}_T"], atol=1e-15)
