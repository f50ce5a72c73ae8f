"""GPU: LiDAR cloud -> pixels inside the detection mask (csrc/raster.cu project_select_kernel) against vectors produced by the
reference's own project_count (cv2.projectPoints) and selection code (tests/golden/formats_project_*.npz).  Bar: identical index
and pixel arrays."""
import glob
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "formats_project_*.npz")))


@pytest.mark.parametrize("path", GOLD, ids=[os.path.basename(p) for p in GOLD])
def test_project_select_matches_reference(path):
    from poseestimator_b200.formats import LidarProjector
    g = np.load(path)
    lp = LidarProjector(0)
    W, H = int(g["W"]), int(g["H"])
    r = lp.select_in_mask(g["pts"], g["T_cam_from_lidar"], g["K"], g["D"], W, H, mask=None)
    assert r["n_front"] == int(g["n_front"]) and r["n_in_image"] == int(g["n_in"])
    assert np.array_equal(r["indices"].cpu().numpy(), g["idx_inimg"]) and np.array_equal(r["uv"].cpu().numpy(), g["uv_in"])
    m = lp.select_in_mask(g["pts"], g["T_cam_from_lidar"], g["K"], g["D"], W, H, mask=g["mask"])
    assert np.array_equal(m["indices"].cpu().numpy(), g["idx_inmask"]) and np.array_equal(m["uv"].cpu().numpy(), g["uv_inmask"])
    assert np.array_equal(m["points"].cpu().numpy(), g["pts"][g["idx_inmask"]])


def test_project_select_edge_cases():
    from poseestimator_b200.formats import LidarProjector
    lp = LidarProjector(0)
    K = np.array([[500.0, 0, 320], [0, 500.0, 240], [0, 0, 1]])
    behind = np.array([[0.0, 0.0, -1.0], [0.1, 0.1, 0.05]])
    r = lp.select_in_mask(behind, np.eye(4), K, None, 640, 480)
    assert r["n_front"] == 0 and r["indices"].shape[0] == 0
    r = lp.select_in_mask(np.zeros((0, 3)), np.eye(4), K, None, 640, 480)
    assert r["indices"].shape[0] == 0
    edge = np.array([[-0.64, -0.48, 1.0], [0.6398, 0.4798, 1.0], [0.64, 0.0, 1.0], [0.0, 0.48, 1.0]])   # u = 0, 639.9, 640, v = 480
    r = lp.select_in_mask(edge, np.eye(4), K, None, 640, 480)
    assert r["indices"].cpu().numpy().tolist() == [0, 1] and r["uv"].cpu().numpy().tolist() == [[0, 0], [639, 479]]
