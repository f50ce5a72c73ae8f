#!/usr/bin/env python
"""Regenerates tests/golden/formats_*.npz by executing the reference's own functions (this container only; needs cv2, yaml, scipy):
    main_seibersdorf.py:55-68   load_calib        (on two calib.yaml texts written to a temp dir)
    main_seibersdorf.py:70-81   project_count     (cv2.projectPoints on float32 points, int32 truncation, image test)
    main_seibersdorf.py:149-162 the index / mask selection that follows it (restated inline below, five lines of numpy)
(paths under /root/reference/src/pose_estimator/).  PINS poseestimator_b200/formats.py and the CUDA project_select kernel.

    python tests/golden/make_golden_formats.py
"""
import ast
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/src/pose_estimator/main_seibersdorf.py"

CALIB_A = "K: [600.0, 0.0, 320.5, 0.0, 610.0, 241.25, 0.0, 0.0, 1.0]\nD: [0.08, -0.04, 0.001, -0.002, 0.01]\nxyz: [0.1, -0.05, 0.2]\nrpy: [0.3, -0.2, 1.1]\n"
CALIB_B = "K: [[500.0, 0.0, 300.0], [0.0, 505.0, 200.0], [0.0, 0.0, 1.0]]\nT: [[0.0, -1.0, 0.0, 0.1], [0.0, 0.0, -1.0, 0.2], [1.0, 0.0, 0.0, -0.3], [0.0, 0.0, 0.0, 1.0]]\n"


def load(names):
    import cv2
    import yaml
    from scipy.spatial.transform import Rotation as R
    ns = {"np": np, "cv2": cv2, "yaml": yaml, "R": R}
    for node in ast.parse(open(SRC).read()).body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            exec(compile(ast.Module([node], []), SRC, "exec"), ns)
    return ns


def main():
    if not os.path.exists(SRC):
        raise SystemExit("the reference is not mounted here; the committed .npz files are the product of this script")
    ns = load(["load_calib", "project_count"])
    out = {"calib_a_text": np.array(CALIB_A), "calib_b_text": np.array(CALIB_B)}
    with tempfile.TemporaryDirectory() as d:
        for tag, text in (("a", CALIB_A), ("b", CALIB_B)):
            p = os.path.join(d, f"calib_{tag}.yaml")
            open(p, "w").write(text)
            K, D, T = ns["load_calib"](p)
            out[f"calib_{tag}_K"], out[f"calib_{tag}_D"], out[f"calib_{tag}_T"] = K, D, T
    np.savez_compressed(os.path.join(HERE, "formats_calib.npz"), **out)
    rng = np.random.default_rng(17)
    W, H = 640, 480
    for tag, n, D in (("d5", 6000, np.array([0.08, -0.04, 0.001, -0.002, 0.01])), ("d0", 2500, np.array([])),
                      ("d8", 3000, np.array([0.05, -0.02, 0.001, 0.0005, 0.003, 0.01, -0.004, 0.0007])), ("d3", 1500, np.array([0.1, 0.2, 0.3]))):
        pts = rng.uniform(-6, 6, (n, 3)) * np.array([1.0, 1.0, 0.4])
        K = np.array([[600.0, 0, 320.5], [0, 610.0, 241.25], [0, 0, 1.0]])
        Tc = np.eye(4); Q = np.linalg.qr(rng.normal(size=(3, 3)))[0]; Q *= np.sign(np.linalg.det(Q)); Tc[:3, :3] = Q; Tc[:3, 3] = rng.uniform(-0.5, 0.5, 3)
        Tinv = np.linalg.inv(Tc)
        Rinv, tinv = Tinv[:3, :3].astype(np.float64), Tinv[:3, 3].astype(np.float64)
        n_in, front, uv, in_img = ns["project_count"](pts, Rinv, tinv, K, D, W, H)
        vv, uu = np.mgrid[0:H, 0:W]
        mask = (((uu - 300) ** 2 + (vv - 250) ** 2) < 150 ** 2).astype(np.uint8) * 255
        idx_front = np.where(front)[0]; idx_inimg = idx_front[in_img]; uv_in = uv[in_img]       # main_seibersdorf.py:150-160
        inside = mask.astype(bool)[uv_in[:, 1], uv_in[:, 0]]
        np.savez_compressed(os.path.join(HERE, f"formats_project_{tag}.npz"), pts=pts, T_cam_from_lidar=Tinv, K=K, D=D, W=W, H=H, mask=mask,
                            n_in=n_in, n_front=int(front.sum()), idx_inimg=idx_inimg.astype(np.int32), uv_in=uv_in.astype(np.int32),
                            idx_inmask=idx_inimg[inside].astype(np.int32), uv_inmask=uv_in[inside].astype(np.int32))
        print(tag, n, "front", int(front.sum()), "in image", n_in, "in mask", int(inside.sum()))


if __name__ == "__main__":
    main()
