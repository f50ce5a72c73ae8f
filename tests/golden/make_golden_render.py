#!/usr/bin/env python
"""Regenerates tests/golden/render_helpers.npz by executing the reference's own pure-numpy camera helpers (this container only):
    template_creation.py:7-9, 52-62, 65-132   fx_from_fov, o3d_lookat, get_reduced_camera_positions
    HelpersRealtime.py:226-247                camera_eye_lookat_up_from_H
    registration_utils.py:21-23               get_angular_error
(paths under /root/reference/src/pose_estimator/EstimHelpers/), cut out with `ast` as in make_golden_refine.py.  These vectors PIN
the host-side geometry of poseestimator_b200/render.py; the rendered pixels themselves (Filament) stay unpinned.

    python tests/golden/make_golden_render.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden_refine import REF, load_functions  # noqa: E402


def main():
    if not os.path.isdir(REF):
        raise SystemExit("the reference is not mounted here; the committed .npz is the product of this script")
    tc = load_functions(os.path.join(REF, "template_creation.py"), ["fx_from_fov", "o3d_lookat", "get_reduced_camera_positions"])
    hr = load_functions(os.path.join(REF, "HelpersRealtime.py"), ["camera_eye_lookat_up_from_H"])
    rng = np.random.default_rng(42)
    out = {"fov": np.array([60.0, 45.0, 90.0]), "width": np.array([640.0, 1280.0, 320.0])}
    out["fx"] = np.array([tc["fx_from_fov"](f, w) for f, w in zip(out["fov"], out["width"])])
    cams = tc["get_reduced_camera_positions"](distance=0.73, center=np.array([0, 0, 0]))
    out["cam_eye"] = np.array([c["eye"] for c in cams]); out["cam_target"] = np.array([c["target"] for c in cams], float)
    out["cam_up"] = np.array([c["up"] for c in cams], float); out["cam_type"] = np.array([c["type"] for c in cams])
    out["cam_lookat"] = np.array([tc["o3d_lookat"](c["eye"], c["target"], c["up"]) for c in cams])
    eyes = rng.uniform(-2, 2, (6, 3)); tgts = rng.uniform(-0.3, 0.3, (6, 3)); ups = rng.normal(size=(6, 3))
    out["la_eye"], out["la_target"], out["la_up"] = eyes, tgts, ups
    out["la_T"] = np.array([tc["o3d_lookat"](e, t, u) for e, t, u in zip(eyes, tgts, ups)])
    Hs = []
    for _ in range(5):
        Q = np.linalg.qr(rng.normal(size=(3, 3)))[0]
        Q *= np.sign(np.linalg.det(Q))
        H = np.eye(4); H[:3, :3] = Q; H[:3, 3] = rng.uniform(-1, 1, 3)
        Hs.append(H)
    out["H"] = np.array(Hs)
    elu = [hr["camera_eye_lookat_up_from_H"](H) for H in Hs]
    out["H_eye"] = np.array([e[0] for e in elu]); out["H_target"] = np.array([e[1] for e in elu]); out["H_up"] = np.array([e[2] for e in elu])
    np.savez_compressed(os.path.join(HERE, "render_helpers.npz"), **out)
    print({k: np.shape(v) for k, v in out.items()})


if __name__ == "__main__":
    main()
