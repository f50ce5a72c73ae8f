"""CAD depth rendering and template creation on the GPU.

The reference renders with Open3D's Filament OffscreenRenderer (OpenGL), which cannot run here or on the GPU box
("unavailable on this box"); the depth image -- the only output that feeds the registration path -- comes from the CUDA
rasteriser of csrc/raster.cu instead.  Host-side geometry helpers keep the reference's names and conventions:

    fx_from_fov, o3d_lookat, get_reduced_camera_positions, sample_n_points     template_creation.py:7-9, 34-62, 65-132
    render_templates                                                            template_creation.py:139-250
    camera_eye_lookat_up_from_H                                                 HelpersRealtime.py:226-247
    PoseEstimator.create_template_from_H                                        PoseEstimator.py:118-141
(paths under /root/reference/src/pose_estimator/EstimHelpers/).  Colour images, windows and file dialogs are left out.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .presteps import PreSteps, _torch

# OpenCV camera (x right, y down, z forward: what back-projection assumes) <-> OpenGL camera (x right, y up, z backward)
_FLIP_YZ = np.diag([1.0, -1.0, -1.0, 1.0])


def fx_from_fov(fov_deg: float, width: float) -> float:
    """Focal length in pixels of a pinhole camera whose horizontal field of view is `fov_deg`."""
    return 0.5 * width / np.tan(np.deg2rad(fov_deg) / 2.0)


def o3d_lookat(eye, target, up) -> np.ndarray:
    """World -> camera 4x4 of an OpenGL-style camera at `eye` looking at `target` (+Z points out of the camera, towards the viewer)."""
    eye = np.asarray(eye, float); target = np.asarray(target, float); up = np.asarray(up, float)
    back = eye - target
    back = back / (np.linalg.norm(back) + 1e-12)
    right = np.cross(up, back)
    right = right / (np.linalg.norm(right) + 1e-12)
    true_up = np.cross(back, right)
    T = np.eye(4)
    T[:3, :3] = np.vstack([right, true_up, back])
    T[:3, 3] = -T[:3, :3] @ eye
    return T


def get_reduced_camera_positions(distance: float, center=None):
    """The five template views of the reference: three edge-on directions and two corner directions, `distance` from the
    origin, target the origin, +Y up.  `type` keeps the reference's labels ('6', '7', '8', '11', '12')."""
    dirs = [((0, 1, 1), "6"), ((0, -1, 1), "7"), ((1, 0, 1), "8"), ((1, 1, 1), "11"), ((1, -1, 1), "12")]
    out = []
    for d, name in dirs:
        d = np.array(d, float)
        d = d / np.linalg.norm(d)
        out.append({"eye": d * distance, "target": np.array([0, 0, 0]), "up": np.array([0, 1, 0]), "type": name})
    return out


def camera_eye_lookat_up_from_H(H):
    """Model -> camera pose H (OpenCV camera) -> (eye, target, up) in model coordinates."""
    H = np.asarray(H, float)
    R, t = H[:3, :3], H[:3, 3]
    eye = -R.T @ t
    forward = R.T @ np.array([0.0, 0.0, 1.0])
    up = R.T @ np.array([0.0, -1.0, 0.0])
    return eye, eye + forward, up / (np.linalg.norm(up) + 1e-12)


def model_to_camera_from_lookat(eye, target, up) -> np.ndarray:
    """Model -> OpenCV-camera pose of the view `scene.camera.look_at(target, eye, up)` sets: flip(y, z) . o3d_lookat."""
    return _FLIP_YZ @ o3d_lookat(eye, target, up)


def sample_n_points(points, n=200, rng=None):
    """Random subset of n points without replacement (np.random.choice in the reference; pass `rng` for a seeded draw).
    Returns (subset, indices); clouds with <= n points come back unchanged."""
    m = int(points.shape[0])
    if m <= n:
        return points, np.arange(m)
    idx = (rng or np.random).choice(m, n, replace=False)
    if hasattr(points, "is_cuda"):
        torch = _torch()
        return points[torch.from_numpy(np.asarray(idx, np.int64)).to(points.device)], idx
    return points[idx], idx


class DepthRenderer(PreSteps):
    """A triangle mesh resident on the device + a pinhole camera; every call renders one depth image (float32, metres, 0 = empty)."""

    def __init__(self, vertices, triangles, width=640, height=480, fx=None, fy=None, cx=None, cy=None, fov_deg=60.0, device: int = 0):
        super().__init__(device)
        torch = _torch()
        self.W, self.H = int(width), int(height)
        self.fx = float(fx) if fx is not None else float(fx_from_fov(fov_deg, self.W))   # template_creation.py:189-191
        self.fy = float(fy) if fy is not None else self.fx
        self.cx = float(cx) if cx is not None else self.W / 2.0
        self.cy = float(cy) if cy is not None else self.H / 2.0
        self.set_mesh(vertices, triangles)
        self._torch = torch

    def set_mesh(self, vertices, triangles):
        torch = _torch()
        v = np.ascontiguousarray(vertices, np.float64); t = np.ascontiguousarray(triangles, np.int32)
        if v.ndim != 2 or v.shape[1] != 3 or t.ndim != 2 or t.shape[1] != 3:
            raise TypeError("vertices [nv,3] float64 and triangles [nt,3] int32 expected")
        if t.size and (t.min() < 0 or t.max() >= v.shape[0]):
            raise ValueError("triangle index out of range")
        self.vertices = torch.from_numpy(v).to(self._dev)
        self.triangles = torch.from_numpy(t).to(self._dev)

    def render_depth(self, T_m2c, z_near=0.01, z_far=5.0):
        """render_to_depth_image(z_in_view_space=True) at the model -> camera pose T_m2c -> CUDA float32 [H,W]."""
        torch = _torch()
        T = np.ascontiguousarray(T_m2c, np.float64)
        assert T.shape == (4, 4)
        depth = self._empty((self.H, self.W), torch.float32)
        _lib.check(self._h, _lib.lib().b2r_render_depth(self._h, self.vertices.data_ptr(), self.vertices.shape[0], self.triangles.data_ptr(),
                                                        self.triangles.shape[0], T.ctypes.data, self.fx, self.fy, self.cx, self.cy, self.W, self.H,
                                                        float(z_near), float(z_far), depth.data_ptr()))
        return depth

    def backproject_depth(self, depth, depth_scale=1.0, depth_trunc=0.0, want_pixels=False):
        """create_from_color_and_depth(depth_scale, depth_trunc) + create_from_rgbd_image -> points [M,3] f64 (camera frame)."""
        torch = _torch()
        if not (isinstance(depth, torch.Tensor) and depth.is_cuda and depth.dtype == torch.float32):
            depth = torch.from_numpy(np.ascontiguousarray(depth, np.float32)).to(self._dev)
        depth = depth.contiguous()
        H, W = depth.shape
        cap = H * W
        pts = self._empty((cap, 3), torch.float64)
        pix = self._empty((cap,), torch.int32) if want_pixels else None
        cnt = C.c_int(0)
        _lib.check(self._h, _lib.lib().b2r_backproject_depth32(self._h, depth.data_ptr(), H, W, self.fx, self.fy, self.cx, self.cy, float(depth_scale),
                                                               float(depth_trunc), pts.data_ptr(), pix.data_ptr() if want_pixels else None, cap,
                                                               C.byref(cnt)))
        m = cnt.value
        return (pts[:m], pix[:m]) if want_pixels else pts[:m]

    def create_template_from_H(self, T_m2c, target_points, rng=None, near=0.01, far=5.0):
        """PoseEstimator.create_template_from_H: render the CAD at the current pose, back-project (depth_trunc = far), keep at
        most `target_points` random points.  -> points [m,3] f64 CUDA tensor in the CAMERA frame."""
        depth = self.render_depth(T_m2c, near, far)
        pts = self.backproject_depth(depth, 1.0, far)
        if pts.shape[0] > target_points:
            pts, _ = sample_n_points(pts, target_points, rng)
        return pts


def render_templates(vertices, triangles, n_points=10000, rng=None, device: int = 0, width=640, height=480, fov_deg=60.0):
    """template_creation.render_templates without the files: millimetre meshes are scaled to metres, the mesh is centred, five
    views are rendered from 2x the bounding-box diagonal, every depth image is back-projected, sub-sampled to `n_points` and
    moved back into the model frame.  -> list of dicts {points (numpy [m,3], model frame), type, T_m2c, depth (numpy)}."""
    v = np.array(vertices, np.float64)
    size = v.max(0) - v.min(0)
    if not np.max(size) < 1:                   # template_creation.py:157-162: anything >= 1 is taken to be in millimetres
        v = v * 0.001
    ext = v.max(0) - v.min(0)
    distance = float(np.linalg.norm(ext)) * 2.0
    # TriangleMesh.get_center(): mean of the vertices
    center = v.mean(0)
    v = v - center
    r = DepthRenderer(v, triangles, width, height, fov_deg=fov_deg, device=device)
    near, far = 0.001, distance * 3
    out = []
    for cam in get_reduced_camera_positions(distance):
        T_w2gl = o3d_lookat(cam["eye"], cam["target"], cam["up"])
        T_m2c = _FLIP_YZ @ T_w2gl
        depth = r.render_depth(T_m2c, near, far)
        pts = r.backproject_depth(depth, 1.0, far)
        pts, _ = sample_n_points(pts, n_points, rng)
        p = pts.cpu().numpy()
        # camera (OpenCV) -> GL camera -> model, then undo the centring (template_creation.py:236-245)
        p = p @ _FLIP_YZ[:3, :3].T
        T_gl2w = np.linalg.inv(T_w2gl)
        p = p @ T_gl2w[:3, :3].T + T_gl2w[:3, 3] + center
        out.append({"points": p, "type": cam["type"], "T_m2c": T_m2c, "depth": depth.cpu().numpy(), "center": center})
    return out
