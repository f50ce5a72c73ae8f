"""On-disk formats and the LiDAR -> image step of the reference's callers, headless:

    PLY template clouds / CAD meshes     o3d.io.read_point_cloud / read_triangle_mesh / write_point_cloud
                                         PoseEstimator.py:29, 66-80; template_creation.py:151, 250
    BOP scene_camera.json, scene_gt.json registration_utils.py:26-52 (load_camera_intrinsics), main_image.py:98-115
    calib.yaml (K, D, T | xyz + rpy)     main_seibersdorf.py:55-68 (load_calib)
    LiDAR cloud -> pixels inside a mask  main_seibersdorf.py:70-81 (project_count), 149-162
(paths under /root/reference/src/pose_estimator/).  File parsing is host code (numpy); the projection / selection of the cloud
runs on the GPU (csrc/raster.cu project_select_kernel).  GUI calls of the callers are dropped.
"""
from __future__ import annotations

import ctypes as C
import json

import numpy as np

from . import _lib
from .presteps import PreSteps, _torch

_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2", "ushort": "u2", "uint16": "u2",
              "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4", "float": "f4", "float32": "f4", "double": "f8", "float64": "f8"}


def read_ply(path):
    """ASCII / binary_little_endian / binary_big_endian PLY -> dict(points [n,3] f64, normals, colors (0..1), triangles [m,3] i32),
    absent parts None.  Faces with more than three vertices are fan-triangulated (as Open3D's reader does)."""
    with open(path, "rb") as f:
        data = f.read()
    end = data.find(b"end_header")
    if not data.startswith(b"ply") or end < 0:
        raise ValueError(f"{path}: not a PLY file")
    nl = data.find(b"\n", end)
    header = data[:end].decode("ascii", "replace").split("\n")
    body = data[nl + 1:]
    fmt, elements = None, []
    for line in header:
        tok = line.strip().split()
        if not tok:
            continue
        if tok[0] == "format":
            fmt = tok[1]
        elif tok[0] == "element":
            elements.append({"name": tok[1], "count": int(tok[2]), "props": []})
        elif tok[0] == "property" and elements:
            if tok[1] == "list":
                elements[-1]["props"].append(("list", tok[2], tok[3], tok[4]))
            else:
                elements[-1]["props"].append(("scalar", tok[1], tok[2]))
    if fmt not in ("ascii", "binary_little_endian", "binary_big_endian"):
        raise ValueError(f"{path}: unsupported PLY format {fmt}")
    bo = ">" if fmt == "binary_big_endian" else "<"
    out = {"points": None, "normals": None, "colors": None, "triangles": None}
    pos = 0
    tokens = body.split() if fmt == "ascii" else None
    for el in elements:
        n = el["count"]
        has_list = any(p[0] == "list" for p in el["props"])
        if not has_list:
            names = [p[2] for p in el["props"]]
            if fmt == "ascii":
                k = len(names)
                arr = np.array(tokens[pos:pos + n * k], dtype=np.float64).reshape(n, k)
                pos += n * k
                cols = {nm: arr[:, i] for i, nm in enumerate(names)}
                int_cols = {p[2] for p in el["props"] if _PLY_TYPES[p[1]][0] in "iu"}
            else:
                dt = np.dtype([(p[2], bo + _PLY_TYPES[p[1]]) for p in el["props"]])
                rec = np.frombuffer(body, dt, n, pos)
                pos += n * dt.itemsize
                cols = {nm: rec[nm] for nm in names}
                int_cols = {nm for nm in names if rec[nm].dtype.kind in "iu"}
            if el["name"] == "vertex":
                if all(c in cols for c in ("x", "y", "z")):
                    out["points"] = np.stack([cols["x"], cols["y"], cols["z"]], 1).astype(np.float64)
                if all(c in cols for c in ("nx", "ny", "nz")):
                    out["normals"] = np.stack([cols["nx"], cols["ny"], cols["nz"]], 1).astype(np.float64)
                if all(c in cols for c in ("red", "green", "blue")):
                    rgb = np.stack([cols["red"], cols["green"], cols["blue"]], 1).astype(np.float64)
                    out["colors"] = rgb / 255.0 if "red" in int_cols else rgb
        else:
            tris = []
            for _ in range(n):
                for p in el["props"]:
                    if p[0] == "list":
                        if fmt == "ascii":
                            c = int(tokens[pos]); idx = [int(float(t)) for t in tokens[pos + 1:pos + 1 + c]]; pos += 1 + c
                        else:
                            cd, it = np.dtype(bo + _PLY_TYPES[p[1]]), np.dtype(bo + _PLY_TYPES[p[2]])
                            c = int(np.frombuffer(body, cd, 1, pos)[0]); pos += cd.itemsize
                            idx = np.frombuffer(body, it, c, pos).astype(np.int64).tolist(); pos += c * it.itemsize
                        if el["name"] == "face" and p[3] in ("vertex_indices", "vertex_index"):
                            tris += [(idx[0], idx[k], idx[k + 1]) for k in range(1, c - 1)]
                    else:
                        if fmt == "ascii":
                            pos += 1
                        else:
                            pos += np.dtype(_PLY_TYPES[p[1]]).itemsize
            if el["name"] == "face":
                out["triangles"] = np.array(tris, np.int32).reshape(-1, 3)
    return out


def write_ply(path, points, normals=None, colors=None, triangles=None, ascii=False):
    """Point cloud (or mesh) -> PLY with double coordinates, like o3d.io.write_point_cloud's default binary little-endian file."""
    p = np.ascontiguousarray(np.asarray(points.cpu() if hasattr(points, "cpu") else points), np.float64)
    n = p.shape[0]
    fields = [("x", "<f8"), ("y", "<f8"), ("z", "<f8")]
    cols = [p[:, 0], p[:, 1], p[:, 2]]
    if normals is not None:
        q = np.asarray(normals, np.float64)
        fields += [("nx", "<f8"), ("ny", "<f8"), ("nz", "<f8")]; cols += [q[:, 0], q[:, 1], q[:, 2]]
    if colors is not None:
        c = np.clip(np.rint(np.asarray(colors, np.float64) * 255.0), 0, 255).astype(np.uint8)
        fields += [("red", "u1"), ("green", "u1"), ("blue", "u1")]; cols += [c[:, 0], c[:, 1], c[:, 2]]
    tnames = {"<f8": "double", "u1": "uchar"}
    hdr = ["ply", "format " + ("ascii 1.0" if ascii else "binary_little_endian 1.0"), "comment Created by poseestimator_b200",
           f"element vertex {n}"] + [f"property {tnames[t]} {nm}" for nm, t in fields]
    t = None
    if triangles is not None:
        t = np.ascontiguousarray(triangles, np.int32)
        hdr += [f"element face {t.shape[0]}", "property list uchar int vertex_indices"]
    hdr.append("end_header")
    with open(path, "wb") as f:
        f.write(("\n".join(hdr) + "\n").encode("ascii"))
        if ascii:
            for i in range(n):
                f.write((" ".join(repr(float(c[i])) if c.dtype.kind == "f" else str(int(c[i])) for c in cols) + "\n").encode("ascii"))
            if t is not None:
                for a, b, c in t:
                    f.write(f"3 {a} {b} {c}\n".encode("ascii"))
        else:
            rec = np.empty(n, np.dtype(fields))
            for (nm, _), c in zip(fields, cols):
                rec[nm] = c
            f.write(rec.tobytes())
            if t is not None:
                fr = np.empty(t.shape[0], np.dtype([("c", "u1"), ("i", "<i4", 3)]))
                fr["c"] = 3; fr["i"] = t
                f.write(fr.tobytes())


def load_camera_intrinsics(scene_camera_path, frame_id):
    """BOP / BlenderProc scene_camera.json -> (fx, fy, cx, cy, depth_scale, cam_K list)  (registration_utils.py:26-52)."""
    frame_id = f"{frame_id}" if isinstance(frame_id, int) else frame_id
    with open(scene_camera_path, "r") as f:
        cam = json.load(f)
    if frame_id not in cam:
        raise ValueError(f"Frame ID {frame_id} not found in scene_camera.json")
    K = cam[frame_id]["cam_K"]
    return K[0], K[4], K[2], K[5], cam[frame_id]["depth_scale"], K


def load_scene_gt(scene_gt_path, frame_id=None, obj_index=0, unit_scale=1.0):
    """BOP scene_gt.json -> 4x4 model -> camera pose of one object of one frame (main_image.py:98-115; first frame by default).
    `unit_scale` converts the file's translation unit (BOP: millimetres) when the caller works in metres."""
    with open(scene_gt_path, "r") as f:
        data = json.load(f)
    key = sorted(data.keys())[0] if frame_id is None else f"{frame_id}"
    obj = data[key][obj_index]
    T = np.eye(4)
    T[:3, :3] = np.array(obj["cam_R_m2c"], float).reshape(3, 3)
    T[:3, 3] = np.array(obj["cam_t_m2c"], float).reshape(3) * unit_scale
    return T


def _euler_xyz_matrix(rpy):
    """scipy Rotation.from_euler("xyz", rpy): extrinsic rotations about x, then y, then z  ->  Rz(y) @ Ry(p) @ Rx(r)."""
    r, p, y = [float(v) for v in rpy]
    cx, sx, cy, sy, cz, sz = np.cos(r), np.sin(r), np.cos(p), np.sin(p), np.cos(y), np.sin(y)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]]); Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def load_calib(path):
    """calib.yaml -> (K [3,3], D [k], T [4,4]); T either given or built from xyz + rpy (main_seibersdorf.py:55-68)."""
    import yaml
    with open(path, "r") as f:
        c = yaml.safe_load(f)
    K = np.asarray(c["K"], dtype=float).reshape(3, 3)
    D = np.asarray(c.get("D", []), dtype=float).reshape(-1)
    if "T" in c:
        T = np.asarray(c["T"], dtype=float).reshape(4, 4)
    else:
        assert "xyz" in c and "rpy" in c, "calib.yaml must have T or (xyz+rpy)"
        T = np.eye(4)
        T[:3, :3] = _euler_xyz_matrix(c["rpy"])
        T[:3, 3] = np.array(c["xyz"], dtype=float)
    return K, D, T


class LidarProjector(PreSteps):
    """project_count + the mask selection of project_and_colorize on the GPU."""

    def select_in_mask(self, points, T_cam_from_lidar, K, D, width, height, mask=None, z_min=0.1):
        """points [n,3] (LiDAR frame), T_cam_from_lidar 4x4 (the reference passes inv(T) of the calibration), mask uint8 [H,W] or None.
        -> dict(indices [m] int32 ascending, uv [m,2] int32, points [m,3] CUDA tensor, n_front, n_in_image)."""
        torch = _torch()
        p = self._pts(points)
        n = p.shape[0]
        T = np.asarray(T_cam_from_lidar, float)
        R = np.ascontiguousarray(T[:3, :3], np.float64); t = np.ascontiguousarray(T[:3, 3], np.float64)
        K = np.ascontiguousarray(K, np.float64).reshape(3, 3)
        D = np.ascontiguousarray(D if D is not None else [], np.float64).reshape(-1)
        m_dev = None
        if mask is not None:
            m_dev = mask if isinstance(mask, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(mask).astype(bool).astype(np.uint8)).to(self._dev)
            m_dev = (m_dev != 0).to(torch.uint8).contiguous()
            assert tuple(m_dev.shape) == (height, width)
        idx = self._empty((max(n, 1),), torch.int32)
        uv = self._empty((max(n, 1), 2), torch.int32)
        if n == 0:
            return {"indices": idx[:0], "uv": uv[:0], "points": p, "n_front": 0, "n_in_image": 0}
        nf, ni, ns = C.c_int(0), C.c_int(0), C.c_int(0)
        _lib.check(self._h, _lib.lib().b2r_project_select(self._h, p.data_ptr(), n, R.ctypes.data, t.ctypes.data, K.ctypes.data,
                                                          D.ctypes.data if D.size else None, int(D.size), int(width), int(height),
                                                          m_dev.data_ptr() if m_dev is not None else None, float(z_min), idx.data_ptr(),
                                                          uv.data_ptr(), C.byref(nf), C.byref(ni), C.byref(ns)))
        idx = idx[: ns.value]
        return {"indices": idx, "uv": uv[: ns.value], "points": p[idx.long()], "n_front": nf.value, "n_in_image": ni.value}
