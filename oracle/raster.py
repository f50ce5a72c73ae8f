"""CPU oracle for the CAD depth render -- TEST INFRASTRUCTURE ONLY.  ctypes front-end of
oracle/_build/liboracle_raster.so (oracle/raster_oracle.cpp).  PARITY UNPINNED (Filament / OpenGL: unavailable on this box)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _BUILD, build

_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(os.path.join(_BUILD, "liboracle_raster.so"))
        vp, i, d = C.c_void_p, C.c_int, C.c_double
        L.orcx_render_depth.argtypes = [vp, i, vp, i, vp, d, d, d, d, i, i, d, d, vp]
        L.orcx_backproject_depth32.restype = i
        L.orcx_backproject_depth32.argtypes = [vp, i, i, d, d, d, d, d, d, vp, vp]
        _lib = L
    return _lib


def render_depth(verts, tris, T_m2c, fx, fy, cx, cy, W, H, z_near, z_far):
    """-> depth float32 [H,W]: view-space z of the nearest surface along the ray of every pixel, 0 = nothing."""
    v = np.ascontiguousarray(verts, np.float64); t = np.ascontiguousarray(tris, np.int32); T = np.ascontiguousarray(T_m2c, np.float64)
    out = np.zeros((H, W), np.float32)
    lib().orcx_render_depth(v.ctypes.data, v.shape[0], t.ctypes.data, t.shape[0], T.ctypes.data, fx, fy, cx, cy, W, H, z_near, z_far, out.ctypes.data)
    return out


def backproject_depth32(depth, fx, fy, cx, cy, depth_scale=1.0, depth_trunc=0.0):
    """-> (points [M,3] f64, pixel ids [M] i32), row-major order."""
    d = np.ascontiguousarray(depth, np.float32)
    H, W = d.shape
    pts = np.zeros((H * W, 3)); pix = np.zeros(H * W, np.int32)
    m = lib().orcx_backproject_depth32(d.ctypes.data, H, W, fx, fy, cx, cy, depth_scale, depth_trunc, pts.ctypes.data, pix.ctypes.data)
    return pts[:m].copy(), pix[:m].copy()
