"""CPU oracle for the pre-steps (back-projection, outlier removal, voxel, normals, FPFH, mutual NN) -- TEST
INFRASTRUCTURE ONLY.  ctypes front-end of oracle/_build/liboracle_presteps.so (oracle/presteps_oracle.cpp, SURVEY.md
Appendix B).  PARITY UNPINNED (Open3D is not on this machine); see the header of the .cpp file."""
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
        L = C.CDLL(os.path.join(_BUILD, "liboracle_presteps.so"))
        vp, i, d = C.c_void_p, C.c_int, C.c_double
        L.orcp_backproject.restype = i
        L.orcp_backproject.argtypes = [vp, vp, i, i, i, i, i, i, d, d, d, d, d, d, d, d, vp, vp]
        L.orcp_knn.argtypes = [vp, i, vp, i, i, d, vp, vp, vp]
        L.orcp_statistical_outlier.restype = i
        L.orcp_statistical_outlier.argtypes = [vp, i, i, d, vp, vp]
        L.orcp_voxel_down.restype = i
        L.orcp_voxel_down.argtypes = [vp, i, d, vp, vp, vp]
        L.orcp_normals.argtypes = [vp, i, d, i, i, vp, vp]
        L.orcp_fpfh.argtypes = [vp, vp, i, d, i, vp, vp]
        L.orcp_mutual_nn.restype = i
        L.orcp_mutual_nn.argtypes = [vp, i, vp, i, i, vp, vp, vp]
        L.orcp_smallest_eigvec.argtypes = [vp, vp]
        L.orcp_num_threads.restype = i
        _lib = L
    return _lib


def _pts(p):
    p = np.ascontiguousarray(p, np.float64)
    assert p.ndim == 2 and p.shape[1] == 3, p.shape
    return p


def backproject(depth, mask, bbox, fx, fy, cx, cy, depth_scale, depth_trunc=3.0, zmin=0.0, zmax=0.0):
    """-> (points [M,3] f64, pixel ids [M] i32), row-major scan order inside bbox=(u0,v0,u1,v1)."""
    depth = np.ascontiguousarray(depth, np.uint16)
    H, W = depth.shape
    if mask is not None:
        mask = np.ascontiguousarray(mask, np.uint8)
        assert mask.shape == depth.shape
    u0, v0, u1, v1 = bbox if bbox is not None else (0, 0, W, H)
    cap = max(0, (min(u1, W) - max(u0, 0))) * max(0, (min(v1, H) - max(v0, 0)))
    pts = np.zeros((max(cap, 1), 3))
    pix = np.zeros(max(cap, 1), np.int32)
    m = lib().orcp_backproject(depth.ctypes.data, mask.ctypes.data if mask is not None else None, H, W, u0, v0, u1, v1,
                               fx, fy, cx, cy, depth_scale, depth_trunc, zmin, zmax, pts.ctypes.data, pix.ctypes.data)
    return pts[:m].copy(), pix[:m].copy()


def knn(points, queries, max_nn, radius=0.0):
    p, q = _pts(points), _pts(queries)
    K = int(max_nn)
    idx = np.zeros((q.shape[0], K), np.int32)
    d2 = np.zeros((q.shape[0], K))
    cnt = np.zeros(q.shape[0], np.int32)
    lib().orcp_knn(p.ctypes.data, p.shape[0], q.ctypes.data, q.shape[0], K, float(radius), idx.ctypes.data, d2.ctypes.data,
                   cnt.ctypes.data)
    return idx, d2, cnt


def statistical_outlier(points, nb_neighbors=20, std_ratio=1.0):
    """-> (keep mask [n] bool, avg neighbour distance [n])."""
    p = _pts(points)
    keep = np.zeros(p.shape[0], np.uint8)
    avg = np.zeros(p.shape[0])
    lib().orcp_statistical_outlier(p.ctypes.data, p.shape[0], nb_neighbors, std_ratio, keep.ctypes.data, avg.ctypes.data)
    return keep.astype(bool), avg


def voxel_down(points, voxel):
    """-> (means [V,3], voxel id of every point [n], points per voxel [V]); voxels in ascending key order."""
    p = _pts(points)
    n = p.shape[0]
    out = np.zeros((max(n, 1), 3))
    vox = np.zeros(max(n, 1), np.int32)
    cnt = np.zeros(max(n, 1), np.int32)
    V = lib().orcp_voxel_down(p.ctypes.data, n, float(voxel), out.ctypes.data, vox.ctypes.data, cnt.ctypes.data)
    return out[:V].copy(), vox[:n].copy(), cnt[:V].copy()


def normals(points, radius, max_nn, orient=0, view=(0.0, 0.0, 0.0)):
    p = _pts(points)
    out = np.zeros_like(p)
    v = np.ascontiguousarray(view, np.float64)
    lib().orcp_normals(p.ctypes.data, p.shape[0], float(radius), int(max_nn), int(orient), v.ctypes.data, out.ctypes.data)
    return out


def fpfh(points, nrm, radius, max_nn=100, want_spfh=False):
    """-> Feature.data layout [33, n] f64."""
    p, q = _pts(points), _pts(nrm)
    n = p.shape[0]
    out = np.zeros((33, n))
    sp = np.zeros((33, n)) if want_spfh else None
    lib().orcp_fpfh(p.ctypes.data, q.ctypes.data, n, float(radius), int(max_nn), out.ctypes.data,
                    sp.ctypes.data if want_spfh else None)
    return (out, sp) if want_spfh else out


def mutual_nn(fa, fb):
    """fa [na, dim], fb [nb, dim] (one descriptor per row) -> (pairs [c,2], nn01 [na], nn10 [nb])."""
    fa = np.ascontiguousarray(fa, np.float64)
    fb = np.ascontiguousarray(fb, np.float64)
    na, dim = fa.shape
    nb = fb.shape[0]
    assert fb.shape[1] == dim
    nn01 = np.zeros(na, np.int32)
    nn10 = np.zeros(nb, np.int32)
    pairs = np.zeros((max(na, 1), 2), np.int32)
    c = lib().orcp_mutual_nn(fa.ctypes.data, na, fb.ctypes.data, nb, dim, nn01.ctypes.data, nn10.ctypes.data, pairs.ctypes.data)
    return pairs[:c].copy(), nn01, nn10


def smallest_eigvec(Cm):
    Cm = np.ascontiguousarray(Cm, np.float64)
    out = np.zeros(3)
    lib().orcp_smallest_eigvec(Cm.ctypes.data, out.ctypes.data)
    return out
