"""Pre-steps of the registration path on the GPU: what the reference does with Open3D between the detector's mask and
`solver.solve` (SURVEY.md rows a-1a .. a-2).

    back-projection        /root/reference/src/pose_estimator/EstimHelpers/RealSenseClass.py:56-67, registration_utils.py:54-90
    statistical outliers   RealSenseClass.py:68, registration_utils.py:93
    voxel_down_sample      PoseEstimator.py:87,93
    estimate_normals       HelpersRealtime.py:215-219
    compute_fpfh_feature   HelpersRealtime.py:221-224
    correspondences        HelpersRealtime.py:194-213  (north-star: mutual nearest neighbour on FPFH)

Thin host layer over the C ABI (include/b2r.h).  torch is used only to own device memory; every result is a CUDA
tensor.  There is no CPU fallback: without libb2r.so or a CUDA device this module raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


_TORCH = None


def _torch():
    """torch, once a CUDA device has been seen (the check costs ~4 us and a frame's wrappers ask 30 times: cached after success)."""
    global _TORCH
    if _TORCH is None:
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("poseestimator_b200.presteps: no CUDA device is visible and there is no CPU fallback")
        _TORCH = torch
    return _TORCH


class PreSteps:
    """One handle = one device + one stream (the torch current stream of that device at construction)."""

    def __init__(self, device: int = 0, stream=None):
        torch = _torch()
        L = _lib.lib()
        self.device = int(device)
        self._dev = torch.device("cuda", self.device)
        if stream is None:
            # torch's default stream is the legacy stream (handle 0); the C ABI reads NULL as "make a private stream",
            # so name the legacy stream explicitly (cudaStreamLegacy == 0x1) to stay ordered with torch's own work
            stream = torch.cuda.current_stream(self._dev).cuda_stream or 1
        self._h = C.c_void_p()
        st = L.b2r_create(None, self.device, C.c_void_p(int(stream) if stream else None), C.byref(self._h))
        if st != 0:
            raise RuntimeError(f"poseestimator_b200: b2r_create failed ({_lib.STATUS_NAMES.get(st, st)})")

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                _lib.lib().b2r_destroy(h)
            except Exception:  # noqa: BLE001
                pass
            self._h = None

    # ---- helpers ----------------------------------------------------------------------------------------------
    def _pts(self, p):
        torch = _torch()
        if not isinstance(p, torch.Tensor):
            p = torch.from_numpy(np.ascontiguousarray(p, np.float64)).to(self._dev)
        if p.dtype != torch.float64 or p.ndim != 2:
            raise TypeError("expected a float64 [n, d] array")
        if p.device != self._dev:   # a raw pointer of another device would fault inside the handle's kernels
            raise TypeError(f"tensor lives on {p.device} but this handle is bound to {self._dev}")
        return p.contiguous()

    def _empty(self, shape, dtype):
        return _torch().empty(shape, dtype=dtype, device=self._dev)

    def launch_count(self) -> int:
        return int(_lib.lib().b2r_launch_count(self._h))

    # ---- a-1a -------------------------------------------------------------------------------------------------
    def backproject(self, depth, mask, bbox, fx, fy, cx, cy, depth_scale, depth_trunc=3.0, zmin=0.0, zmax=0.0, want_pixels=False):
        """depth u16 [H,W] (numpy host image or CUDA tensor), mask u8 [H,W] or None, bbox=(u0,v0,u1,v1) or None.
        -> points [M,3] float64 CUDA tensor in row-major scan order (and pixel ids [M] int32 when want_pixels)."""
        torch = _torch()
        on_dev = isinstance(depth, torch.Tensor)
        if on_dev:
            if depth.dtype not in (torch.uint16, torch.int16) or not depth.is_cuda:
                raise TypeError("depth tensor must be a CUDA uint16 image")
            depth = depth.contiguous()
            H, W = depth.shape
            dptr = depth.data_ptr()
            if mask is not None:
                mask = mask.contiguous()
                assert mask.is_cuda and mask.dtype == torch.uint8 and tuple(mask.shape) == (H, W)
            mptr = mask.data_ptr() if mask is not None else None
        else:
            depth = np.ascontiguousarray(depth, np.uint16)
            H, W = depth.shape
            dptr = depth.ctypes.data
            if mask is not None:
                mask = np.ascontiguousarray(mask, np.uint8)
                assert mask.shape == depth.shape
            mptr = mask.ctypes.data if mask is not None else None
        u0, v0, u1, v1 = bbox if bbox is not None else (0, 0, W, H)
        cap = max(0, min(u1, W) - max(u0, 0)) * max(0, min(v1, H) - max(v0, 0))
        pts = self._empty((max(cap, 1), 3), torch.float64)
        pix = self._empty((max(cap, 1),), torch.int32) if want_pixels else None
        cam = _lib.B2RCamera(fx, fy, cx, cy, depth_scale, depth_trunc, zmin, zmax)
        cnt = C.c_int(0)
        _lib.check(self._h, _lib.lib().b2r_backproject(self._h, C.c_void_p(dptr), C.c_void_p(mptr) if mptr else None, H, W,
                                                       int(u0), int(v0), int(u1), int(v1), C.byref(cam), int(on_dev),
                                                       C.c_void_p(pts.data_ptr()), C.c_void_p(pix.data_ptr()) if want_pixels else None,
                                                       cap, C.byref(cnt)))
        m = cnt.value
        return (pts[:m], pix[:m]) if want_pixels else pts[:m]

    def knn(self, points, queries, max_nn, radius=0.0):
        torch = _torch()
        p, q = self._pts(points), self._pts(queries)
        K = int(max_nn)
        idx = self._empty((q.shape[0], K), torch.int32)
        d2 = self._empty((q.shape[0], K), torch.float64)
        cnt = self._empty((q.shape[0],), torch.int32)
        _lib.check(self._h, _lib.lib().b2r_knn(self._h, p.data_ptr(), p.shape[0], q.data_ptr(), q.shape[0], K, float(radius),
                                               idx.data_ptr(), d2.data_ptr(), cnt.data_ptr()))
        return idx, d2, cnt

    def statistical_outlier(self, points, nb_neighbors=20, std_ratio=1.0):
        """-> (kept points [k,3], kept indices [k] int32, keep mask [n] uint8, average neighbour distance [n])."""
        torch = _torch()
        p = self._pts(points)
        n = p.shape[0]
        keep = self._empty((max(n, 1),), torch.uint8)
        avg = self._empty((max(n, 1),), torch.float64)
        out = self._empty((max(n, 1), 3), torch.float64)
        idx = self._empty((max(n, 1),), torch.int32)
        k = C.c_int(0)
        _lib.check(self._h, _lib.lib().b2r_statistical_outlier(self._h, p.data_ptr(), n, int(nb_neighbors), float(std_ratio),
                                                               keep.data_ptr(), avg.data_ptr(), out.data_ptr(), idx.data_ptr(), C.byref(k)))
        return out[: k.value], idx[: k.value], keep[:n], avg[:n]

    # ---- a-1b -------------------------------------------------------------------------------------------------
    def voxel_down(self, points, voxel):
        """-> (voxel means [V,3] in ascending key order, voxel id of every point [n], points per voxel [V])."""
        torch = _torch()
        p = self._pts(points)
        n = p.shape[0]
        out = self._empty((max(n, 1), 3), torch.float64)
        vox = self._empty((max(n, 1),), torch.int32)
        cnt = self._empty((max(n, 1),), torch.int32)
        V = C.c_int(0)
        _lib.check(self._h, _lib.lib().b2r_voxel_down(self._h, p.data_ptr(), n, float(voxel), out.data_ptr(), vox.data_ptr(),
                                                      cnt.data_ptr(), C.byref(V)))
        return out[: V.value], vox[:n], cnt[: V.value]

    # ---- a-1c / a-1d ------------------------------------------------------------------------------------------
    def normals(self, points, radius, max_nn, orient=0, view=(0.0, 0.0, 0.0)):
        torch = _torch()
        p = self._pts(points)
        out = self._empty(tuple(p.shape), torch.float64)
        v = (C.c_double * 3)(*[float(x) for x in view])
        _lib.check(self._h, _lib.lib().b2r_normals(self._h, p.data_ptr(), p.shape[0], float(radius), int(max_nn), int(orient), v,
                                                   out.data_ptr()))
        return out

    def fpfh(self, points, normals, radius, max_nn=100, want_spfh=False):
        """-> Feature.data layout [33, n] float64 (and the SPFH in the same layout when want_spfh)."""
        torch = _torch()
        p, q = self._pts(points), self._pts(normals)
        n = p.shape[0]
        out = self._empty((33, n), torch.float64)
        sp = self._empty((33, n), torch.float64) if want_spfh else None
        _lib.check(self._h, _lib.lib().b2r_fpfh(self._h, p.data_ptr(), q.data_ptr(), n, float(radius), int(max_nn), out.data_ptr(),
                                                sp.data_ptr() if want_spfh else None))
        return (out, sp) if want_spfh else out

    # ---- a-2 --------------------------------------------------------------------------------------------------
    def nn_exact(self, fq, fr):
        """fq [nq, dim], fr [nr, dim] (one descriptor per row) -> (nearest row of fr for every row of fq [nq] int32, its d2)."""
        torch = _torch()
        a, b = self._pts(fq), self._pts(fr)
        assert a.shape[1] == b.shape[1]
        nn = self._empty((a.shape[0],), torch.int32)
        dist = self._empty((a.shape[0],), torch.float64)
        _lib.check(self._h, _lib.lib().b2r_nn_exact(self._h, a.data_ptr(), a.shape[0], b.data_ptr(), b.shape[0], a.shape[1],
                                                    nn.data_ptr(), dist.data_ptr()))
        return nn, dist

    def mutual_pairs(self, nn01, nn10):
        torch = _torch()
        na, nb = nn01.shape[0], nn10.shape[0]
        pairs = self._empty((max(na, 1), 2), torch.int32)
        c = C.c_int(0)
        _lib.check(self._h, _lib.lib().b2r_mutual_pairs(self._h, nn01.data_ptr(), na, nn10.data_ptr(), nb, pairs.data_ptr(), C.byref(c)))
        return pairs[: c.value]

    def nn_gemm(self, fq, fr):
        """Same result as nn_exact through the tcgen05 3xTF32 GEMM shortlist + exact re-rank.  -> (nn, d2)."""
        torch = _torch()
        a, b = self._pts(fq), self._pts(fr)
        assert a.shape[1] == b.shape[1]
        nn = self._empty((a.shape[0],), torch.int32)
        dist = self._empty((a.shape[0],), torch.float64)
        fb = C.c_int(0)
        _lib.check(self._h, _lib.lib().b2r_nn_gemm(self._h, a.data_ptr(), a.shape[0], b.data_ptr(), b.shape[0], a.shape[1],
                                                   nn.data_ptr(), dist.data_ptr(), C.byref(fb)))
        self.last_fallback_rows = fb.value
        return nn, dist

    def nn_gemm_error(self, fq, fr):
        """Test hook (b2r_debug.h): nn_gemm with the error of the tensor-core distances measured on every kept candidate pair.
        -> (nn, dict(ratio, bound, fallback_rows))."""
        torch = _torch()
        a, b = self._pts(fq), self._pts(fr)
        nn = self._empty((a.shape[0],), torch.int32)
        out = (C.c_double * 3)()
        _lib.check(self._h, _lib.lib().b2r_debug_nn_gemm_error(self._h, a.data_ptr(), a.shape[0], b.data_ptr(), b.shape[0], a.shape[1],
                                                               nn.data_ptr(), out))
        return nn, dict(ratio=float(out[0]), bound=float(out[1]), fallback_rows=int(out[2]))

    def mutual_nn(self, fa, fb, method="gemm"):
        """Mutual nearest neighbours of two descriptor sets (one descriptor per row).
        -> (pairs [c,2] int32 ascending i, nn01 [na], nn10 [nb]).  method: "gemm" (tensor cores + exact re-rank) or
        "exact" (fp64 brute force); both return the oracle's indices."""
        torch = _torch()
        a, b = self._pts(fa), self._pts(fb)
        assert a.shape[1] == b.shape[1]
        if method == "exact":
            nn01, _ = self.nn_exact(a, b)
            nn10, _ = self.nn_exact(b, a)
            return self.mutual_pairs(nn01, nn10), nn01, nn10
        na, nb = a.shape[0], b.shape[0]
        nn01 = self._empty((max(na, 1),), torch.int32)
        nn10 = self._empty((max(nb, 1),), torch.int32)
        pairs = self._empty((max(na, 1), 2), torch.int32)
        c, fbk = C.c_int(0), C.c_int(0)
        _lib.check(self._h, _lib.lib().b2r_mutual_nn(self._h, a.data_ptr(), na, b.data_ptr(), nb, a.shape[1], nn01.data_ptr(),
                                                     nn10.data_ptr(), pairs.data_ptr(), C.byref(c), C.byref(fbk)))
        self.last_fallback_rows = fbk.value
        return pairs[: c.value], nn01[:na], nn10[:nb]

    def gather_correspondences(self, a_pts, b_pts, pairs):
        """-> (src, dst) packed [N,3] CUDA tensors (the 3xN column-major operands of solve)."""
        torch = _torch()
        a, b = self._pts(a_pts), self._pts(b_pts)
        n = pairs.shape[0]
        src = self._empty((max(n, 1), 3), torch.float64)
        dst = self._empty((max(n, 1), 3), torch.float64)
        pairs = pairs.contiguous()
        _lib.check(self._h, _lib.lib().b2r_gather_correspondences(self._h, a.data_ptr(), b.data_ptr(), pairs.data_ptr(), n,
                                                                  src.data_ptr(), dst.data_ptr()))
        return src[:n], dst[:n]
