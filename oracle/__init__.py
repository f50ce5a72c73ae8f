"""CPU oracle -- TEST INFRASTRUCTURE ONLY (see oracle/teaser_oracle.cpp header).

ctypes front-end for ``oracle/_build/liboracle_teaser.so`` (TEASER++ restatement,
SURVEY.md Appendix A) and ``liboracle_presteps.so`` (Open3D pre-steps, Appendix B).

PARITY UNPINNED: the reference (patricksko/PoseEstimator) holds no tests or
golden vectors for this path and its native dependencies (TEASER++, Open3D)
are absent from this machine, so the oracle is a restatement validated by
ground-truth recovery and independent cross-checks (tests/test_oracle_*.py).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs import this package.  The product
(``poseestimator_b200``) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_BUILD = os.path.join(_HERE, "_build")


def build(force: bool = False) -> None:
    """Compile the oracle with the committed Makefile (g++ only, seconds)."""
    libs = [os.path.join(_BUILD, n) for n in ("liboracle_teaser.so", "liboracle_presteps.so", "liboracle_refine.so", "liboracle_raster.so")]
    srcs = [os.path.join(_HERE, n) for n in ("teaser_oracle.cpp", "presteps_oracle.cpp", "refine_oracle.cpp", "raster_oracle.cpp")]
    stale = force or any(
        (not os.path.exists(l)) or os.path.getmtime(l) < os.path.getmtime(s) for l, s in zip(libs, srcs) if os.path.exists(s)
    )
    if stale:
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))


class OrcParams(C.Structure):
    # mirrors `struct Params` in teaser_oracle.cpp (SURVEY A.1 defaults)
    _fields_ = [
        ("noise_bound", C.c_double),
        ("cbar2", C.c_double),
        ("estimate_scaling", C.c_int),
        ("rotation_estimation_algorithm", C.c_int),
        ("rotation_gnc_factor", C.c_double),
        ("rotation_max_iterations", C.c_int),
        ("rotation_cost_threshold", C.c_double),
        ("rotation_tim_graph", C.c_int),
        ("inlier_selection_mode", C.c_int),
        ("kcore_heuristic_threshold", C.c_double),
        ("max_clique_time_limit", C.c_double),
        ("max_clique_num_threads", C.c_int),
        ("rotation_noise_bound", C.c_double),
        ("materialize_tims", C.c_int),
    ]


def make_params(**kw) -> OrcParams:
    p = OrcParams(
        noise_bound=0.01, cbar2=1.0, estimate_scaling=1, rotation_estimation_algorithm=0,
        rotation_gnc_factor=1.4, rotation_max_iterations=100, rotation_cost_threshold=1e-6,
        rotation_tim_graph=0, inlier_selection_mode=0, kcore_heuristic_threshold=0.5,
        max_clique_time_limit=3000.0, max_clique_num_threads=0, rotation_noise_bound=0.0,
        materialize_tims=1,
    )
    for k, v in kw.items():
        if not hasattr(p, k):
            raise AttributeError(k)
        setattr(p, k, v)
    return p


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(os.path.join(_BUILD, "liboracle_teaser.so"))
        L.orc_solve.restype = C.c_void_p
        L.orc_solve.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(OrcParams)]
        L.orc_free.argtypes = [C.c_void_p]
        for name in ("orc_sizes", "orc_timing", "orc_clique", "orc_rot_mask", "orc_trans_mask", "orc_scale_mask",
                     "orc_gnc_weights", "orc_cores"):
            getattr(L, name).argtypes = [C.c_void_p, C.c_void_p]
        L.orc_solution.argtypes = [C.c_void_p] + [C.c_void_p] * 5
        L.orc_adjacency.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.orc_rot_tims.argtypes = [C.c_void_p] * 4
        L.orc_svd3.argtypes = [C.c_void_p] * 4
        L.orc_tls_estimate.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_gnc_tls.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_int, C.c_double,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_max_clique_bits.restype = C.c_int
        L.orc_max_clique_bits.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_void_p]
        L.orc_tim_graph.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_void_p, C.c_int, C.c_int]
        L.orc_num_threads.restype = C.c_int
        _lib = L
    return _lib


def _colmajor(a: np.ndarray) -> np.ndarray:
    """3xN array -> contiguous buffer of N (x,y,z) triples (column-major 3xN)."""
    a = np.asarray(a, dtype=np.float64)
    assert a.ndim == 2 and a.shape[0] == 3, a.shape
    return np.ascontiguousarray(a.T)  # (N,3) C-order == 3xN column-major


def words32(n: int) -> int:
    return (n + 31) // 32


@dataclass
class OracleResult:
    n: int
    n_edges: int
    scale: float
    rotation: np.ndarray
    translation: np.ndarray
    valid: bool
    clique: np.ndarray
    rot_mask: np.ndarray
    trans_mask: np.ndarray
    adjacency: np.ndarray | None
    gnc_iters: int
    gnc_cost: float
    heuristic_lb: int
    max_core: int
    timing: dict
    scale_mask: np.ndarray | None = None
    gnc_weights: np.ndarray | None = None

    @property
    def rotation_inliers(self):  # A.3: indices into the clique / pruned TIM list
        return np.flatnonzero(self.rot_mask).astype(np.int32)

    @property
    def translation_inliers(self):  # A.3: indices into the rotation-inlier list
        return np.flatnonzero(self.trans_mask).astype(np.int32)


def solve(src, dst, params: OrcParams | None = None, want_adjacency=True, want_scale_mask=False, **kw) -> OracleResult:
    """Full TEASER++ solve on the CPU (A.8)."""
    L = lib()
    p = params if params is not None else make_params(**kw)
    s, d = _colmajor(src), _colmajor(dst)
    n = s.shape[0]
    assert d.shape[0] == n
    h = L.orc_solve(s.ctypes.data, d.ctypes.data, n, C.byref(p))
    try:
        sizes = np.zeros(16, np.int64)
        L.orc_sizes(h, sizes.ctypes.data)
        scale = C.c_double(); valid = C.c_int(); cost = C.c_double()
        R = np.zeros(9); t = np.zeros(3)
        L.orc_solution(h, C.addressof(scale), R.ctypes.data, t.ctypes.data, C.addressof(valid), C.addressof(cost))
        tim = np.zeros(8)
        L.orc_timing(h, tim.ctypes.data)
        clique = np.zeros(int(sizes[3]), np.int32)
        if clique.size:
            L.orc_clique(h, clique.ctypes.data)
        rot_mask = np.zeros(int(sizes[4]), np.uint8)
        if rot_mask.size:
            L.orc_rot_mask(h, rot_mask.ctypes.data)
        trans_mask = np.zeros(int(sizes[6]), np.uint8)
        if trans_mask.size:
            L.orc_trans_mask(h, trans_mask.ctypes.data)
        adj = None
        if want_adjacency:
            W = words32(n)
            adj = np.zeros((n, W), np.uint32)
            L.orc_adjacency(h, adj.ctypes.data, W)
        smask = None
        if want_scale_mask and p.materialize_tims:
            smask = np.zeros(int(sizes[1]), np.uint8)
            L.orc_scale_mask(h, smask.ctypes.data)
        w = np.zeros(int(sizes[4]))
        if w.size:
            L.orc_gnc_weights(h, w.ctypes.data)
        return OracleResult(
            n=n, n_edges=int(sizes[2]), scale=scale.value, rotation=R.reshape(3, 3).copy(), translation=t,
            valid=bool(valid.value), clique=clique, rot_mask=rot_mask.astype(bool), trans_mask=trans_mask.astype(bool),
            adjacency=adj, gnc_iters=int(sizes[7]), gnc_cost=cost.value, heuristic_lb=int(sizes[8]),
            max_core=int(sizes[9]),
            timing=dict(zip(("tims", "scale", "graph", "clique", "rotation", "translation", "total"), tim[:7].tolist())),
            scale_mask=smask, gnc_weights=w,
        )
    finally:
        L.orc_free(h)


def tim_graph(src, dst, noise_bound, cbar2=1.0, variant=0) -> np.ndarray:
    """Bit-packed consistency graph only (A.4), u32 [N, ceil(N/32)]."""
    L = lib()
    s, d = _colmajor(src), _colmajor(dst)
    n = s.shape[0]
    W = words32(n)
    out = np.zeros((n, W), np.uint32)
    L.orc_tim_graph(s.ctypes.data, d.ctypes.data, n, float(noise_bound), float(cbar2), out.ctypes.data, W, variant)
    return out


def max_clique_bits(adj: np.ndarray, mode=0, kcore_thr=0.5) -> np.ndarray:
    L = lib()
    adj = np.ascontiguousarray(adj, np.uint32)
    n, W = adj.shape
    out = np.zeros(n, np.int32)
    c = L.orc_max_clique_bits(adj.ctypes.data, n, W, mode, kcore_thr, out.ctypes.data)
    return out[:c].copy()


def svd3(A):
    L = lib()
    A = np.ascontiguousarray(A, np.float64)
    U = np.zeros((3, 3)); S = np.zeros(3); V = np.zeros((3, 3))
    L.orc_svd3(A.ctypes.data, U.ctypes.data, S.ctypes.data, V.ctypes.data)
    return U, S, V


def tls_estimate(X, ranges):
    L = lib()
    X = np.ascontiguousarray(X, np.float64); r = np.ascontiguousarray(ranges, np.float64)
    est = C.c_double(); inl = np.zeros(X.size, np.uint8)
    L.orc_tls_estimate(X.ctypes.data, r.ctypes.data, X.size, C.addressof(est), inl.ctypes.data)
    return est.value, inl.astype(bool)


def gnc_tls(X, Y, noise_bound, factor=1.4, max_iter=100, thr=1e-6):
    L = lib()
    x, y = _colmajor(X), _colmajor(Y)
    n = x.shape[0]
    R = np.zeros(9); inl = np.zeros(n, np.uint8); w = np.zeros(n); it = C.c_int(); cost = C.c_double()
    L.orc_gnc_tls(x.ctypes.data, y.ctypes.data, n, noise_bound, factor, max_iter, thr, R.ctypes.data,
                  inl.ctypes.data, w.ctypes.data, C.addressof(it), C.addressof(cost))
    return R.reshape(3, 3), inl.astype(bool), w, it.value, cost.value


def num_threads() -> int:
    return lib().orc_num_threads()


def set_num_threads(n: int) -> int:
    """OpenMP threads of the oracle (torchrun sets OMP_NUM_THREADS=1 for its ranks); returns the count now in force."""
    L = lib()
    L.orc_set_num_threads.restype = None
    L.orc_set_num_threads.argtypes = [C.c_int]
    L.orc_set_num_threads(int(n))
    return L.orc_num_threads()
