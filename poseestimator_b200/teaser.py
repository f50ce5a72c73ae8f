"""`teaserpp_python`-compatible front-end over the B200 registration library.

Mirrors the pybind11 surface the reference uses:
  /root/reference/src/pose_estimator/EstimHelpers/HelpersRealtime.py:158-171
  /root/reference/src/pose_estimator/EstimHelpers/registration_utils.py:251-266
Same names, argument meaning and error behaviour; the arithmetic runs in hand-written sm_100a kernels behind the
C ABI in include/b2r.h.  There is no CPU fallback: constructing a solver without a CUDA device raises.
"""
from __future__ import annotations

import ctypes as C
import enum

import numpy as np

from . import _lib


class RotationEstimationAlgorithm(enum.IntEnum):
    GNC_TLS = 0
    FGR = 1


class InlierSelectionMode(enum.IntEnum):
    PMC_EXACT = 0
    PMC_HEU = 1
    KCORE_HEU = 2
    NONE = 3


class InlierGraphFormulation(enum.IntEnum):
    CHAIN = 0
    COMPLETE = 1


class RegistrationSolution:
    """teaser::RegistrationSolution: scale, 3x3 rotation, (3,) translation, valid."""

    def __init__(self, scale=1.0, rotation=None, translation=None, valid=False):
        self.scale = float(scale)
        self.rotation = np.eye(3) if rotation is None else np.asarray(rotation, dtype=np.float64).reshape(3, 3)
        self.translation = np.zeros(3) if translation is None else np.asarray(translation, dtype=np.float64).reshape(3)
        self.valid = bool(valid)

    def __repr__(self):
        return f"<RegistrationSolution scale={self.scale} valid={self.valid}\nrotation=\n{self.rotation}\ntranslation={self.translation}>"


class _Params:
    """RobustRegistrationSolver.Params -- read/write attributes with upstream's defaults."""

    __slots__ = ("noise_bound", "cbar2", "estimate_scaling", "rotation_estimation_algorithm", "rotation_gnc_factor",
                 "rotation_max_iterations", "rotation_cost_threshold", "rotation_tim_graph", "inlier_selection_mode",
                 "kcore_heuristic_threshold", "use_max_clique", "max_clique_exact_solution", "max_clique_time_limit",
                 "max_clique_num_threads", "graph_exact_only", "graph_filter_tau")

    def __init__(self):
        self.noise_bound = 0.01
        self.cbar2 = 1.0
        self.estimate_scaling = True
        self.rotation_estimation_algorithm = RotationEstimationAlgorithm.GNC_TLS
        self.rotation_gnc_factor = 1.4
        self.rotation_max_iterations = 100
        self.rotation_cost_threshold = 1e-6
        self.rotation_tim_graph = InlierGraphFormulation.CHAIN
        self.inlier_selection_mode = InlierSelectionMode.PMC_EXACT
        self.kcore_heuristic_threshold = 0.5
        self.use_max_clique = True             # deprecated upstream: False -> NONE
        self.max_clique_exact_solution = True  # deprecated upstream: False -> PMC_HEU
        self.max_clique_time_limit = 3000.0
        self.max_clique_num_threads = 0
        # test hooks of this implementation (not in upstream, not in b2r_params: applied through include/b2r_debug.h)
        self.graph_exact_only = False
        self.graph_filter_tau = 0.0

    def _copy(self):
        p = _Params()
        for k in self.__slots__:
            setattr(p, k, getattr(self, k))
        return p

    def _to_c(self, rotation_noise_bound=0.0) -> _lib.B2RParams:
        mode = int(self.inlier_selection_mode)
        if not self.use_max_clique:
            mode = int(InlierSelectionMode.NONE)
        elif not self.max_clique_exact_solution and mode == int(InlierSelectionMode.PMC_EXACT):
            mode = int(InlierSelectionMode.PMC_HEU)
        return _lib.B2RParams(
            noise_bound=float(self.noise_bound), cbar2=float(self.cbar2), estimate_scaling=int(bool(self.estimate_scaling)),
            rotation_estimation_algorithm=int(self.rotation_estimation_algorithm),
            rotation_gnc_factor=float(self.rotation_gnc_factor), rotation_max_iterations=int(self.rotation_max_iterations),
            rotation_cost_threshold=float(self.rotation_cost_threshold), rotation_tim_graph=int(self.rotation_tim_graph),
            inlier_selection_mode=mode, kcore_heuristic_threshold=float(self.kcore_heuristic_threshold),
            max_clique_time_limit=float(self.max_clique_time_limit), max_clique_num_threads=int(self.max_clique_num_threads),
            rotation_noise_bound=float(rotation_noise_bound),
        )

    def _apply_debug(self, handle) -> None:
        """graph_exact_only / graph_filter_tau are test hooks: b2r_debug_set_graph_filter (include/b2r_debug.h)."""
        if int(self.graph_exact_only) or float(self.graph_filter_tau) > 0:
            _lib.check(handle, _lib.lib().b2r_debug_set_graph_filter(handle, int(self.graph_exact_only), float(self.graph_filter_tau)))

    def __repr__(self):
        return "<Params " + " ".join(f"{k}={getattr(self, k)!r}" for k in self.__slots__) + ">"


def _is_torch_cuda(x) -> bool:
    return type(x).__module__.startswith("torch") and hasattr(x, "is_cuda") and x.is_cuda


def _as_3xn(a, name):
    """pybind11/Eigen semantics: anything convertible to a float64 3xN matrix, else TypeError."""
    try:
        arr = np.asarray(a)
        if arr.dtype != np.float64:
            arr = arr.astype(np.float64)
    except Exception as e:  # noqa: BLE001
        raise TypeError(f"solve(): {name} is not convertible to a float64 matrix") from e
    if arr.ndim != 2 or arr.shape[0] != 3:
        raise TypeError(f"solve(): incompatible function arguments. {name} must be a float64 array of shape (3, N), got {arr.shape}")
    return arr


def _tim_map(n: int) -> np.ndarray:
    i, j = np.triu_indices(n, 1)  # row-major upper triangle == TIM order (A.4-1)
    return np.stack([i, j]).astype(np.int32)


class RobustRegistrationSolver:
    """Drop-in for teaserpp_python.RobustRegistrationSolver."""

    Params = _Params
    ROTATION_ESTIMATION_ALGORITHM = RotationEstimationAlgorithm
    INLIER_SELECTION_MODE = InlierSelectionMode
    INLIER_GRAPH_FORMULATION = InlierGraphFormulation

    def __init__(self, params: "_Params | None" = None, device: "int | None" = None, stream=None):
        L = _lib.lib()
        if L.b2r_device_count() <= 0:
            raise RuntimeError("poseestimator_b200: no CUDA device is visible and there is no CPU fallback")
        if device is None:
            device = 0
            try:
                import torch
                if torch.cuda.is_available():
                    device = torch.cuda.current_device()
            except Exception:  # noqa: BLE001
                pass
        self._device = int(device)
        self._params = (params or _Params())._copy()  # value semantics: later edits of `params` have no effect
        self._rot_nb = float(self._params.noise_bound)  # A.6-2: compounds across solve() calls until reset()
        self._h = C.c_void_p()
        cp = self._params._to_c(self._rot_nb)
        st = L.b2r_create(C.byref(cp), self._device, C.c_void_p(int(stream) if stream else None), C.byref(self._h))
        if st != 0:
            raise RuntimeError(f"poseestimator_b200: b2r_create failed ({_lib.STATUS_NAMES.get(st, st)})")
        self._params._apply_debug(self._h)
        self._src = self._dst = None
        self._n = 0
        self._solved = False
        self._solution = RegistrationSolution()
        self._pending_params = None
        self._cache = {}

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                _lib.lib().b2r_destroy(h)
            except Exception:  # noqa: BLE001
                pass
            self._h = None

    # ---- configuration ------------------------------------------------------------------------------------
    def getParams(self):
        return self._params._copy()

    def reset(self, params: "_Params"):
        self._params = params._copy()
        self._rot_nb = float(self._params.noise_bound)
        cp = self._params._to_c(self._rot_nb)
        _lib.check(self._h, _lib.lib().b2r_reset(self._h, C.byref(cp)))
        _lib.check(self._h, _lib.lib().b2r_debug_set_graph_filter(self._h, int(self._params.graph_exact_only), float(self._params.graph_filter_tau)))
        self._solved = False
        self._solution = RegistrationSolution()
        self._pending_params = None

    # ---- solve --------------------------------------------------------------------------------------------
    def solve(self, src, dst):
        """solve(src, dst): 3xN float64 (any strides).  Also accepts float64 CUDA tensors (zero copy)."""
        L = _lib.lib()
        if self._pending_params is not None:  # compounded rotation noise bound from the previous solve (A.6-2)
            _lib.check(self._h, L.b2r_reset(self._h, C.byref(self._pending_params)))
            self._pending_params = None
        if _is_torch_cuda(src) or _is_torch_cuda(dst):
            if not (_is_torch_cuda(src) and _is_torch_cuda(dst)):
                raise TypeError("solve(): src and dst must live on the same device")
            import torch
            if (src.dtype != torch.float64 or dst.dtype != torch.float64 or src.dim() != 2 or src.shape[0] != 3
                    or src.shape != dst.shape):
                raise TypeError("solve(): CUDA tensors must be float64 of identical shape (3, N)")
            if src.device.index != self._device or dst.device.index != self._device:
                raise TypeError(f"solve(): tensors live on {src.device} / {dst.device} but this solver is bound to cuda:{self._device}")
            if src.stride() != dst.stride():  # bring both to the packed column-major layout
                src, dst = src.T.contiguous().T, dst.T.contiguous().T
            torch.cuda.current_stream(src.device).synchronize()  # the library runs on its own stream
            n = int(src.shape[1])
            rs, cs = src.stride()
            self._src, self._dst = src, dst
            st = L.b2r_solve(self._h, C.c_void_p(src.data_ptr()), C.c_void_p(dst.data_ptr()), n, rs, cs, 1)
        else:
            s = _as_3xn(src, "src")
            d = _as_3xn(dst, "dst")
            if s.shape != d.shape:
                raise TypeError(f"solve(): src {s.shape} and dst {d.shape} must have the same shape")
            n = int(s.shape[1])
            if d.strides != s.strides:
                s, d = np.asfortranarray(s), np.asfortranarray(d)
            self._src, self._dst = s, d
            rs, cs = s.strides[0] // 8, s.strides[1] // 8
            st = L.b2r_solve(self._h, C.c_void_p(s.ctypes.data), C.c_void_p(d.ctypes.data), n, rs, cs, 0)
        _lib.check(self._h, st)
        self._n = n
        self._solved = True
        self._cache = {}
        sol = _lib.B2RSolution()
        _lib.check(self._h, L.b2r_get_solution(self._h, 0, C.byref(sol)))
        self._solution = RegistrationSolution(sol.scale, np.array(sol.rotation[:]).reshape(3, 3),
                                              np.array(sol.translation[:]), sol.valid)
        # upstream multiplies the rotation solver's stored noise bound in place on every solve (A.6-2);
        # mirrored here so a re-used solver object behaves like upstream's.  reset() restores it.
        self._rot_nb *= 2.0 / self._solution.scale
        self._pending_params = self._params._to_c(self._rot_nb)
        return self._solution

    def getSolution(self):
        return self._solution

    # ---- results ------------------------------------------------------------------------------------------
    def _require(self):
        if not self._solved:
            raise RuntimeError("solve() has not been called")

    def _indices(self, kind) -> list:
        if not self._solved:
            return []
        key = ("idx", kind)
        if key not in self._cache:
            L = _lib.lib()
            cnt = C.c_int32()
            _lib.check(self._h, L.b2r_get_indices(self._h, 0, kind, None, C.byref(cnt)))
            out = np.zeros(max(cnt.value, 1), np.int32)
            _lib.check(self._h, L.b2r_get_indices(self._h, 0, kind, C.c_void_p(out.ctypes.data), C.byref(cnt)))
            self._cache[key] = out[: cnt.value].tolist()
        return list(self._cache[key])

    def stats(self) -> dict:
        """Per-solve statistics (extension)."""
        self._require()
        st = _lib.B2RStats()
        _lib.check(self._h, _lib.lib().b2r_get_stats(self._h, 0, C.byref(st)))
        return st.as_dict()

    def getInlierMaxClique(self):
        return self._indices(_lib.IDX_MAX_CLIQUE)

    def getRotationInliers(self):
        return self._indices(_lib.IDX_ROT_INLIERS)

    def getRotationInliersMap(self):
        return np.asarray(self._indices(_lib.IDX_ROT_INLIERS_MAP), np.int32).reshape(1, -1)

    def getTranslationInliers(self):
        return self._indices(_lib.IDX_TRANS_INLIERS)

    def getTranslationInliersMap(self):
        return np.asarray(self._indices(_lib.IDX_TRANS_INLIERS_MAP), np.int32).reshape(1, -1)

    def _mask(self, idx, length):
        m = np.zeros((1, length), bool)
        m[0, idx] = True
        return m

    def getRotationInliersMask(self):
        self._require()
        return self._mask(self._indices(_lib.IDX_ROT_INLIERS), self.stats()["n_rot_tims"])

    def getTranslationInliersMask(self):
        self._require()
        return self._mask(self._indices(_lib.IDX_TRANS_INLIERS), self.stats()["n_rot_inliers"])

    def getGNCRotationCostAtTermination(self):
        self._require()
        return float(self.stats()["gnc_cost"])

    def _adjacency_bits(self) -> np.ndarray:
        self._require()
        if "adj" not in self._cache:
            W = (self._n + 31) // 32
            out = np.zeros((self._n, W), np.uint32)
            _lib.check(self._h, _lib.lib().b2r_get_adjacency(self._h, 0, C.c_void_p(out.ctypes.data), W))
            self._cache["adj"] = out
        return self._cache["adj"]

    def _tc_values(self, row_block: int, unit: int):
        """Test hook: raw Dm~ / L~ (128 x 64 fp32 each) of one unit of the tensor-core graph filter and its constants."""
        self._require()
        dm = np.zeros((128, 64), np.float32)
        lv = np.zeros((128, 64), np.float32)
        c = np.zeros(12, np.float64)
        _lib.check(self._h, _lib.lib().b2r_debug_tc_values(self._h, 0, row_block, unit, C.c_void_p(dm.ctypes.data), C.c_void_p(lv.ctypes.data),
                                                            C.c_void_p(c.ctypes.data)))
        return dm, lv, dict(kappa=c[0], ctr=c[1:7].copy(), t1=c[7], t0=c[8], lcut=c[9], k=c[10], use32=int(c[11]))

    def _tc_sweep(self) -> dict:
        """Test hook (b2r_debug.h): EVERY unit of the tensor-core graph filter recomputed and compared with fp64 on the device."""
        self._require()
        out = (C.c_double * 8)()
        _lib.check(self._h, _lib.lib().b2r_debug_tc_sweep(self._h, 0, out))
        names = ("tc_accum_ratio", "dm_over_ED", "l_over_EL", "pairs", "Pi", "E_D", "E_L", "use32")
        return dict(zip(names, [float(x) for x in out]))

    def _adjacency_dense(self) -> np.ndarray:
        bits = self._adjacency_bits()
        return np.unpackbits(bits.view(np.uint8), axis=1, bitorder="little")[:, : self._n].astype(bool)

    def getInlierGraph(self):
        A = self._adjacency_dense()
        return [np.flatnonzero(r).tolist() for r in A]

    def getScaleInliersMask(self):
        A = self._adjacency_dense()
        i, j = np.triu_indices(self._n, 1)
        return A[i, j].reshape(1, -1)

    def getScaleInliersMap(self):
        self._require()
        return _tim_map(self._n)

    def getScaleInliers(self):
        """One (i, j) tuple per TIM that passed the scale test, in TIM order == the edge list (A.3). Lazy."""
        if not self._solved:
            return []
        A = self._adjacency_dense()
        i, j = np.nonzero(np.triu(A, 1))
        return list(zip(i.tolist(), j.tolist()))

    def getSrcTIMsMap(self):
        self._require()
        return _tim_map(self._n)

    getDstTIMsMap = getSrcTIMsMap

    def _host_points(self):
        s, d = self._src, self._dst
        if _is_torch_cuda(s):
            s, d = s.cpu().numpy(), d.cpu().numpy()
        return np.asarray(s), np.asarray(d)

    def getSrcTIMs(self):
        self._require()
        s, _ = self._host_points()
        m = _tim_map(self._n)
        return s[:, m[1]] - s[:, m[0]]

    def getDstTIMs(self):
        self._require()
        _, d = self._host_points()
        m = _tim_map(self._n)
        return d[:, m[1]] - d[:, m[0]]

    def getSrcTIMsMapForRotation(self):
        self._require()
        c = np.asarray(self.getInlierMaxClique(), np.int32)
        if c.size == 0:
            return np.zeros((2, 0), np.int32)
        if int(self._params.rotation_tim_graph) == int(InlierGraphFormulation.CHAIN):
            return np.stack([c, np.roll(c, -1)])
        i, j = np.triu_indices(c.size, 1)
        return np.stack([c[i], c[j]])

    getDstTIMsMapForRotation = getSrcTIMsMapForRotation

    def getMaxCliqueSrcTIMs(self):
        s, _ = self._host_points()
        m = self.getSrcTIMsMapForRotation()
        return s[:, m[1]] - s[:, m[0]]

    def getMaxCliqueDstTIMs(self):
        _, d = self._host_points()
        m = self.getSrcTIMsMapForRotation()
        return (d[:, m[1]] - d[:, m[0]]) * (1.0 / self._solution.scale)
