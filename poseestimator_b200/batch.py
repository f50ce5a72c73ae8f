"""Batched registration of independent detections / templates / frames (BASELINE.json config 4).

The reference loops over templates one `run_teaser` call at a time
(/root/reference/src/pose_estimator/EstimHelpers/PoseEstimator.py:90-114, registration_utils.py:359-408).
Here the whole batch goes through one set of launches (grid.y = problem) on one stream, with one H2D copy in and
one small D2H copy out.  Independent units shard across GPUs with no collective (`shard_round_robin`).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .teaser import RegistrationSolution, _Params


from .sharded import shard_round_robin  # noqa: E402,F401  (kept here for callers of the batch module)


def _pack(a) -> np.ndarray:
    """3xN (any strides) -> packed column-major buffer (N,3) C-order."""
    a = np.asarray(a, dtype=np.float64)
    if a.ndim != 2 or a.shape[0] != 3:
        raise TypeError(f"expected a (3, N) float64 array, got {a.shape}")
    return np.ascontiguousarray(a.T)


class PinnedBuffer:
    """Page-locked host memory from b2r_alloc_pinned, viewed as numpy arrays (for BatchSolver.solve_pinned_async)."""

    def __init__(self, nbytes: int):
        self._p = C.c_void_p()
        st = _lib.lib().b2r_alloc_pinned(int(nbytes), C.byref(self._p))
        if st != 0 or not self._p.value:
            raise MemoryError(f"poseestimator_b200: b2r_alloc_pinned({nbytes}) failed ({_lib.STATUS_NAMES.get(st, st)})")
        self.nbytes = int(nbytes)
        self._raw = (C.c_char * self.nbytes).from_address(self._p.value)
        self._raw._owner = self  # views reach the block through their .base chain; the memory lives as long as any view

    def array(self, shape, dtype=np.float64, offset: int = 0) -> np.ndarray:
        """A numpy view of the block (keeps the block alive through its .base chain)."""
        count = int(np.prod(shape))
        a = np.frombuffer(self._raw, dtype=dtype, count=count, offset=offset).reshape(shape)
        return a

    @staticmethod
    def fill(dst: np.ndarray, src: np.ndarray) -> None:
        """dst[...] = src with streaming stores (b2r_copy_to_pinned): the lines end up in DRAM, where the copy engine reads
        them at full rate, instead of dirty in the CPU caches."""
        src = np.ascontiguousarray(src, dtype=dst.dtype)
        if not dst.flags.c_contiguous or dst.nbytes != src.nbytes:
            raise ValueError("fill needs a C-contiguous destination of the source's size")
        st = _lib.lib().b2r_copy_to_pinned(C.c_void_p(dst.ctypes.data), C.c_void_p(src.ctypes.data), src.nbytes)
        if st != 0:
            raise RuntimeError("b2r_copy_to_pinned failed")

    def __del__(self):
        p = getattr(self, "_p", None)
        if p is not None and p.value:
            try:
                _lib.lib().b2r_free_pinned(p)
            except Exception:  # noqa: BLE001
                pass
            self._p = C.c_void_p()


def pinned_problem_block(batch: int, n: int):
    """(src, dst): two page-locked (batch, n, 3) float64 arrays, each contiguous, so that a whole batch uploads in two copies."""
    nb = batch * n * 3 * 8
    buf = PinnedBuffer(2 * nb)
    return buf.array((batch, n, 3)), buf.array((batch, n, 3), offset=nb)


class HostBatch:
    """Pointer tables (n, src, dst) of a batch of packed (N,3) float64 operands in host memory, built ONCE: a caller that rewrites
    the same arrays every frame (a page-locked block, typically) submits with BatchSolver.solve_host_batch_async and pays no
    per-step Python cost for the 2 x B pointers.  page_locked=True: the arrays live in page-locked memory (on_device = 2)."""

    def __init__(self, packed, page_locked: bool = False):
        B = len(packed)
        for a, b in packed:
            if a.dtype != np.float64 or b.dtype != np.float64 or not a.flags.c_contiguous or not b.flags.c_contiguous \
                    or a.ndim != 2 or a.shape[1] != 3 or a.shape != b.shape:
                raise TypeError("HostBatch needs C-contiguous (N, 3) float64 pairs of equal shape")
        self.batch = B
        self.n = (C.c_int * B)(*[p[0].shape[0] for p in packed])
        self.sp = (C.c_void_p * B)(*[p[0].ctypes.data for p in packed])
        self.dp = (C.c_void_p * B)(*[p[1].ctypes.data for p in packed])
        self.mode = 2 if page_locked else 0
        self._keep = packed


class BatchSolver:
    """Many registrations per call through b2r_solve_batch (one handle = one device + one stream)."""

    def __init__(self, params: "_Params | None" = None, device: int = 0, stream=None, profiling: bool = False):
        L = _lib.lib()
        if L.b2r_device_count() <= 0:
            raise RuntimeError("poseestimator_b200: no CUDA device is visible and there is no CPU fallback")
        self._params = (params or _Params())._copy()
        self._h = C.c_void_p()
        cp = self._params._to_c(0.0)
        st = L.b2r_create(C.byref(cp), int(device), C.c_void_p(int(stream) if stream else None), C.byref(self._h))
        if st != 0:
            raise RuntimeError(f"poseestimator_b200: b2r_create failed ({_lib.STATUS_NAMES.get(st, st)})")
        self._params._apply_debug(self._h)
        if profiling:
            L.b2r_set_profiling(self._h, 1)
        self.device = int(device)
        self._keep = None
        self._batch = 0

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                _lib.lib().b2r_destroy(h)
            except Exception:  # noqa: BLE001
                pass
            self._h = None

    # ---- host inputs ------------------------------------------------------------------------------------
    def solve_async(self, problems):
        """problems: sequence of (src, dst) 3xN arrays.  Enqueues H2D + all kernels + D2H; returns immediately."""
        packed = [(_pack(s), _pack(d)) for s, d in problems]
        return self.solve_packed_async(packed)

    def solve_packed_async(self, packed):
        """packed: sequence of (src, dst) already packed (N,3) C-order float64 arrays (no conversion cost)."""
        B = len(packed)
        n = (C.c_int * B)(*[p[0].shape[0] for p in packed])
        sp = (C.c_void_p * B)(*[p[0].ctypes.data for p in packed])
        dp = (C.c_void_p * B)(*[p[1].ctypes.data for p in packed])
        self._keep = (packed, n, sp, dp)
        self._batch = B
        _lib.check(self._h, _lib.lib().b2r_solve_batch_async(self._h, B, n, sp, dp, 0))

    def solve_pinned_async(self, packed):
        """Like solve_packed_async for arrays that live in page-locked memory (PinnedBuffer, torch pin_memory, ...): no
        staging pass on the host, the copy engine reads them directly; they must stay untouched until sync()."""
        B = len(packed)
        n = (C.c_int * B)(*[p[0].shape[0] for p in packed])
        sp = (C.c_void_p * B)(*[p[0].ctypes.data for p in packed])
        dp = (C.c_void_p * B)(*[p[1].ctypes.data for p in packed])
        self._keep = (packed, n, sp, dp)
        self._batch = B
        _lib.check(self._h, _lib.lib().b2r_solve_batch_async(self._h, B, n, sp, dp, 2))

    def solve_host_batch_async(self, hb: "HostBatch"):
        """Submit a prepared HostBatch (pointer tables built once; the arrays' current contents are what gets uploaded)."""
        self._keep = hb
        self._batch = hb.batch
        _lib.check(self._h, _lib.lib().b2r_solve_batch_async(self._h, hb.batch, hb.n, hb.sp, hb.dp, hb.mode))

    # ---- device inputs (zero copy) ----------------------------------------------------------------------
    def solve_device_async(self, n_list, src_ptrs, dst_ptrs, keep_alive=None):
        """Device pointers to packed 3xN column-major float64 operands (e.g. torch tensors' data_ptr()) ON THIS HANDLE'S DEVICE;
        pass the tensors as keep_alive and their device is checked (raw pointers cannot be)."""
        if keep_alive is not None:
            for t in (keep_alive if isinstance(keep_alive, (tuple, list)) else (keep_alive,)):
                dev = getattr(t, "device", None)
                if dev is not None and getattr(dev, "type", "") == "cuda" and dev.index is not None and dev.index != self.device:
                    raise TypeError(f"tensor lives on {dev} but this solver is bound to cuda:{self.device}")
        B = len(n_list)
        n = (C.c_int * B)(*[int(x) for x in n_list])
        sp = (C.c_void_p * B)(*[int(x) for x in src_ptrs])
        dp = (C.c_void_p * B)(*[int(x) for x in dst_ptrs])
        self._keep = (keep_alive, n, sp, dp)
        self._batch = B
        _lib.check(self._h, _lib.lib().b2r_solve_batch_async(self._h, B, n, sp, dp, 1))

    def sync(self):
        _lib.check(self._h, _lib.lib().b2r_sync(self._h))

    def solve(self, problems):
        self.solve_async(problems)
        self.sync()
        return self.solutions()

    # ---- results ----------------------------------------------------------------------------------------
    def solution(self, i: int) -> RegistrationSolution:
        sol = _lib.B2RSolution()
        _lib.check(self._h, _lib.lib().b2r_get_solution(self._h, i, C.byref(sol)))
        return RegistrationSolution(sol.scale, np.array(sol.rotation[:]).reshape(3, 3), np.array(sol.translation[:]), sol.valid)

    def solutions_array(self) -> np.ndarray:
        """All solutions of the last batch in ONE C call: structured array (scale, rotation [3,3], translation [3], valid)."""
        out = np.zeros(max(self._batch, 1), _lib.SOLUTION_DTYPE)
        _lib.check(self._h, _lib.lib().b2r_get_solutions(self._h, C.c_void_p(out.ctypes.data), self._batch))
        return out[: self._batch]

    def solutions(self):
        a = self.solutions_array()
        return [RegistrationSolution(float(r["scale"]), r["rotation"], r["translation"], bool(r["valid"])) for r in a]

    def stats(self, i: int) -> dict:
        st = _lib.B2RStats()
        _lib.check(self._h, _lib.lib().b2r_get_stats(self._h, i, C.byref(st)))
        return st.as_dict()

    def indices(self, i: int, kind: int) -> np.ndarray:
        L = _lib.lib()
        cnt = C.c_int32()
        _lib.check(self._h, L.b2r_get_indices(self._h, i, kind, None, C.byref(cnt)))
        out = np.zeros(max(cnt.value, 1), np.int32)
        _lib.check(self._h, L.b2r_get_indices(self._h, i, kind, C.c_void_p(out.ctypes.data), C.byref(cnt)))
        return out[: cnt.value].copy()

    def max_clique(self, i: int) -> np.ndarray:
        return self.indices(i, _lib.IDX_MAX_CLIQUE)

    def adjacency(self, i: int, n: int) -> np.ndarray:
        W = (n + 31) // 32
        out = np.zeros((n, W), np.uint32)
        _lib.check(self._h, _lib.lib().b2r_get_adjacency(self._h, i, C.c_void_p(out.ctypes.data), W))
        return out

    def gnc_weights(self, i: int) -> np.ndarray:
        T = self.stats(i)["n_rot_tims"]
        out = np.zeros(max(T, 1))
        _lib.check(self._h, _lib.lib().b2r_get_gnc_weights(self._h, i, C.c_void_p(out.ctypes.data)))
        return out[:T]

    def timings_ms(self) -> dict:
        t = (C.c_float * 12)()
        _lib.check(self._h, _lib.lib().b2r_get_timings(self._h, t))
        names = ("upload", "graph", "degree", "heuristic", "prune_compact", "exact_search", "canonical", "rotation",
                 "translation", "total", "scale_vote", "tail_fused")
        return dict(zip(names, [float(x) for x in t]))

    def launch_count(self) -> int:
        return int(_lib.lib().b2r_launch_count(self._h))

    def tail_cycles(self, i: int) -> dict:
        """Test hook (b2r_debug.h): SM cycles of the fused tail kernel at the end of each of its stages, problem i."""
        out = (C.c_uint32 * 32)()
        _lib.check(self._h, _lib.lib().b2r_debug_tail_cycles(self._h, int(i), out))
        names = ("fix", "degrees", "heuristic", "prune", "clique", "rotation", "translation", "start")
        r = dict(zip(names, [int(x) for x in out]))
        r["marks"] = {k: int(out[k]) for k in range(8, 32) if out[k]}
        return r

    def tc_sweep(self, i: int) -> dict:
        """Test hook (b2r_debug.h): every unit of problem i recomputed on the tensor cores and compared with fp64 on the device."""
        out = (C.c_double * 8)()
        _lib.check(self._h, _lib.lib().b2r_debug_tc_sweep(self._h, int(i), out))
        names = ("tc_accum_ratio", "dm_over_ED", "l_over_EL", "pairs", "Pi", "E_D", "E_L", "use32")
        return dict(zip(names, [float(x) for x in out]))

    # ---- row-sharded large problem (config 5) -----------------------------------------------------------
    def graph_rows(self, src_ptr, dst_ptr, n, row_begin, row_end, out_ptr, pitch_words):
        _lib.check(self._h, _lib.lib().b2r_graph_rows(self._h, C.c_void_p(src_ptr), C.c_void_p(dst_ptr), n, row_begin,
                                                      row_end, C.c_void_p(out_ptr), pitch_words))

    def solve_from_adjacency(self, src_ptr, dst_ptr, n, adj_ptr, pitch_words):
        self._batch = 1
        _lib.check(self._h, _lib.lib().b2r_solve_from_adjacency(self._h, C.c_void_p(src_ptr), C.c_void_p(dst_ptr), n,
                                                                C.c_void_p(adj_ptr), pitch_words))
        return self.solution(0)


    # ---- tile-sharded large problem with NVLink peer stores (b2r_solve_sharded) ---------------------------
    def shard_export(self, n: int) -> bytes:
        buf = (C.c_ubyte * _lib.PEER_HANDLE_BYTES)()
        _lib.check(self._h, _lib.lib().b2r_shard_export(self._h, int(n), buf))
        return bytes(buf)

    def shard_attach(self, rank: int, world: int, handles) -> None:
        """handles: the shard_export() bytes of every rank, in rank order."""
        if len(handles) != world or any(len(h) != _lib.PEER_HANDLE_BYTES for h in handles):
            raise ValueError("shard_attach needs one exported handle per rank")
        blob = b"".join(handles)
        buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        _lib.check(self._h, _lib.lib().b2r_shard_attach(self._h, int(rank), int(world), buf))

    def solve_sharded(self, src_ptr, dst_ptr, n):
        self._batch = 1
        _lib.check(self._h, _lib.lib().b2r_solve_sharded(self._h, C.c_void_p(src_ptr), C.c_void_p(dst_ptr), int(n)))
        return self.solution(0)

    def shard_fell_back(self) -> bool:
        return bool(_lib.lib().b2r_shard_last_fell_back(self._h))

    def shard_detach(self) -> None:
        _lib.check(self._h, _lib.lib().b2r_shard_detach(self._h))


def pitch_words(n: int) -> int:
    return int(_lib.lib().b2r_pitch_words(n))
