"""b2r_voxel_coverage_noalloc (include/b2r.h; no counterpart in the reference): the voxel-coverage count with its hash table in a
file-scope static device array, called through the C ABI and compared with b2r_voxel_coverage on the same points.  The count is an
integer: the bar is equality.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _cloud(seed, n, spread):
    rng = np.random.default_rng(seed)
    return np.ascontiguousarray(rng.uniform(-spread, spread, size=(n, 3)))


def test_no_alloc_voxel_coverage_matches_workspace_variant_twice():
    import torch
    from poseestimator_b200 import _lib
    from poseestimator_b200.refine import Refiner
    L = _lib.lib()
    rf = Refiner(0)
    h = rf._h
    count_dev = torch.full((1,), -7, dtype=torch.int32, device="cuda:0")
    # two calls in a row on the shared table: the second, smaller cloud must not see the first call's entries; then the first again
    cases = [(11, 50000, 0.30, 0.01), (12, 3000, 0.05, 0.004), (11, 50000, 0.30, 0.01), (13, 1, 1.0, 0.1), (14, 0, 1.0, 0.1)]
    for seed, n, spread, voxel in cases:
        pts = torch.from_numpy(_cloud(seed, max(n, 1), spread)).to("cuda:0")
        want = C.c_int(-1)
        _lib.check(h, L.b2r_voxel_coverage(h, pts.data_ptr(), n, voxel, C.byref(want)))
        _lib.check(h, L.b2r_voxel_coverage_noalloc(h, pts.data_ptr(), n, voxel, count_dev.data_ptr()))
        torch.cuda.synchronize()
        got = int(count_dev.cpu()[0])
        assert got == want.value, (seed, n, got, want.value)
        if n > 1:   # and against numpy on the host (floor of the fp64 quotient, the reference's expression)
            ref = len(np.unique(np.floor(pts.cpu().numpy()[:n] / voxel).astype(np.int32), axis=0))
            assert got == ref, (seed, n, got, ref)
    # python wrapper = the same entry point
    pts = torch.from_numpy(_cloud(15, 20000, 0.2)).to("cuda:0")
    rf.voxel_coverage_noalloc(pts, 0.01, count_dev)
    torch.cuda.synchronize()
    assert int(count_dev.cpu()[0]) == rf.voxel_coverage(pts, 0.01)


def test_no_alloc_rejects_sizes_beyond_the_static_table():
    import torch
    from poseestimator_b200 import _lib
    from poseestimator_b200.refine import Refiner
    L = _lib.lib()
    rf = Refiner(0)
    h = rf._h
    count_dev = torch.zeros(1, dtype=torch.int32, device="cuda:0")
    pts = torch.zeros((8, 3), dtype=torch.float64, device="cuda:0")   # never read: the size check comes first
    rc = L.b2r_voxel_coverage_noalloc(h, pts.data_ptr(), Refiner.VOXEL_NOALLOC_MAX_POINTS + 1, 0.01, count_dev.data_ptr())
    assert rc == 5   # B2R_ERR_LIMIT
    msg = L.b2r_last_error(h).decode()
    assert "B2R_VOXEL_NOALLOC_MAX_POINTS" in msg and str(Refiner.VOXEL_NOALLOC_MAX_POINTS) in msg
    # the largest size the table holds
    n = Refiner.VOXEL_NOALLOC_MAX_POINTS
    big = torch.from_numpy(_cloud(16, n, 1.0)).to("cuda:0")
    _lib.check(h, L.b2r_voxel_coverage_noalloc(h, big.data_ptr(), n, 0.02, count_dev.data_ptr()))
    torch.cuda.synchronize()
    assert int(count_dev.cpu()[0]) == rf.voxel_coverage(big, 0.02)
