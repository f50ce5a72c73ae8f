"""GPU tests at BASELINE.json's largest sizes: config 5 (N = 20 000) against the oracle's graph, and the documented maximum
(N = 65 535) through size-independent properties (planted clique recovered, symmetry, degrees = row popcounts, sampled rows
bit-equal to the reference arithmetic restated in numpy: registration.cc scale-invariant test, fp64, (x^2 + y^2) + z^2)."""
import numpy as np
import pytest

from poseestimator_b200 import synth

pytestmark = pytest.mark.gpu


def _params(tpp, nb, **kw):
    p = tpp.RobustRegistrationSolver.Params()
    p.noise_bound = nb
    p.estimate_scaling = False
    p.rotation_cost_threshold = 1e-12
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def _row_bits_numpy(src, dst, i, beta):
    """Row i of the consistency graph with the reference's fp64 order of operations (numpy never fuses multiply-adds)."""
    a = src - src[:, i : i + 1]
    b = dst - dst[:, i : i + 1]
    A = (a[0] * a[0] + a[1] * a[1]) + a[2] * a[2]
    B = (b[0] * b[0] + b[1] * b[1]) + b[2] * b[2]
    e = np.abs(np.sqrt(A) - np.sqrt(B)) <= beta
    e[i] = False
    return e


def _popcount_rows(bits):
    return np.unpackbits(bits.view(np.uint8), axis=1).sum(axis=1, dtype=np.int64)


def test_config5_graph_and_solution_at_n20000(oracle_mod):
    """Config 5's single-GPU leg: 2e8 pair tests bit-equal to the oracle, planted clique, pose."""
    import teaserpp_python as tpp
    pb = synth.config5()
    s = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound))
    sol = s.solve(pb.src, pb.dst)
    ref = oracle_mod.tim_graph(pb.src, pb.dst, pb.noise_bound)
    got = s._adjacency_bits()
    assert np.array_equal(got, ref)
    assert s.getInlierMaxClique() == pb.inliers.tolist()
    assert sol.valid and synth.rotation_error_rad(sol.rotation, pb.R) < 0.01
    assert np.linalg.norm(sol.translation - pb.t) < 0.01
    assert s.stats()["n_edges"] * 2 == int(_popcount_rows(got).sum())


@pytest.mark.parametrize("mode", [0, 2])
def test_maximum_size_properties(mode):
    """N = 65 535 (kMaxN), 98 % outliers: default tensor-core filter and the fp64 filter v2."""
    import teaserpp_python as tpp
    from poseestimator_b200 import _lib
    n = 65535
    pb = synth.make_problem(n, 0.98, 65535)
    s = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound, graph_exact_only=mode))
    sol = s.solve(pb.src, pb.dst)
    assert sol.valid
    clique = s.getInlierMaxClique()
    assert set(pb.inliers.tolist()) <= set(clique) and len(clique) <= len(pb.inliers) + 8
    assert synth.rotation_error_rad(sol.rotation, pb.R) < 0.01 and np.linalg.norm(sol.translation - pb.t) < 0.01
    bits = s._adjacency_bits()
    beta = 2 * pb.noise_bound
    rng = np.random.default_rng(3)
    rows = np.concatenate([[0, 127, 128, n - 1, n - 64], rng.integers(0, n, 19), pb.inliers[:4]])
    dense_cols = {}
    for i in rows:
        want = _row_bits_numpy(pb.src, pb.dst, int(i), beta)
        got = np.unpackbits(bits[int(i)].view(np.uint8), bitorder="little")[:n].astype(bool)
        assert np.array_equal(got, want), int(i)
        # symmetry: column i of the matrix equals row i
        col = (bits[:, int(i) >> 5] >> np.uint32(int(i) & 31)) & np.uint32(1)
        assert np.array_equal(col.astype(bool), want), int(i)
        dense_cols[int(i)] = int(want.sum())
    # padding bits beyond N are zero, the diagonal is empty
    assert not (bits[:, -1] >> np.uint32(n & 31)).any()
    deg = _popcount_rows(bits)
    assert s.stats()["n_edges"] * 2 == int(deg.sum())
    import ctypes as C
    cnt = C.c_int32(0)
    out = np.zeros(n, np.int32)
    _lib.check(s._h, _lib.lib().b2r_get_indices(s._h, 0, _lib.IDX_DEGREES, C.c_void_p(out.ctypes.data), C.byref(cnt)))
    assert cnt.value == n and np.array_equal(out.astype(np.int64), deg)
    for i, d in dense_cols.items():
        assert deg[i] == d
