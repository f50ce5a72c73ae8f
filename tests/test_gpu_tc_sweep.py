"""The tensor-core graph filter's exactness argument, measured over WHOLE matrices (graph_tc.cu header).

The band of the filter assumes the tcgen05 accumulation error stays below 2^-20 of the sum of |products| and bounds the total
error of Dm~ / L~ by E_D / E_L.  b2r_debug_tc_sweep recomputes every 128 x 64 unit of a problem on the tensor cores and checks both
against fp64 on the device; the adjacency is then compared with the oracle bit for bit.  Cases: config 3, config 5, and
constructed clouds at the edge of the fp32 path's gate (extent / beta up to 2900), anisotropic, and offset by 1e6.
"""
import numpy as np
import pytest

from poseestimator_b200 import synth

pytestmark = pytest.mark.gpu

ACCUM_BOUND = 2.0 ** -21     # half of what the band assumes


def _params(tpp, nb):
    p = tpp.RobustRegistrationSolver.Params()
    p.noise_bound = nb
    p.estimate_scaling = False
    p.rotation_cost_threshold = 1e-12
    return p


def _solve_and_sweep(tpp, src, dst, nb):
    s = tpp.RobustRegistrationSolver(_params(tpp, nb))
    s.solve(src, dst)
    return s, s._tc_sweep()


def _report(name, sw):
    print(f"{name}: pairs {int(sw['pairs'])}, accumulation error / sum|products| = 2^{np.log2(max(sw['tc_accum_ratio'], 1e-300)):.2f}, "
          f"|Dm~-Dm| / E_D = {sw['dm_over_ED']:.4f}, |L~-L| / E_L = {sw['l_over_EL']:.4f}, Pi = {sw['Pi']:.4g}")


def _assert_sweep(sw, n):
    assert sw["use32"] == 1.0
    T = (n + 127) // 128
    assert sw["pairs"] >= n * (n - 1) / 2          # every pair of the upper triangle (diagonal tiles twice)
    assert sw["pairs"] <= n * n
    assert sw["tc_accum_ratio"] <= ACCUM_BOUND, "tcgen05 accumulation error exceeds half of the assumed bound"
    assert sw["dm_over_ED"] <= 0.5 and sw["l_over_EL"] <= 0.5
    assert T >= 1


@pytest.mark.parametrize("name", ["config3", "config5"])
def test_sweep_every_unit_of_the_benchmark_problems(oracle_mod, name):
    import teaserpp_python as tpp
    pb = synth.config3() if name == "config3" else synth.config5()
    n = pb.src.shape[1]
    s, sw = _solve_and_sweep(tpp, pb.src, pb.dst, pb.noise_bound)
    _report(name, sw)
    _assert_sweep(sw, n)
    assert np.array_equal(s._adjacency_bits(), oracle_mod.tim_graph(pb.src, pb.dst, pb.noise_bound))


def _cloud(rng, n, half_extent, aniso, offset, beta):
    """src uniform in a box of the given half extents (in units of beta), dst = rigid copy + noise with half of it replaced by
    outliers drawn from the same box (so that the bounding boxes, hence Pi, stay at the requested size)."""
    he = np.array([half_extent, half_extent / aniso, half_extent / aniso]) * beta
    src = rng.uniform(-1, 1, (3, n)) * he[:, None]
    dst = src + rng.normal(0, beta / 8, (3, n))
    out = rng.choice(n, n // 2, replace=False)
    dst[:, out] = rng.uniform(-1, 1, (3, len(out))) * he[:, None]
    return src + offset, dst + offset


@pytest.mark.parametrize("half_extent,aniso,offset", [
    (100.0, 1.0, 0.0), (1000.0, 1.0, 0.0), (2900.0, 1.0, 0.0),      # extent / beta up to just inside the E_L < 1e3 gate
    (2900.0, 1000.0, 0.0), (1000.0, 1000.0, 0.0),                    # one axis 1000 x the others
    (100.0, 1.0, 1.0e6), (1000.0, 1000.0, 1.0e6),                    # offset by 1e6 before centring
])
def test_sweep_constructed_clouds(oracle_mod, half_extent, aniso, offset):
    import teaserpp_python as tpp
    rng = np.random.default_rng(int(half_extent) + int(aniso) + (1 if offset else 0))
    n, nb = 1700, 0.01
    beta = 2.0 * nb
    src, dst = _cloud(rng, n, half_extent, aniso, offset, beta)
    s, sw = _solve_and_sweep(tpp, src, dst, nb)
    _report(f"half extent {half_extent} beta, aniso {aniso}, offset {offset}", sw)
    _assert_sweep(sw, n)
    assert sw["E_L"] < 1.0e3
    assert np.array_equal(s._adjacency_bits(), oracle_mod.tim_graph(src, dst, nb))


def test_beyond_the_gate_the_fp64_filter_decides(oracle_mod):
    """extent / beta = 6000: E_L >= 1e3, the fp32 path refuses the problem (use32 = 0) and the bits are still the reference's."""
    import teaserpp_python as tpp
    rng = np.random.default_rng(9)
    n, nb = 900, 0.01
    src, dst = _cloud(rng, n, 6000.0, 1.0, 0.0, 2.0 * nb)
    s, sw = _solve_and_sweep(tpp, src, dst, nb)
    assert sw["use32"] == 0.0
    assert np.array_equal(s._adjacency_bits(), oracle_mod.tim_graph(src, dst, nb))
