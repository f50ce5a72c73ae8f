"""GPU parity tests: CUDA path (through the C ABI / teaserpp_python surface) vs the CPU oracle.

Bars (BASELINE.json north_star): adjacency, max-clique and index lists bit-exact; rotation within 1e-5 rad;
translation within 1e-6 of the scene scale.
"""
import itertools

import numpy as np
import pytest

from poseestimator_b200 import synth

pytestmark = pytest.mark.gpu

ROT_TOL = 1e-5          # rad
TRANS_TOL_REL = 1e-6    # x scene scale


def _params(tpp, nb, **kw):
    p = tpp.RobustRegistrationSolver.Params()
    p.noise_bound = nb
    p.estimate_scaling = False
    p.rotation_cost_threshold = 1e-12
    for k, v in kw.items():
        setattr(p, k, v)
    return p


def _check_against_oracle(solver, sol, ref, pb):
    assert np.array_equal(solver._adjacency_bits(), ref.adjacency)
    assert solver.getInlierMaxClique() == ref.clique.tolist()
    assert sol.valid == ref.valid
    if ref.valid:
        assert synth.rotation_error_rad(sol.rotation, ref.rotation) < ROT_TOL
        assert np.linalg.norm(sol.translation - ref.translation) < TRANS_TOL_REL * pb.scene_scale
        assert solver.getRotationInliers() == ref.rotation_inliers.tolist()
        assert solver.getTranslationInliers() == ref.translation_inliers.tolist()
        assert abs(solver.stats()["gnc_iterations"] - ref.gnc_iters) <= 1
    assert solver.stats()["n_edges"] == ref.n_edges


@pytest.mark.parametrize("n,ratio,seed", [(50, 0.3, 1), (129, 0.6, 2), (400, 0.6, 7), (1000, 0.5, 1000), (1000, 0.9, 3), (2049, 0.8, 4)])
def test_solve_matches_oracle(oracle_mod, n, ratio, seed):
    import teaserpp_python as tpp
    pb = synth.make_problem(n, ratio, seed)
    solver = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound))
    sol = solver.solve(pb.src, pb.dst)
    ref = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=0, rotation_cost_threshold=1e-12)
    _check_against_oracle(solver, sol, ref, pb)
    assert set(pb.inliers.tolist()) <= set(solver.getInlierMaxClique())


def test_config3_full_size(oracle_mod):
    """BASELINE config 3: N=5000, 95 % outliers."""
    import teaserpp_python as tpp
    pb = synth.config3()
    solver = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound))
    sol = solver.solve(pb.src, pb.dst)
    ref = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=0, rotation_cost_threshold=1e-12,
                           materialize_tims=0)
    _check_against_oracle(solver, sol, ref, pb)
    assert solver.getInlierMaxClique() == pb.inliers.tolist()
    assert synth.rotation_error_rad(sol.rotation, pb.R) < 0.01


def test_noisy_gnc_iterates(oracle_mod):
    """Tight noise bound so that GNC really iterates and rejects TIMs."""
    import teaserpp_python as tpp
    pb = synth.make_problem(300, 0.3, 11, sigma=0.02)
    nb = 0.03
    solver = tpp.RobustRegistrationSolver(_params(tpp, nb))
    sol = solver.solve(pb.src, pb.dst)
    ref = oracle_mod.solve(pb.src, pb.dst, noise_bound=nb, estimate_scaling=0, rotation_cost_threshold=1e-12)
    assert ref.gnc_iters > 3
    _check_against_oracle(solver, sol, ref, pb)
    w = np.zeros(solver.stats()["n_rot_tims"])
    assert np.allclose(solver.getRotationInliersMask()[0], ref.rot_mask)


@pytest.mark.parametrize("tau", [0.0, 1e-3, 0.5])
def test_graph_filter_equals_exact_kernel(oracle_mod, tau):
    """The sqrt-free filter (any ambiguity band) and the two-sqrt kernel give the same bits."""
    import teaserpp_python as tpp
    pb = synth.make_problem(777, 0.7, 21)
    ref = oracle_mod.tim_graph(pb.src, pb.dst, pb.noise_bound)
    a = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound, graph_exact_only=True))
    a.solve(pb.src, pb.dst)
    b = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound, graph_filter_tau=tau))
    b.solve(pb.src, pb.dst)
    assert np.array_equal(a._adjacency_bits(), ref)
    assert np.array_equal(b._adjacency_bits(), ref)
    if tau >= 1e-3:
        assert b.stats()["graph_slow_pairs"] > 0


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4, 5])
@pytest.mark.parametrize("n,ratio,seed", [(300, 0.5, 31), (1111, 0.8, 32)])
def test_every_filter_generation_matches_oracle(oracle_mod, mode, n, ratio, seed):
    """graph_exact_only = 0 (packed-fp32 norm-expansion filter v4), 1 (two square roots), 2 / 3 (fp64 filters v2 / v1), 4 (packed-fp32 filter v3): same bits."""
    import teaserpp_python as tpp
    pb = synth.make_problem(n, ratio, seed)
    ref = oracle_mod.tim_graph(pb.src, pb.dst, pb.noise_bound)
    s = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound, graph_exact_only=mode))
    s.solve(pb.src, pb.dst)
    assert np.array_equal(s._adjacency_bits(), ref)


@pytest.mark.parametrize("offset,extent,nb", [
    (0.0, 1.0, 0.05),          # ordinary
    (1.0e4, 1.0, 0.05),        # cloud far from the origin: the fp32 copies are centred, the filter still decides
    (0.0, 50.0, 1e-3),         # extent / noise bound = 5e4: wide band, most chunks redone in fp64
    (3.0e6, 2.0e3, 1e-4),      # band saturates -> everything through the exact path
    (0.0, 1e-3, 1e-7),         # beta below the fp32 filter's range -> fp64 filter
    (0.0, 0.0, 0.05),          # all points identical (M = 0)
])
def test_fp32_filter_bands_and_fallbacks(oracle_mod, offset, extent, nb):
    """The packed-fp32 filter is only a filter: whatever its band does, the bits are the fp64 reference's."""
    import teaserpp_python as tpp
    rng = np.random.default_rng(77)
    n = 333
    src = offset + rng.uniform(-extent, extent, (3, n))
    R = synth.random_rotation(rng)
    dst = R @ src + rng.uniform(-1, 1, (3, 1)) * extent + rng.normal(0, nb / 4, (3, n))
    dst[:, : n // 2] = offset + rng.uniform(-2 * extent, 2 * extent, (3, n // 2))
    ref = oracle_mod.tim_graph(src, dst, nb)
    s = tpp.RobustRegistrationSolver(_params(tpp, nb))
    s.solve(src, dst)
    assert np.array_equal(s._adjacency_bits(), ref)


def test_fp32_filter_non_finite_input(oracle_mod):
    """inf / NaN coordinates: the prep kernel routes the problem to the fp64 path; verdicts follow IEEE compares."""
    import teaserpp_python as tpp
    rng = np.random.default_rng(78)
    n = 140
    src = rng.uniform(-1, 1, (3, n)); dst = src + rng.normal(0, 0.01, (3, n))
    src[0, 5] = np.inf; dst[1, 77] = np.nan
    ref = oracle_mod.tim_graph(src, dst, 0.05)
    s = tpp.RobustRegistrationSolver(_params(tpp, 0.05, inlier_selection_mode=tpp.RobustRegistrationSolver.INLIER_SELECTION_MODE.NONE))
    s.solve(src, dst)
    assert np.array_equal(s._adjacency_bits(), ref)


def test_borderline_pairs_follow_reference_rounding(oracle_mod):
    """Pairs constructed to sit exactly on / one ulp around the bound: verdicts must follow the fp64 reference order."""
    import teaserpp_python as tpp
    rng = np.random.default_rng(5)
    n = 256
    src = rng.uniform(-1, 1, (3, n))
    dst = src.copy()
    nb = 0.05
    beta = 2 * nb
    # stretch selected points so that | |dst_j-dst_0| - |src_j-src_0| | is beta +- few ulps
    for j in range(1, n):
        v = src[:, j] - src[:, 0]
        L = np.linalg.norm(v)
        target = L + beta * (1 + (j % 5 - 2) * 2.2e-16)
        dst[:, j] = dst[:, 0] + v / L * target
    ref = oracle_mod.tim_graph(src, dst, nb)
    for kw in (dict(graph_exact_only=True), dict(), dict(graph_exact_only=2), dict(graph_exact_only=3), dict(graph_exact_only=4), dict(graph_exact_only=5)):
        s = tpp.RobustRegistrationSolver(_params(tpp, nb, **kw))
        s.solve(src, dst)
        assert np.array_equal(s._adjacency_bits(), ref)
    alt = oracle_mod.tim_graph(src, dst, nb, variant=1)  # borderline counter: other association
    print("borderline flips under x2+(y2+z2):", int((np.unpackbits((alt ^ ref).view(np.uint8))).sum()))


def test_degenerate_inputs(oracle_mod):
    import teaserpp_python as tpp
    rng = np.random.default_rng(1)
    src = rng.uniform(-1, 1, (3, 30)); dst = rng.uniform(-1, 1, (3, 30))
    s = tpp.RobustRegistrationSolver(_params(tpp, 1e-9))
    sol = s.solve(src, dst)
    assert not sol.valid and np.array_equal(sol.rotation, np.eye(3)) and np.array_equal(sol.translation, np.zeros(3))
    assert len(s.getInlierMaxClique()) <= 1 and s.getRotationInliers() == [] and s.getTranslationInliers() == []
    assert list(s.getScaleInliers() or []) == []
    R = synth.random_rotation(rng); t = np.array([0.3, -0.2, 0.1])
    dst = R @ src + t[:, None]
    s = tpp.RobustRegistrationSolver(_params(tpp, 0.01))
    sol = s.solve(src, dst)
    assert sol.valid and s.stats()["gnc_iterations"] == 0 and len(s.getInlierMaxClique()) == 30
    assert synth.rotation_error_rad(sol.rotation, R) < 1e-7 and np.allclose(sol.translation, t, atol=1e-7)
    src2 = src.copy(); src2[:, 1] = src2[:, 0]
    sol = s.solve(src2, R @ src2 + t[:, None])
    assert sol.valid and len(s.getInlierMaxClique()) == 30
    for n in (2, 3):
        s = tpp.RobustRegistrationSolver(_params(tpp, 0.01))
        sol = s.solve(src[:, :n], dst[:, :n])
        ref = oracle_mod.solve(src[:, :n], dst[:, :n], noise_bound=0.01, estimate_scaling=0, rotation_cost_threshold=1e-12)
        assert sol.valid and len(s.getInlierMaxClique()) == n
        assert synth.rotation_error_rad(sol.rotation, ref.rotation) < 1e-5 or n == 2
    with pytest.raises(TypeError):
        s.solve(np.zeros((4, 10)), np.zeros((4, 10)))
    with pytest.raises(TypeError):
        s.solve(np.zeros((3, 10)), np.zeros((3, 11)))


def test_argument_layouts_and_reference_call_pattern(oracle_mod):
    """The reference passes np.array([...]).T (F-order view); also C-order, float32 and a re-used solver."""
    from teaserpp_python import _teaserpp as tpp
    pb = synth.make_problem(200, 0.5, 31)
    p = tpp.RobustRegistrationSolver.Params()
    p.cbar2 = 1
    p.noise_bound = float(pb.noise_bound)
    p.estimate_scaling = False
    p.rotation_estimation_algorithm = tpp.RotationEstimationAlgorithm.GNC_TLS
    p.rotation_gnc_factor = 1.4
    p.rotation_max_iterations = 100
    p.rotation_cost_threshold = 1e-12
    outs = []
    for src, dst in ((pb.src, pb.dst), (np.ascontiguousarray(pb.src), np.ascontiguousarray(pb.dst)),
                     (pb.src[:, ::1], np.asfortranarray(pb.dst))):
        solver = tpp.RobustRegistrationSolver(p)
        solver.solve(src, dst)
        sol = solver.getSolution()
        R_inl = list(solver.getRotationInliers() or [])
        s_inl = list(solver.getScaleInliers() or [])
        t_inl = list(solver.getTranslationInliers() or [])
        outs.append((np.asarray(sol.rotation).copy(), np.asarray(sol.translation).reshape(3).copy(), R_inl, len(s_inl), t_inl))
    for o in outs[1:]:
        assert np.array_equal(o[0], outs[0][0]) and np.array_equal(o[1], outs[0][1]) and o[2:] == outs[0][2:]
    ref = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=0, rotation_cost_threshold=1e-12,
                           want_scale_mask=True)
    assert outs[0][3] == ref.n_edges
    solver = tpp.RobustRegistrationSolver(p)
    solver.solve(pb.src, pb.dst)
    A = np.unpackbits(ref.adjacency.view(np.uint8), axis=1, bitorder="little")[:, :200].astype(bool)
    assert solver.getScaleInliers() == list(zip(*[x.tolist() for x in np.nonzero(np.triu(A, 1))]))
    assert np.array_equal(solver.getScaleInliersMask()[0], ref.scale_mask.astype(bool))
    assert solver.getInlierGraph() == [np.flatnonzero(r).tolist() for r in A]
    assert np.array_equal(solver.getRotationInliersMap()[0], ref.clique[ref.rotation_inliers])
    assert np.array_equal(solver.getTranslationInliersMap()[0], ref.clique[ref.rotation_inliers][ref.translation_inliers])


def test_torch_cuda_inputs_zero_copy(oracle_mod):
    import torch
    import teaserpp_python as tpp
    pb = synth.make_problem(300, 0.5, 41)
    s = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound))
    a = s.solve(pb.src, pb.dst)
    ts = torch.from_numpy(np.ascontiguousarray(pb.src.T)).cuda().T   # (3,N) view with strides (1,3)
    td = torch.from_numpy(np.ascontiguousarray(pb.dst.T)).cuda().T
    s2 = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound))
    b = s2.solve(ts, td)
    assert np.array_equal(a.rotation, b.rotation) and np.array_equal(a.translation, b.translation)
    tc = torch.from_numpy(np.ascontiguousarray(pb.src)).cuda()      # C-order (3,N): strided gather path
    tdc = torch.from_numpy(np.ascontiguousarray(pb.dst)).cuda()
    s3 = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound))
    c = s3.solve(tc, tdc)
    assert np.array_equal(a.rotation, c.rotation) and s3.getInlierMaxClique() == s.getInlierMaxClique()


def _bits_from_dense(A):
    n = A.shape[0]
    W = (n + 31) // 32
    pad = np.zeros((n, W * 32), np.uint8)
    pad[:, :n] = A
    return np.packbits(pad, axis=1, bitorder="little").view(np.uint32).reshape(n, W)


def _clique_of_graph(A, mode=0, kcore_thr=0.5):
    """Feed an explicit graph to the GPU clique stages through b2r_solve_from_adjacency."""
    import torch
    import teaserpp_python as tpp
    from poseestimator_b200 import batch
    n = A.shape[0]
    pitch = batch.pitch_words(n)
    bits = np.zeros((n, pitch), np.uint32)
    bits[:, : (n + 31) // 32] = _bits_from_dense(A)
    p = _params(tpp, 0.01, inlier_selection_mode=mode, kcore_heuristic_threshold=kcore_thr)
    bs = batch.BatchSolver(p)
    pts = torch.zeros((n, 3), dtype=torch.float64, device="cuda")
    adj = torch.from_numpy(bits.view(np.int32)).cuda()
    torch.cuda.synchronize()
    bs.solve_from_adjacency(pts.data_ptr(), pts.data_ptr(), n, adj.data_ptr(), pitch)
    return bs.max_clique(0).tolist(), bs.stats(0)


@pytest.mark.parametrize("seed", range(10))
def test_max_clique_random_graphs_with_ties(oracle_mod, seed):
    """Exact search + canonical tie-break on graphs with many maximum cliques (heuristic alone is not enough)."""
    rng = np.random.default_rng(100 + seed)
    n = int(rng.integers(20, 400))
    p = rng.uniform(0.05, 0.7) if n < 150 else rng.uniform(0.05, 0.3)
    A = rng.random((n, n)) < p
    A = np.triu(A, 1); A = A | A.T
    want = oracle_mod.max_clique_bits(_bits_from_dense(A)).tolist()
    got, st = _clique_of_graph(A)
    assert got == want, (n, p, st)
    assert st["exact_proven"] == 1


@pytest.mark.parametrize("seed", range(12))
def test_near_clique_core_decided_by_fused_tail(oracle_mod, seed):
    """A planted clique plus a few vertices adjacent to almost all of it: the (lb-1)-core is larger than the greedy clique, several
    maximum cliques may exist.  The fused tail settles it itself (minimum vertex cover of the core's non-adjacent pairs) and must
    return the oracle's canonical (lexicographically smallest) maximum clique."""
    rng = np.random.default_rng(900 + seed)
    n = int(rng.integers(200, 700))
    c = int(rng.integers(30, 90))
    e = 1 + seed % 8
    A = rng.random((n, n)) < 0.03
    A = np.triu(A, 1); A = A | A.T
    ids = rng.permutation(n)
    planted, extra = ids[:c], ids[c:c + e]
    A[np.ix_(planted, planted)] = True
    for x in extra:                           # adjacent to all of the clique but 1 .. 2 members, and to the other extras at random
        A[x, planted] = True; A[planted, x] = True
        miss = rng.choice(planted, size=int(rng.integers(1, 3)), replace=False)
        A[x, miss] = False; A[miss, x] = False
        for y in extra:
            if y != x:
                v = bool(rng.random() < 0.5); A[x, y] = v; A[y, x] = v
    if seed % 3 == 0:                         # and one member that lost an edge inside the clique (two maximum cliques through it)
        a, b = planted[:2]
        A[a, b] = False; A[b, a] = False
    np.fill_diagonal(A, False)
    want = oracle_mod.max_clique_bits(_bits_from_dense(A)).tolist()
    got, st = _clique_of_graph(A)
    assert got == want, (n, c, e, st)
    assert st["omega"] == len(want) and st["exact_proven"] == 1
    if st["n_pruned"] - st["heuristic_lb"] <= 8:
        assert st["fused_tail"] == 1, st      # near-clique core: no hand-over to the multi-kernel chain


def test_max_clique_planted_and_modes(oracle_mod):
    rng = np.random.default_rng(3)
    n = 600
    A = rng.random((n, n)) < 0.05
    A = np.triu(A, 1); A = A | A.T
    planted = np.sort(rng.permutation(n)[:40])
    A[np.ix_(planted, planted)] = True
    np.fill_diagonal(A, False)
    bits = _bits_from_dense(A)
    got, st = _clique_of_graph(A)
    assert got == planted.tolist() and st["omega"] == 40
    for mode in (1, 3):
        got, _ = _clique_of_graph(A, mode=mode)
        assert got == oracle_mod.max_clique_bits(bits, mode=mode).tolist()
    got, _ = _clique_of_graph(A, mode=2, kcore_thr=0.05)
    assert got == oracle_mod.max_clique_bits(bits, mode=2, kcore_thr=0.05).tolist()
    got, _ = _clique_of_graph(A, mode=2, kcore_thr=0.9)   # threshold not met -> heuristic fall-through
    assert got == oracle_mod.max_clique_bits(bits, mode=2, kcore_thr=0.9).tolist()


def test_batch_mixed_sizes_matches_single_solves(oracle_mod):
    import teaserpp_python as tpp
    from poseestimator_b200 import batch
    pbs = [synth.config4_detection(0, d) for d in range(12)]
    nb = pbs[0].noise_bound
    bs = batch.BatchSolver(_params(tpp, nb))
    sols = bs.solve([(p.src, p.dst) for p in pbs])
    for i, pb in enumerate(pbs):
        ref = oracle_mod.solve(pb.src, pb.dst, noise_bound=nb, estimate_scaling=0, rotation_cost_threshold=1e-12)
        assert bs.max_clique(i).tolist() == ref.clique.tolist()
        assert np.array_equal(bs.adjacency(i, pb.src.shape[1]), ref.adjacency)
        assert synth.rotation_error_rad(sols[i].rotation, ref.rotation) < ROT_TOL
        assert np.linalg.norm(sols[i].translation - ref.translation) < TRANS_TOL_REL * pb.scene_scale
    # a second batch on the same handle (workspace re-use) and determinism
    sols2 = bs.solve([(p.src, p.dst) for p in pbs[:5]])
    for a, b in zip(sols2, sols[:5]):
        assert np.array_equal(a.rotation, b.rotation) and np.array_equal(a.translation, b.translation)


def test_complete_tim_graph_and_none_mode(oracle_mod):
    import teaserpp_python as tpp
    pb = synth.make_problem(150, 0.4, 51)
    for kw, okw in ((dict(rotation_tim_graph=tpp.RobustRegistrationSolver.INLIER_GRAPH_FORMULATION.COMPLETE), dict(rotation_tim_graph=1)),
                    (dict(inlier_selection_mode=tpp.RobustRegistrationSolver.INLIER_SELECTION_MODE.NONE), dict(inlier_selection_mode=3))):
        s = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound, **kw))
        sol = s.solve(pb.src, pb.dst)
        ref = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=0, rotation_cost_threshold=1e-12, **okw)
        assert s.getInlierMaxClique() == ref.clique.tolist()
        assert synth.rotation_error_rad(sol.rotation, ref.rotation) < ROT_TOL
        assert np.linalg.norm(sol.translation - ref.translation) < TRANS_TOL_REL * pb.scene_scale
        assert s.getRotationInliers() == ref.rotation_inliers.tolist()


def test_row_sharded_graph_equals_full(oracle_mod):
    """Config-5 building block: rows computed in shards (as 4 ranks would) == full graph; solve continues from it."""
    import torch
    import teaserpp_python as tpp
    from poseestimator_b200 import batch
    pb = synth.make_problem(1500, 0.9, 61)
    n = 1500
    pitch = batch.pitch_words(n)
    bs = batch.BatchSolver(_params(tpp, pb.noise_bound))
    src = torch.from_numpy(np.ascontiguousarray(pb.src.T)).cuda()
    dst = torch.from_numpy(np.ascontiguousarray(pb.dst.T)).cuda()
    adj = torch.zeros((n, pitch), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    bounds = [0, 375, 750, 1125, 1500]
    for a, b in zip(bounds[:-1], bounds[1:]):
        bs.graph_rows(src.data_ptr(), dst.data_ptr(), n, a, b, adj[a:b].data_ptr(), pitch)
    sol = bs.solve_from_adjacency(src.data_ptr(), dst.data_ptr(), n, adj.data_ptr(), pitch)
    ref = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=0, rotation_cost_threshold=1e-12)
    got = adj.cpu().numpy().view(np.uint32)[:, : (n + 31) // 32]
    assert np.array_equal(got, ref.adjacency)
    assert bs.max_clique(0).tolist() == ref.clique.tolist()
    assert synth.rotation_error_rad(sol.rotation, ref.rotation) < ROT_TOL


# ---- estimate_scaling=True: adaptive voting over all TIM ratios ---------------------
SCALE_TOL_REL = 1e-9


@pytest.mark.parametrize("n,ratio,seed,gen_scale", [(40, 0.2, 21, 1.0), (200, 0.5, 22, 1.7), (333, 0.7, 23, 0.6), (1000, 0.5, 24, 2.5)])
def test_scale_estimation_matches_oracle(oracle_mod, n, ratio, seed, gen_scale):
    import teaserpp_python as tpp
    pb = synth.make_problem(n, ratio, seed, scale=gen_scale)
    solver = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound, estimate_scaling=True))
    sol = solver.solve(pb.src, pb.dst)
    ref = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=1, rotation_cost_threshold=1e-12)
    assert abs(sol.scale - ref.scale) < SCALE_TOL_REL * abs(ref.scale)
    assert abs(sol.scale - gen_scale) < 0.05 * gen_scale
    # the graph is the scale-inlier mask: bit-exact, and so is everything downstream
    assert np.array_equal(solver._adjacency_bits(), ref.adjacency)
    assert solver.getInlierMaxClique() == ref.clique.tolist()
    assert sol.valid and ref.valid
    assert synth.rotation_error_rad(sol.rotation, ref.rotation) < ROT_TOL
    assert np.linalg.norm(sol.translation - ref.translation) < TRANS_TOL_REL * pb.scene_scale * max(1.0, gen_scale)
    assert solver.getRotationInliers() == ref.rotation_inliers.tolist()
    assert solver.getTranslationInliers() == ref.translation_inliers.tolist()
    assert solver.stats()["n_edges"] == ref.n_edges
    assert len(solver.getScaleInliers()) == ref.n_edges


def test_scale_estimation_in_a_batch(oracle_mod):
    """Problems of different sizes in one batch share the scale-vote workspace one after the other."""
    import teaserpp_python as tpp
    from poseestimator_b200 import batch as b2batch
    pbs = [synth.make_problem(n, 0.4, 30 + i, scale=s) for i, (n, s) in enumerate([(150, 1.3), (64, 0.8), (300, 2.0)])]
    bs = b2batch.BatchSolver(_params(tpp, pbs[0].noise_bound, estimate_scaling=True))
    sols = bs.solve([(p.src, p.dst) for p in pbs])
    for i, (pb, sol) in enumerate(zip(pbs, sols)):
        ref = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=1, rotation_cost_threshold=1e-12)
        assert abs(sol.scale - ref.scale) < SCALE_TOL_REL * abs(ref.scale)
        assert bs.max_clique(i).tolist() == ref.clique.tolist()
        assert np.array_equal(bs.adjacency(i, pb.src.shape[1]), ref.adjacency)


def test_non_unique_maximum_clique_is_canonical_and_fast(oracle_mod):
    """Seeds whose consistency graph has several maximum cliques (an outlier consistent with all but one inlier): the
    answer must be the lexicographically smallest one, as in the oracle, and the canonicalisation must not fall back to
    hundreds of serial decision searches (it took 150 ms per batch before the universal-member shortcut)."""
    import teaserpp_python as tpp
    from poseestimator_b200 import batch as b2batch
    pbs = [synth.make_problem(5000, 0.95, 105000 + i) for i in range(16)]
    bs = b2batch.BatchSolver(_params(tpp, pbs[0].noise_bound), profiling=True)
    bs.solve([(p.src, p.dst) for p in pbs])
    sols = bs.solve([(p.src, p.dst) for p in pbs])
    tm = bs.timings_ms()
    n_multi = 0
    for i, pb in enumerate(pbs):
        ref = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=0, rotation_cost_threshold=1e-12,
                               materialize_tims=0, want_adjacency=False)
        assert bs.max_clique(i).tolist() == ref.clique.tolist()
        assert sols[i].valid
        n_multi += bs.stats(i)["n_pruned"] > bs.stats(i)["omega"]
    assert n_multi >= 1, "no problem of this batch has extra vertices in its (omega-1)-core; pick other seeds"
    assert tm["canonical"] < 5.0 and tm["exact_search"] < 5.0, tm
    assert all(bs.stats(i)["fused_tail"] == 1 for i in range(len(pbs))), "a near-clique core was handed to the multi-kernel chain"


def test_batched_solution_getter_equals_per_problem_getter():
    from poseestimator_b200 import batch as b2batch
    import teaserpp_python as tpp
    pbs = [synth.make_problem(200 + 10 * i, 0.5, 40 + i) for i in range(5)]
    bs = b2batch.BatchSolver(_params(tpp, pbs[0].noise_bound))
    sols = bs.solve([(p.src, p.dst) for p in pbs])
    arr = bs.solutions_array()
    assert arr.shape == (5,) and arr["valid"].all()
    for i, s in enumerate(sols):
        one = bs.solution(i)
        assert np.array_equal(arr["rotation"][i], one.rotation) and np.array_equal(arr["translation"][i], one.translation)
        assert np.array_equal(s.rotation, one.rotation) and s.valid == one.valid and s.scale == one.scale


def _same_solutions(a, b):
    return len(a) == len(b) and all(
        np.ascontiguousarray(a[f]).tobytes() == np.ascontiguousarray(b[f]).tobytes() for f in ("scale", "rotation", "translation", "valid"))


def test_page_locked_inputs_equal_staged_inputs():
    """on_device == 2 (the copy engine reads the caller's page-locked arrays): one merged copy per operand for a contiguous
    block, one copy per problem for scattered arrays, mixed sizes, and the strided single-solve entry; same bits as staging."""
    from poseestimator_b200 import batch as b2batch
    import teaserpp_python as tpp
    B, n = 6, 300
    pbs = [synth.make_problem(n, 0.6, 70 + i) for i in range(B)]
    nb = pbs[0].noise_bound
    bs = b2batch.BatchSolver(_params(tpp, nb))
    packed = [(np.ascontiguousarray(p.src.T), np.ascontiguousarray(p.dst.T)) for p in pbs]
    bs.solve_packed_async(packed); bs.sync()
    want = bs.solutions_array().copy()
    # contiguous block
    ps, pd = b2batch.pinned_problem_block(B, n)
    for i, (a, b) in enumerate(packed):
        ps[i] = a; pd[i] = b
    bs.solve_pinned_async([(ps[i], pd[i]) for i in range(B)]); bs.sync()
    got = bs.solutions_array()
    assert _same_solutions(got, want) and bool(got["valid"].all())
    hb = b2batch.HostBatch([(ps[i], pd[i]) for i in range(B)], page_locked=True)   # pointer tables built once, arrays rewritten in place
    ps[...] = ps[::-1].copy(); pd[...] = pd[::-1].copy()
    bs.solve_host_batch_async(hb); bs.sync()
    assert _same_solutions(bs.solutions_array()[::-1], want)
    ps[...] = ps[::-1].copy(); pd[...] = pd[::-1].copy()
    bs.solve_host_batch_async(b2batch.HostBatch(packed)); bs.sync()
    assert _same_solutions(bs.solutions_array(), want)
    # scattered: reversed order inside the block, so no two operands are adjacent in the upload order
    bs.solve_pinned_async([(ps[i], pd[i]) for i in reversed(range(B))]); bs.sync()
    got = bs.solutions_array()
    assert _same_solutions(got[::-1], want)
    # mixed sizes carved from one buffer
    sizes = [50, 333, 128, 2]
    mp = [synth.make_problem(max(m, 8), 0.3, 90 + i) for i, m in enumerate(sizes)]
    mixed = [(np.ascontiguousarray(p.src.T[:m]), np.ascontiguousarray(p.dst.T[:m])) for p, m in zip(mp, sizes)]
    bs2 = b2batch.BatchSolver(_params(tpp, mp[0].noise_bound))
    bs2.solve_packed_async(mixed); bs2.sync()
    want2 = bs2.solutions_array().copy()
    buf = b2batch.PinnedBuffer(2 * 24 * sum(sizes))
    off = 0
    pinned = []
    for a, b in mixed:
        va = buf.array(a.shape, offset=off); off += a.nbytes
        vb = buf.array(b.shape, offset=off); off += b.nbytes
        b2batch.PinnedBuffer.fill(va, a); vb[...] = b
        pinned.append((va, vb))
    bs2.solve_pinned_async(pinned); bs2.sync()
    assert _same_solutions(bs2.solutions_array(), want2)
    del pinned, buf


def _tf32_rna(x32):
    """cvt.rna.tf32.f32: round to 10 explicit mantissa bits, ties away from zero."""
    b = x32.astype(np.float32).view(np.uint32).astype(np.uint64)
    b = (b + 0x1000) & 0xFFFFE000
    return b.astype(np.uint32).view(np.float32)


@pytest.mark.parametrize("n,ratio,seed,blocks", [(1000, 0.5, 41, ((0, 0), (0, 5), (3, 2), (7, 0))), (5000, 0.95, 5000, ((0, 0), (5, 33), (39, 0)))])
def test_tensor_core_values_within_assumed_error(n, ratio, seed, blocks):
    """The band of the tensor-core filter ASSUMES |tcgen05 result - exact sum of products| <= 2^-20 * sum |products| and bounds the
    total error of Dm~ / L~ by E_D / E_L (graph_tc.cu).  Measure both on the device."""
    import teaserpp_python as tpp
    pb = synth.make_problem(n, ratio, seed)
    s = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound))
    s.solve(pb.src, pb.dst)
    worst_tc = worst_d = worst_l = 0.0
    for bi, unit in blocks:
        dm, lv, c = s._tc_values(bi, unit)
        assert c["use32"] == 1
        kappa, ctr = c["kappa"], c["ctr"]
        i0 = 128 * bi
        j0 = 128 * (bi + unit // 2) + 64 * (unit % 2)
        ri = np.arange(i0, min(i0 + 128, n)); cj = np.arange(j0, min(j0 + 64, n))
        # exact (fp64) scaled quantities
        S = (pb.src - ctr[:3, None]) * kappa; D = (pb.dst - ctr[3:, None]) * kappa
        A = ((S[:, ri, None] - S[:, None, cj]) ** 2).sum(0); B = ((D[:, ri, None] - D[:, None, cj]) ** 2).sum(0)
        Dm, L = A - B, 1.0 - 2.0 * (A + B)
        # the bounds the kernel uses
        he = np.concatenate([np.max(np.abs(pb.src - ctr[:3, None]), 1), np.max(np.abs(pb.dst - ctr[3:, None]), 1)]) * kappa * (1 + 1e-6)
        Pi = float((he ** 2).sum()); w = 2.0 ** -22
        E_L, E_D = w * (71.8 * Pi + 4.3), 45.9 * w * Pi
        worst_d = max(worst_d, float(np.abs(dm[: len(ri), : len(cj)] - Dm).max()) / E_D)
        worst_l = max(worst_l, float(np.abs(lv[: len(ri), : len(cj)] - L).max()) / E_L)
        # the tensor-core accumulation alone: operands as the device builds them (22-bit copies, 3xTF32 split), products summed in fp64
        def split(X):
            x32 = X.astype(np.float32); hi = _tf32_rna(x32); lo = _tf32_rna(x32 - hi)
            v = (hi + lo).astype(np.float32); h2 = _tf32_rna(v)
            return v.astype(np.float64), h2.astype(np.float64), (v - h2).astype(np.float64)
        sv, sh, sl = split(S); dv, dh, dl = split(D)
        rs, rd = (sv ** 2).sum(0), (dv ** 2).sum(0)
        p = (0.5 - 2.0 * (rs + rd)).astype(np.float32).astype(np.float64); m = (rs - rd).astype(np.float32).astype(np.float64)
        def tri(h1, l1, h2, l2):  # hi.hi + hi.lo + lo.hi summed over the three coordinates, and the sum of |products|
            t = h1[:, ri, None] * h2[:, None, cj] + h1[:, ri, None] * l2[:, None, cj] + l1[:, ri, None] * h2[:, None, cj]
            a = np.abs(h1[:, ri, None] * h2[:, None, cj]) + np.abs(h1[:, ri, None] * l2[:, None, cj]) + np.abs(l1[:, ri, None] * h2[:, None, cj])
            return t.sum(0), a.sum(0)
        ts, as_ = tri(sh, sl, sh, sl); td, ad = tri(dh, dl, dh, dl)
        L_ops = p[ri, None] + p[None, cj] + 4 * ts + 4 * td
        D_ops = m[ri, None] + m[None, cj] - 2 * ts + 2 * td
        L_abs = np.abs(p[ri, None]) + np.abs(p[None, cj]) + 4 * as_ + 4 * ad
        D_abs = np.abs(m[ri, None]) + np.abs(m[None, cj]) + 2 * as_ + 2 * ad
        worst_tc = max(worst_tc, float((np.abs(lv[: len(ri), : len(cj)] - L_ops) / L_abs).max()), float((np.abs(dm[: len(ri), : len(cj)] - D_ops) / D_abs).max()))
    print(f"tensor-core accumulation error / sum|products|: 2^{np.log2(max(worst_tc, 1e-30)):.2f} (assumed 2^-20); "
          f"|Dm~-Dm| / E_D = {worst_d:.3f}, |L~-L| / E_L = {worst_l:.3f}")
    assert worst_tc <= 2.0 ** -21, "tcgen05 accumulation error exceeds half of the assumed bound"
    assert worst_d <= 0.5 and worst_l <= 0.5


def test_tensor_core_filter_list_overflow_falls_back(oracle_mod):
    """A band so wide that the ambiguous-pair list overflows: the problem is redone by the fp64 filter, bits unchanged."""
    import teaserpp_python as tpp
    pb = synth.make_problem(1500, 0.7, 43)
    ref = oracle_mod.tim_graph(pb.src, pb.dst, pb.noise_bound)
    s = tpp.RobustRegistrationSolver(_params(tpp, pb.noise_bound, graph_filter_tau=1e-2))
    s.solve(pb.src, pb.dst)
    assert np.array_equal(s._adjacency_bits(), ref)
    assert s.stats()["graph_slow_pairs"] > 32768


@pytest.mark.parametrize("n", [2, 3, 63, 64, 65, 127, 128, 129, 191, 192, 193, 257, 640, 1153])
def test_tensor_core_filter_tile_edges_and_fused_degrees(oracle_mod, n):
    """Sizes around the 64-column unit and 128-row tile edges: bits equal the oracle's, and the degrees the filter kernel
    accumulates on the fly (corrected by the fix-up of the listed pairs) equal the popcounts of the final rows."""
    import teaserpp_python as tpp
    rng = np.random.default_rng(1000 + n)
    src = rng.uniform(-1, 1, (3, n))
    dst = synth.random_rotation(rng) @ src + rng.uniform(-1, 1, (3, 1)) + rng.normal(0, 0.01, (3, n))
    k = n // 2
    dst[:, :k] = rng.uniform(-2, 2, (3, k))
    nb = 0.05
    ref = oracle_mod.tim_graph(src, dst, nb)
    s = tpp.RobustRegistrationSolver(_params(tpp, nb, inlier_selection_mode=tpp.RobustRegistrationSolver.INLIER_SELECTION_MODE.NONE))
    s.solve(src, dst)
    bits = s._adjacency_bits()
    assert np.array_equal(bits, ref)
    assert s.stats()["n_edges"] == int(np.unpackbits(ref.view(np.uint8)).sum()) // 2


@pytest.mark.parametrize("n,ratio", [(300, 0.5), (3000, 0.9), (5000, 0.95)])
def test_tile_sharded_solve_world_of_one(oracle_mod, n, ratio):
    """b2r_solve_sharded with a world of one rank: the PEERS form of the tensor-core filter (item striding, bit-only fix-up, degrees
    recomputed from the matrix, READY / DONE flag words) must give exactly the ordinary solve.  Two and more ranks: bench.py config5."""
    import torch
    import teaserpp_python as tpp
    from poseestimator_b200 import batch as b2batch
    pb = synth.make_problem(n, ratio, seed=31 + n)
    ref = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=0, rotation_cost_threshold=1e-12)
    bs = b2batch.BatchSolver(_params(tpp, pb.noise_bound))
    src = torch.from_numpy(np.ascontiguousarray(pb.src.T)).cuda(); dst = torch.from_numpy(np.ascontiguousarray(pb.dst.T)).cuda()
    torch.cuda.synchronize()
    bs.shard_attach(0, 1, [bs.shard_export(n)])
    for _ in range(3):        # epochs
        sol = bs.solve_sharded(src.data_ptr(), dst.data_ptr(), n)
        assert not bs.shard_fell_back()
        assert np.array_equal(bs.adjacency(0, n), ref.adjacency)
        assert bs.max_clique(0).tolist() == ref.clique.tolist()
        assert sol.valid and synth.rotation_error_rad(sol.rotation, ref.rotation) < 1e-5
        assert np.linalg.norm(sol.translation - ref.translation) < 1e-6 * pb.scene_scale
    # an ill-conditioned problem (extent / beta beyond the tensor-core filter's gate): every rank falls back to the ordinary path
    far = synth.make_problem(400, 0.5, seed=5)
    bs2 = b2batch.BatchSolver(_params(tpp, far.noise_bound * 1e-5))
    s2 = torch.from_numpy(np.ascontiguousarray(far.src.T)).cuda(); d2 = torch.from_numpy(np.ascontiguousarray(far.dst.T)).cuda()
    torch.cuda.synchronize()
    bs2.shard_attach(0, 1, [bs2.shard_export(400)])
    bs2.solve_sharded(s2.data_ptr(), d2.data_ptr(), 400)
    assert bs2.shard_fell_back()
    ref2 = oracle_mod.solve(far.src, far.dst, noise_bound=far.noise_bound * 1e-5, estimate_scaling=0, rotation_cost_threshold=1e-12)
    assert np.array_equal(bs2.adjacency(0, 400), ref2.adjacency)
