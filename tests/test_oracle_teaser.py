"""Validation of the CPU oracle itself (SURVEY.md section 8c "oracle self-validation").

The reference pins nothing (no tests / golden vectors), so the oracle is checked by
 * ground-truth recovery on seeded synthetic problems,
 * networkx exact max clique + enumeration of all maximum cliques (canonical tie-break),
 * numpy SVD, a brute-force numpy restatement of adaptive voting and of the edge test,
 * invariances (rigid transform of dst, symmetric / zero-diagonal adjacency).
"""
import itertools

import networkx as nx
import numpy as np
import pytest

from poseestimator_b200 import synth


def unpack(adj, n):
    bits = np.unpackbits(adj.view(np.uint8), axis=1, bitorder="little")[:, :n]
    return bits.astype(bool)


def numpy_edge_test(src, dst, noise_bound, cbar2=1.0):
    """Vectorised fp64 restatement of A.4-2, association (x^2+y^2)+z^2, no FMA in numpy."""
    s, d = src.T, dst.T
    n = s.shape[0]
    i, j = np.triu_indices(n, 1)
    ds, dd = s[j] - s[i], d[j] - d[i]
    a = np.sqrt((ds[:, 0] * ds[:, 0] + ds[:, 1] * ds[:, 1]) + ds[:, 2] * ds[:, 2])
    b = np.sqrt((dd[:, 0] * dd[:, 0] + dd[:, 1] * dd[:, 1]) + dd[:, 2] * dd[:, 2])
    beta = 2 * noise_bound * np.sqrt(cbar2)
    ok = np.abs(a - b) <= beta
    A = np.zeros((n, n), bool)
    A[i[ok], j[ok]] = True
    return A | A.T


@pytest.mark.parametrize("n,ratio,seed", [(60, 0.5, 1), (300, 0.8, 2), (1000, 0.5, 1000)])
def test_ground_truth_recovery_and_edge_test(oracle_mod, n, ratio, seed):
    pb = synth.make_problem(n, ratio, seed)
    r = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=0, rotation_cost_threshold=1e-12)
    assert r.valid
    A = unpack(r.adjacency, n)
    assert np.array_equal(A, A.T) and not A.diagonal().any()
    assert np.array_equal(A, numpy_edge_test(pb.src, pb.dst, pb.noise_bound))
    assert r.n_edges == A.sum() // 2
    # true inliers are mutually consistent with overwhelming probability -> contained in the clique
    assert set(pb.inliers.tolist()) <= set(r.clique.tolist())
    assert synth.rotation_error_rad(r.rotation, pb.R) < 0.02
    assert np.linalg.norm(r.translation - pb.t) < 0.02
    assert abs(np.linalg.det(r.rotation) - 1) < 1e-12
    # streaming variant gives the identical graph and solution
    r2 = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=0, rotation_cost_threshold=1e-12,
                          materialize_tims=0)
    assert np.array_equal(r.adjacency, r2.adjacency) and np.array_equal(r.clique, r2.clique)
    assert np.array_equal(r.adjacency, oracle_mod.tim_graph(pb.src, pb.dst, pb.noise_bound))


def test_scale_mask_order_is_tim_order(oracle_mod):
    pb = synth.make_problem(40, 0.5, 5)
    r = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=0, want_scale_mask=True)
    A = unpack(r.adjacency, 40)
    i, j = np.triu_indices(40, 1)  # row-major upper triangle == A.4-1 TIM order
    assert np.array_equal(r.scale_mask.astype(bool), A[i, j])


def test_adjacency_invariant_under_rigid_motion_of_dst(oracle_mod):
    pb = synth.make_problem(200, 0.7, 11)
    rng = np.random.default_rng(0)
    R2 = synth.random_rotation(rng)
    a1 = unpack(oracle_mod.tim_graph(pb.src, pb.dst, pb.noise_bound), 200)
    a2 = unpack(oracle_mod.tim_graph(pb.src, R2 @ pb.dst + rng.normal(size=(3, 1)), pb.noise_bound), 200)
    # exact-arithmetic invariant; in fp64 only pairs within ~1e-15 of the bound may flip
    assert (a1 != a2).sum() <= 2


def canonical_max_clique_nx(A):
    G = nx.from_numpy_array(A.astype(int))
    best, size = None, 0
    for c in nx.find_cliques(G):
        c = sorted(c)
        if len(c) > size or (len(c) == size and c < best):
            best, size = c, len(c)
    return best


def bits_from_dense(A):
    n = A.shape[0]
    W = (n + 31) // 32
    pad = np.zeros((n, W * 32), np.uint8)
    pad[:, :n] = A
    return np.packbits(pad, axis=1, bitorder="little").view(np.uint32).reshape(n, W)


@pytest.mark.parametrize("seed", range(12))
def test_max_clique_matches_networkx_with_ties(oracle_mod, seed):
    rng = np.random.default_rng(seed)
    n = int(rng.integers(5, 90))
    p = rng.uniform(0.1, 0.8)
    A = rng.random((n, n)) < p
    A = np.triu(A, 1)
    A = A | A.T
    want = canonical_max_clique_nx(A)
    got = oracle_mod.max_clique_bits(bits_from_dense(A)).tolist()
    assert got == want, (n, p)


def test_max_clique_planted_and_heuristic_modes(oracle_mod):
    rng = np.random.default_rng(3)
    n = 200
    A = rng.random((n, n)) < 0.05
    A = np.triu(A, 1)
    A = A | A.T
    planted = np.sort(rng.permutation(n)[:25])
    A[np.ix_(planted, planted)] = True
    np.fill_diagonal(A, False)
    bits = bits_from_dense(A)
    assert oracle_mod.max_clique_bits(bits, mode=0).tolist() == planted.tolist()
    heu = oracle_mod.max_clique_bits(bits, mode=1).tolist()
    assert all(A[a, b] for a, b in itertools.combinations(heu, 2)) and len(heu) == 25
    assert oracle_mod.max_clique_bits(bits, mode=3).tolist() == list(range(n))  # NONE -> all (A.5-7)
    # KCORE_HEU with a threshold the max core exceeds -> vertices of the max core (A.5-3)
    kc = oracle_mod.max_clique_bits(bits, mode=2, kcore_thr=0.05).tolist()
    G = nx.from_numpy_array(A.astype(int))
    cn = nx.core_number(G)
    mc = max(cn.values())
    assert kc == sorted(v for v, c in cn.items() if c >= mc)


def test_hand_checkable_n8_known_5_clique(oracle_mod):
    A = np.zeros((8, 8), bool)
    for a, b in itertools.combinations([1, 3, 4, 6, 7], 2):
        A[a, b] = A[b, a] = True
    for a, b in [(0, 1), (0, 2), (2, 5), (5, 6), (0, 5)]:
        A[a, b] = A[b, a] = True
    assert oracle_mod.max_clique_bits(bits_from_dense(A)).tolist() == [1, 3, 4, 6, 7]


@pytest.mark.parametrize("seed", range(6))
def test_svd3_against_numpy(oracle_mod, seed):
    rng = np.random.default_rng(seed)
    A = rng.normal(size=(3, 3))
    if seed == 4:
        A[:, 2] = A[:, 0] * 2  # rank deficient
    if seed == 5:
        A = np.outer(rng.normal(size=3), rng.normal(size=3))  # rank one
    U, S, V = oracle_mod.svd3(A)
    assert np.allclose(U @ np.diag(S) @ V.T, A, atol=1e-12)
    assert np.allclose(U.T @ U, np.eye(3), atol=1e-10) and np.allclose(V.T @ V, np.eye(3), atol=1e-10)
    assert np.allclose(S, np.linalg.svd(A, compute_uv=False), atol=1e-12)


def tls_brute(X, r):
    """Direct restatement of A.7 with O(M^2) consensus sets (no running sums)."""
    M = len(X)
    ends = sorted([(X[i] - r[i], 0, i) for i in range(M)] + [(X[i] + r[i], 1, i) for i in range(M)])
    inside = np.zeros(M, bool)
    best = (np.inf, None)
    w = 1 / r ** 2
    for v, kind, i in ends:
        inside[i] = kind == 0
        if not inside.any():
            continue
        xh = (w[inside] * X[inside]).sum() / w[inside].sum()
        cost = ((X[inside] - xh) ** 2).sum() + r[~inside].sum()
        if cost < best[0] - 1e-12:
            best = (cost, xh)
    return best[1]


@pytest.mark.parametrize("seed", range(5))
def test_tls_estimate_against_bruteforce(oracle_mod, seed):
    rng = np.random.default_rng(seed)
    M = 40
    X = np.concatenate([rng.normal(1.0, 0.01, 25), rng.uniform(-3, 3, 15)])
    r = np.full(M, 0.05) if seed % 2 == 0 else rng.uniform(0.02, 0.08, M)
    est, inl = oracle_mod.tls_estimate(X, r)
    assert abs(est - tls_brute(X, r)) < 1e-9
    assert np.array_equal(inl, np.abs(X - est) <= r)
    assert abs(est - 1.0) < 0.02


def test_gnc_tls_rejects_outliers(oracle_mod):
    rng = np.random.default_rng(7)
    R = synth.random_rotation(rng)
    X = rng.normal(size=(3, 60))
    Y = R @ X + rng.normal(0, 0.01, size=(3, 60))
    Y[:, :20] = rng.normal(size=(3, 20)) * 2
    Rh, inl, w, it, cost = oracle_mod.gnc_tls(X, Y, noise_bound=0.05, thr=1e-12)
    assert synth.rotation_error_rad(Rh, R) < 0.02
    assert inl[20:].all() and inl[:20].sum() <= 2 and 1 < it <= 100


def test_degenerate_cases(oracle_mod):
    rng = np.random.default_rng(1)
    # all outliers, tiny noise bound -> no edges -> clique <= 1 -> valid False, identity (A.5-6)
    src = rng.uniform(-1, 1, (3, 30)); dst = rng.uniform(-1, 1, (3, 30))
    r = oracle_mod.solve(src, dst, noise_bound=1e-9, estimate_scaling=0)
    assert not r.valid and len(r.clique) <= 1 and np.array_equal(r.rotation, np.eye(3))
    # zero noise -> mu <= 0 early exit (A.6-3c), exact recovery
    R = synth.random_rotation(rng); t = np.array([0.3, -0.2, 0.1])
    dst = R @ src + t[:, None]
    r = oracle_mod.solve(src, dst, noise_bound=0.01, estimate_scaling=0)
    assert r.valid and r.gnc_iters == 0 and len(r.clique) == 30
    assert synth.rotation_error_rad(r.rotation, R) < 1e-7 and np.allclose(r.translation, t, atol=1e-7)
    # duplicate points (TIM norm 0) and N = 2, 3
    src2 = src.copy(); src2[:, 1] = src2[:, 0]
    dst2 = R @ src2 + t[:, None]
    r = oracle_mod.solve(src2, dst2, noise_bound=0.01, estimate_scaling=0)
    assert r.valid and len(r.clique) == 30
    for n in (2, 3):
        r = oracle_mod.solve(src[:, :n], dst[:, :n], noise_bound=0.01, estimate_scaling=0)
        assert r.valid and len(r.clique) == n


def test_scale_estimation(oracle_mod):
    pb = synth.make_problem(120, 0.4, 21, scale=1.7)
    r = oracle_mod.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=1)
    assert r.valid and abs(r.scale - 1.7) < 0.02
    assert synth.rotation_error_rad(r.rotation, pb.R) < 0.03
    assert np.linalg.norm(r.translation - pb.t) < 0.05
