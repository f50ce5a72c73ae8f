"""Committed golden fixtures (tests/golden/*.npz, made by tests/golden/make_golden.py from the oracle -- the reference
ships none, parity unpinned).  CPU: the oracle still reproduces them.  GPU: the CUDA path reproduces them."""
import glob
import hashlib
import os

import numpy as np
import pytest

from poseestimator_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
FILES = sorted(f for f in glob.glob(os.path.join(HERE, "golden", "teaser_*.npz")) if "hand" not in f)


def _sha(adj):
    return hashlib.sha256(np.ascontiguousarray(adj, np.uint32).tobytes()).hexdigest()


def _problem(g):
    return synth.make_problem(int(g["n"]), float(g["ratio"]), int(g["seed"]), sigma=float(g["sigma"]), scale=float(g["gen_scale"]))


def test_fixtures_exist():
    assert len(FILES) >= 6


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[7:-4] for f in FILES])
def test_oracle_reproduces_golden(oracle_mod, path):
    g = np.load(path)
    pb = _problem(g)
    r = oracle_mod.solve(pb.src, pb.dst, noise_bound=float(g["noise_bound"]), estimate_scaling=int(g["params"][0]),
                         rotation_tim_graph=int(g["params"][1]), rotation_cost_threshold=1e-12)
    assert _sha(r.adjacency) == str(g["adjacency_sha256"])
    assert r.n_edges == int(g["n_edges"])
    assert r.clique.tolist() == g["clique"].tolist()
    assert np.array_equal(r.rotation, g["rotation"]) and np.array_equal(r.translation, g["translation"])
    assert r.scale == float(g["scale"])
    assert r.rotation_inliers.tolist() == g["rot_inliers"].tolist()
    assert r.translation_inliers.tolist() == g["trans_inliers"].tolist()


def test_oracle_hand_case(oracle_mod):
    g = np.load(os.path.join(HERE, "golden", "teaser_hand_n8.npz"))
    r = oracle_mod.solve(g["src"], g["dst"], noise_bound=float(g["noise_bound"]), estimate_scaling=0, rotation_cost_threshold=1e-12)
    assert r.clique.tolist() == [0, 2, 3, 5, 7]
    assert np.array_equal(r.adjacency, g["adjacency"])
    assert synth.rotation_error_rad(r.rotation, g["R"]) < 1e-9
    assert np.linalg.norm(r.translation - g["t"]) < 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f)[7:-4] for f in FILES])
def test_gpu_reproduces_golden(path):
    import teaserpp_python as tpp
    g = np.load(path)
    pb = _problem(g)
    p = tpp.RobustRegistrationSolver.Params()
    p.noise_bound = float(g["noise_bound"])
    p.estimate_scaling = bool(int(g["params"][0]))
    p.rotation_tim_graph = tpp.RobustRegistrationSolver.INLIER_GRAPH_FORMULATION(int(g["params"][1]))
    p.rotation_cost_threshold = 1e-12
    s = tpp.RobustRegistrationSolver(p)
    sol = s.solve(pb.src, pb.dst)
    assert _sha(s._adjacency_bits()) == str(g["adjacency_sha256"])
    assert s.getInlierMaxClique() == g["clique"].tolist()
    assert sol.valid == bool(g["valid"])
    assert synth.rotation_error_rad(sol.rotation, g["rotation"]) < 1e-5
    assert np.linalg.norm(sol.translation - g["translation"]) < 1e-6 * pb.scene_scale * max(1.0, float(g["scale"]))
    assert abs(sol.scale - float(g["scale"])) < 1e-9 * max(1.0, float(g["scale"]))
    assert s.getRotationInliers() == g["rot_inliers"].tolist()
    assert s.getTranslationInliers() == g["trans_inliers"].tolist()


@pytest.mark.gpu
def test_gpu_hand_case():
    import teaserpp_python as tpp
    g = np.load(os.path.join(HERE, "golden", "teaser_hand_n8.npz"))
    p = tpp.RobustRegistrationSolver.Params()
    p.noise_bound = float(g["noise_bound"])
    p.estimate_scaling = False
    s = tpp.RobustRegistrationSolver(p)
    sol = s.solve(g["src"], g["dst"])
    assert s.getInlierMaxClique() == [0, 2, 3, 5, 7]
    assert np.array_equal(s._adjacency_bits(), g["adjacency"])
    assert synth.rotation_error_rad(sol.rotation, g["R"]) < 1e-9
    assert np.linalg.norm(sol.translation - g["t"]) < 1e-9
