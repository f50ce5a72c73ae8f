#!/usr/bin/env python
"""Regenerates tests/golden/teaser_*.npz.

PARITY UNPINNED: the reference ships no golden vectors and its native dependencies (TEASER++, Open3D) are not
importable here, so these fixtures are OUTPUTS OF THE ORACLE (oracle/teaser_oracle.cpp) on seeded synthetic inputs
(SURVEY.md 8c "golden vectors to create").  They pin the oracle against drift and give the GPU tests a second,
file-based target; they are not outputs of upstream TEASER++.

    python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle  # noqa: E402
from poseestimator_b200 import synth  # noqa: E402

CASES = [  # name, n, outlier ratio, seed, extra oracle params
    ("cfg1_n1000_o50", 1000, 0.5, 1000, {}),
    ("n400_o60", 400, 0.6, 7, {}),
    ("n129_o60", 129, 0.6, 2, {}),
    ("n300_noisy_gnc", 300, 0.3, 11, {"sigma": 0.02, "noise_bound": 0.03}),
    ("n200_scale", 200, 0.5, 21, {"estimate_scaling": 1, "scale": 1.7}),
    ("n200_complete", 200, 0.7, 31, {"rotation_tim_graph": 1}),
]


def adjacency_sha(adj: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(adj, np.uint32).tobytes()).hexdigest()


def hand_case():
    """N=8 with a known 5-clique {0,2,3,5,7}: those five are an exact rigid copy, the rest are far off."""
    rng = np.random.default_rng(8)
    src = rng.uniform(-1, 1, (8, 3))
    R = synth.random_rotation(rng)
    t = np.array([0.3, -0.2, 0.5])
    dst = src @ R.T + t
    for i in (1, 4, 6):
        dst[i] = rng.uniform(5, 9, 3) * (1 if i != 4 else -1)
    return src.T, dst.T, R, t


def main():
    oracle.build()
    for name, n, ratio, seed, extra in CASES:
        extra = dict(extra)
        sigma = extra.pop("sigma", 0.01)
        scale = extra.pop("scale", 1.0)
        pb = synth.make_problem(n, ratio, seed, sigma=sigma, scale=scale)
        nb = extra.pop("noise_bound", pb.noise_bound)
        kw = dict(noise_bound=nb, estimate_scaling=0, rotation_cost_threshold=1e-12)
        kw.update(extra)
        r = oracle.solve(pb.src, pb.dst, **kw)
        np.savez_compressed(
            os.path.join(HERE, f"teaser_{name}.npz"),
            n=n, ratio=ratio, seed=seed, sigma=sigma, gen_scale=scale, noise_bound=nb,
            params=np.array([kw.get("estimate_scaling", 0), kw.get("rotation_tim_graph", 0)]),
            adjacency_sha256=adjacency_sha(r.adjacency), n_edges=r.n_edges, clique=r.clique,
            scale=r.scale, rotation=r.rotation, translation=r.translation, valid=r.valid,
            rot_inliers=r.rotation_inliers, trans_inliers=r.translation_inliers, gnc_iters=r.gnc_iters)
        print(name, "edges", r.n_edges, "clique", r.clique.size, "valid", r.valid, "scale", r.scale)
    s, d, R, t = hand_case()
    r = oracle.solve(s, d, noise_bound=1e-3, estimate_scaling=0, rotation_cost_threshold=1e-12)
    assert r.clique.tolist() == [0, 2, 3, 5, 7], r.clique
    np.savez_compressed(os.path.join(HERE, "teaser_hand_n8.npz"), src=s, dst=d, R=R, t=t, noise_bound=1e-3,
                        adjacency=r.adjacency, clique=r.clique, rotation=r.rotation, translation=r.translation)
    print("hand_n8 clique", r.clique.tolist())


if __name__ == "__main__":
    main()
