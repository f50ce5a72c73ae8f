"""Prints per-stage device times for a batch of config-3 problems (quick A/B tool for kernel changes)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from poseestimator_b200 import synth, batch
import teaserpp_python as tpp

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
N = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
ratio = float(sys.argv[3]) if len(sys.argv) > 3 else 0.95
pbs = [synth.make_problem(N, ratio, 5000 + i) for i in range(B)]
p = tpp.RobustRegistrationSolver.Params(); p.noise_bound = pbs[0].noise_bound; p.estimate_scaling = False; p.rotation_cost_threshold = 1e-12
bs = batch.BatchSolver(p, profiling=True)
packed = [(np.ascontiguousarray(q.src.T), np.ascontiguousarray(q.dst.T)) for q in pbs]
acc = {}
R = 10
for it in range(R + 2):
    bs.solve_packed_async(packed); bs.sync()
    if it >= 2:
        for k, v in bs.timings_ms().items(): acc[k] = acc.get(k, 0) + v / R
print(f"B={B} N={N} ratio={ratio}", {k: round(v, 4) for k, v in acc.items()})
print(" stats0", {k: v for k, v in bs.stats(0).items() if k in ("n_edges", "heuristic_lb", "n_pruned", "omega", "clique_size", "bnb_nodes", "graph_slow_pairs", "gnc_iterations")})
if bs.stats(0).get("fused_tail"):
    for i in (0, 1, B - 1):
        print(" tail cycles", i, bs.tail_cycles(i))
