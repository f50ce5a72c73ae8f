"""Which config-3 problems does the fused tail hand to the multi-kernel chain, and why?  Seeds of all eight bench ranks."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from poseestimator_b200 import batch as b2batch
out = []
for r in range(8):
    pbs = bench.make_problems(64, r)
    sv = b2batch.BatchSolver(bench.solver_params(pbs[0].noise_bound), device=0, profiling=True)
    sv.solve_packed_async([(np.ascontiguousarray(p.src.T), np.ascontiguousarray(p.dst.T)) for p in pbs]); sv.sync()
    for i in range(64):
        st = sv.stats(i)
        if st["fused_tail"] != 1:
            out.append({"rank": r, "i": i, **{k: st[k] for k in ("heuristic_lb", "n_pruned", "omega", "clique_size", "bnb_nodes", "graph_slow_pairs")}})
print(json.dumps({"handed_over": len(out), "of": 512, "cases": out}))
