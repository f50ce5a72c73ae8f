import sys, os, json
sys.path.insert(0, '/root/repo')
import numpy as np
import bench
from poseestimator_b200 import batch as b2batch, synth
for name, pbs in (("config4", [synth.config4_detection(0, d) for d in range(64)]), ("config3", [synth.make_problem(5000, 0.95, 5000 + i) for i in range(64)]),
                  ("config5", [synth.make_problem(20000, 0.95, 20000)])):
    packed = [(np.ascontiguousarray(p.src.T), np.ascontiguousarray(p.dst.T)) for p in pbs]
    sv = b2batch.BatchSolver(bench.solver_params(pbs[0].noise_bound), profiling=True)
    for _ in range(3):
        sv.solve_packed_async(packed); sv.sync()
    worst = max(range(len(pbs)), key=lambda i: sv.tail_cycles(i)["translation"])
    c = sv.tail_cycles(worst)
    print(name, "problem", worst, "n", pbs[worst].src.shape[1], "inliers", len(pbs[worst].inliers), json.dumps(c), sv.timings_ms()["tail_fused"])
    st = sv.stats(worst)
    print(name, "marks 20..29", {k: c["marks"].get(k) for k in range(20, 30)}, "gnc_iterations", st["gnc_iterations"], "rot tims", st["n_rot_tims"])
