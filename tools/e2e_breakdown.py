"""Where the end-to-end step goes: host time of the enqueue call (pack into pinned staging + launches), of sync and of the result
getter, for the config-3 batch; plus the pipelined two-handle throughput of bench.py's e2e leg."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from poseestimator_b200 import batch as b2batch

B = 64
pbs = bench.make_problems(B, 0)
packed = [(np.ascontiguousarray(p.src.T), np.ascontiguousarray(p.dst.T)) for p in pbs]
nb = pbs[0].noise_bound
pipes = [b2batch.BatchSolver(bench.solver_params(nb)), b2batch.BatchSolver(bench.solver_params(nb))]
for sv in pipes:
    sv.solve_packed_async(packed); sv.sync()
t = {"enqueue": [], "sync": [], "get": []}
sv = pipes[0]
for _ in range(20):
    t0 = time.perf_counter(); sv.solve_packed_async(packed); t1 = time.perf_counter(); sv.sync(); t2 = time.perf_counter(); a = sv.solutions_array(); t3 = time.perf_counter()
    t["enqueue"].append(t1 - t0); t["sync"].append(t2 - t1); t["get"].append(t3 - t2)
print("serial, one handle (ms):", {k: round(1e3 * float(np.median(v)), 3) for k, v in t.items()})
K = 40
t0 = time.perf_counter()
pipes[0].solve_packed_async(packed)
enq = []
for k in range(K):
    cur, nxt = pipes[k & 1], pipes[(k + 1) & 1]
    if k + 1 < K:
        a0 = time.perf_counter(); nxt.solve_packed_async(packed); enq.append(time.perf_counter() - a0)
    cur.sync(); a = cur.solutions_array()
dt = time.perf_counter() - t0
print(f"pipelined: {1e3 * dt / K:.3f} ms per step, {B * K / dt:.0f} reg/s; enqueue call median {1e3 * float(np.median(enq)):.3f} ms")
