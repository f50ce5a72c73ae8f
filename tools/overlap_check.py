"""Device-resident inputs, K batches through 1, 2 or 3 handles (streams) in flight: does the latency-bound tail of one batch
(clique / rotation / translation kernels, 64 CTAs each) overlap the graph kernel of the next?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from poseestimator_b200 import batch as b2batch

B, K = 64, 60
pbs = bench.make_problems(B, 0)
nb = pbs[0].noise_bound
d_src = [torch.from_numpy(np.ascontiguousarray(p.src.T)).cuda() for p in pbs]
d_dst = [torch.from_numpy(np.ascontiguousarray(p.dst.T)).cuda() for p in pbs]
n_list = [p.src.shape[1] for p in pbs]
sp = [t.data_ptr() for t in d_src]; dp = [t.data_ptr() for t in d_dst]
for NP in (1, 2, 3):
    pipes = [b2batch.BatchSolver(bench.solver_params(nb)) for _ in range(NP)]
    for sv in pipes:
        sv.solve_device_async(n_list, sp, dp); sv.sync()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(min(NP - 1, K)):
        pipes[k % NP].solve_device_async(n_list, sp, dp)
    for k in range(K):
        if k + NP - 1 < K:
            pipes[(k + NP - 1) % NP].solve_device_async(n_list, sp, dp)
        pipes[k % NP].sync()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"handles in flight {NP}: {1e3 * dt / K:.3f} ms per batch, {B * K / dt:.0f} reg/s")

# the same with HOST inputs (the e2e leg of bench.py): pack into pinned staging + one H2D copy per batch
packed = [(np.ascontiguousarray(p.src.T), np.ascontiguousarray(p.dst.T)) for p in pbs]
for NP in (1, 2, 3, 4):
    pipes = [b2batch.BatchSolver(bench.solver_params(nb)) for _ in range(NP)]
    for sv in pipes:
        sv.solve_packed_async(packed); sv.sync()
    torch.cuda.synchronize()
    enq = []; syn = []; get = []
    t0 = time.perf_counter()
    for k in range(min(NP - 1, K)):
        pipes[k % NP].solve_packed_async(packed)
    for k in range(K):
        if k + NP - 1 < K:
            a = time.perf_counter(); pipes[(k + NP - 1) % NP].solve_packed_async(packed); enq.append(time.perf_counter() - a)
        a = time.perf_counter(); pipes[k % NP].sync(); syn.append(time.perf_counter() - a)
        a = time.perf_counter(); sols = pipes[k % NP].solutions_array(); get.append(time.perf_counter() - a)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"host inputs, handles in flight {NP}: {1e3 * dt / K:.3f} ms per batch, {B * K / dt:.0f} reg/s; enqueue median {1e3 * np.median(enq):.3f} / mean {1e3 * np.mean(enq):.3f} ms, sync wait mean {1e3 * np.mean(syn):.3f} ms, get mean {1e3 * np.mean(get):.3f} ms")

# per-stage device times inside the 3-handle host-input pipeline (events on each handle's stream; stages of different handles overlap)
NP = 3
pipes = [b2batch.BatchSolver(bench.solver_params(nb), profiling=True) for _ in range(NP)]
for sv in pipes:
    sv.solve_packed_async(packed); sv.sync()
acc = {}
for k in range(min(NP - 1, K)):
    pipes[k % NP].solve_packed_async(packed)
for k in range(K):
    if k + NP - 1 < K:
        pipes[(k + NP - 1) % NP].solve_packed_async(packed)
    pipes[k % NP].sync()
    for a, b in pipes[k % NP].timings_ms().items():
        acc[a] = acc.get(a, 0.0) + b / K
print("3 handles, host inputs, mean stage ms per batch (overlapped):", {a: round(b, 3) for a, b in acc.items()})
