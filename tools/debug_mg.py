import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import torch.distributed as dist
import bench
from poseestimator_b200 import batch as b2batch
rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
use_dist = world > 1 and os.environ.get("NO_DIST") != "1"
if use_dist:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
pbs = bench.make_problems(16, rank)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
solver = b2batch.BatchSolver(bench.solver_params(pbs[0].noise_bound), device=local, stream=stream.cuda_stream, profiling=True)
packed = [(np.ascontiguousarray(p.src.T), np.ascontiguousarray(p.dst.T)) for p in pbs]
d_src = [torch.from_numpy(a).cuda() for a, _ in packed]; d_dst = [torch.from_numpy(b).cuda() for _, b in packed]
sp = [t.data_ptr() for t in d_src]; dp = [t.data_ptr() for t in d_dst]; nl = [5000] * 16
for _ in range(3):
    solver.solve_device_async(nl, sp, dp); solver.sync()
if use_dist:
    dist.barrier()
torch.cuda.synchronize()
for it in range(6):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record(stream)
    t1 = time.perf_counter(); solver.solve_device_async(nl, sp, dp)
    t2 = time.perf_counter(); e1.record(stream)
    t3 = time.perf_counter(); solver.sync()
    t4 = time.perf_counter()
    print(f"rank {rank} it {it}: rec {1e3*(t1-t0):.3f} enqueue {1e3*(t2-t1):.3f} rec {1e3*(t3-t2):.3f} sync {1e3*(t4-t3):.3f} | events {e0.elapsed_time(e1):.3f} " + str({k: round(v, 2) for k, v in solver.timings_ms().items()}), flush=True)
if use_dist:
    dist.barrier(); dist.destroy_process_group()
