"""Where does a config-4 frame (64 detections, N ~ U{200..1000}) spend its time?  Stage times (CUDA events), tail cycle marks of the
slowest problems, host-side wall time of submit / sync."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from poseestimator_b200 import batch as b2batch, synth
pbs = [synth.config4_detection(0, d) for d in range(64)]
packed = [(np.ascontiguousarray(p.src.T), np.ascontiguousarray(p.dst.T)) for p in pbs]
hb = b2batch.HostBatch(packed)
sv = b2batch.BatchSolver(bench.solver_params(pbs[0].noise_bound), profiling=True)
for _ in range(5):
    sv.solve_host_batch_async(hb); sv.sync()
acc = {}; sub = []; syn = []; tot = []
for _ in range(50):
    t0 = time.perf_counter(); sv.solve_host_batch_async(hb); t1 = time.perf_counter(); sv.sync(); t2 = time.perf_counter(); sv.solutions_array(); t3 = time.perf_counter()
    sub.append((t1 - t0) * 1e3); syn.append((t2 - t1) * 1e3); tot.append((t3 - t0) * 1e3)
    for k, v in sv.timings_ms().items():
        acc[k] = acc.get(k, 0) + v / 50
cyc = sorted(((sv.tail_cycles(i)["translation"], i) for i in range(64)), reverse=True)[:3]
out = {"stage_ms": {k: round(v, 4) for k, v in acc.items() if v}, "host_submit_ms_p50": float(np.median(sub)), "host_sync_ms_p50": float(np.median(syn)),
       "frame_ms_p50": float(np.median(tot)), "slowest_tail_problems": [{"i": i, "n": pbs[i].src.shape[1], "inliers": int(len(pbs[i].inliers)),
       "cycles": {k: v for k, v in sv.tail_cycles(i).items() if k != "marks"}} for _, i in cyc]}
print(json.dumps(out))
