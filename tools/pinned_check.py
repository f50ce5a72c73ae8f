"""Upload time of one 64 x 5000 batch: page-locked caller arrays (on_device = 2) vs staging (on_device = 0).

    python tools/pinned_check.py        (on a GPU box; prints one JSON line per variant)
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from poseestimator_b200 import batch as b2batch  # noqa: E402


def main():
    B, N = 64, 5000
    pbs = bench.make_problems(B, 0)
    nb = pbs[0].noise_bound
    packed = [(np.ascontiguousarray(p.src.T), np.ascontiguousarray(p.dst.T)) for p in pbs]
    ps, pd = b2batch.pinned_problem_block(B, N)
    for i, (a, b) in enumerate(packed):
        b2batch.PinnedBuffer.fill(ps[i], a); b2batch.PinnedBuffer.fill(pd[i], b)
    pinned = [(ps[i], pd[i]) for i in range(B)]
    ps2, pd2 = b2batch.pinned_problem_block(B, N)
    for i, (a, b) in enumerate(packed):
        ps2[i] = a; pd2[i] = b
    pinned_plain = [(ps2[i], pd2[i]) for i in range(B)]
    sv = b2batch.BatchSolver(bench.solver_params(nb), profiling=True)
    for name, fn in (("staged", lambda: sv.solve_packed_async(packed)), ("pinned_nt_filled", lambda: sv.solve_pinned_async(pinned)),
                     ("pinned_plain_filled", lambda: sv.solve_pinned_async(pinned_plain)), ("staged_again", lambda: sv.solve_packed_async(packed))):
        up, call, tot, wall = [], [], [], []
        for it in range(12):
            t0 = time.perf_counter()
            fn()
            t1 = time.perf_counter()
            sv.sync()
            t2 = time.perf_counter()
            tm = sv.timings_ms()
            if it >= 3:
                up.append(tm["upload"]); call.append((t1 - t0) * 1e3); tot.append(tm["total"]); wall.append((t2 - t0) * 1e3)
        print(json.dumps({"variant": name, "upload_ms": float(np.median(up)), "submit_call_ms": float(np.median(call)),
                          "device_total_ms": float(np.median(tot)), "wall_ms": float(np.median(wall))}), flush=True)


if __name__ == "__main__":
    main()
