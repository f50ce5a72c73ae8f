"""A/B of the NN GEMM at a GPU-filling size: B2R_NN_SPLITS=<s> python tools/nn_ab.py [n]  -> per-kernel CUDA-event times."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from poseestimator_b200.presteps import PreSteps
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
ps = PreSteps(0)
rng = np.random.default_rng(0)
fq = torch.from_numpy(rng.gamma(2.0, 5.0, (n, 33))).cuda(); fr = torch.from_numpy(rng.gamma(2.0, 5.0, (n, 33))).cuda()
for _ in range(3): ps.nn_gemm(fq, fr)
torch.cuda.synchronize()
ts = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ps.nn_gemm(fq, fr); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
print(json.dumps({"n": n, "splits_env": os.environ.get("B2R_NN_SPLITS"), "ms_whole_call_median": float(np.median(ts)), "fallback_rows": ps.last_fallback_rows,
                  "algorithmic_tflops": 2.0 * n * n * 33 / np.median(ts) / 1e9}))
