"""Warm per-stage device times of the config-2 pipeline (CUDA events), for DESIGN.md / profiles."""
import json
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from poseestimator_b200 import synth
from poseestimator_b200.pipeline import FrameRegistration, PipelineConfig

f = synth.make_frame(seed=2)
reg = FrameRegistration(f.cad, config=PipelineConfig())
acc = {}
R = 20
for i in range(R + 3):
    res = reg.register(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale, timing=True)
    if i >= 3:
        for k, v in res.timings_ms.items():
            acc[k] = acc.get(k, 0.0) + v / R
print(json.dumps({"stage_ms": {k: round(v, 4) for k, v in acc.items()}, "n_raw": res.n_raw, "n_filtered": res.n_filtered,
                  "n_scene": res.n_scene, "n_template": res.n_template, "n_corr": res.n_corr, "valid": res.solution.valid,
                  "nn_fallback_rows": res.stats.get("nn_fallback_rows"), "launches": reg.ps.launch_count() + reg.solver.launch_count()}))
