import sys, time, json
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from poseestimator_b200 import synth
from poseestimator_b200.pipeline import FrameRegistration, PipelineConfig
f = synth.make_frame(seed=4)
reg = FrameRegistration(f.cad, config=PipelineConfig())
for _ in range(5): res = reg.register(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale, want_pairs=True)
pairs = torch.from_numpy(res.pairs).cuda()
raw, kept = reg.scene_cloud(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
scene, _, _ = reg.ps.voxel_down(kept, reg.cfg.voxel)
torch.cuda.synchronize()
T = {k: [] for k in ("gather", "submit", "sync", "solution", "stats")}
reg.solver._lib_prof = True
from poseestimator_b200 import _lib
_lib.lib().b2r_set_profiling(reg.solver._h, 1)
for _ in range(30):
    t0 = time.perf_counter(); src, dst = reg.ps.gather_correspondences(reg.template, scene, pairs); t1 = time.perf_counter()
    reg.solver.solve_device_async([pairs.shape[0]], [src.data_ptr()], [dst.data_ptr()], keep_alive=(src, dst)); t2 = time.perf_counter()
    reg.solver.sync(); t3 = time.perf_counter()
    sol = reg.solver.solution(0); t4 = time.perf_counter()
    st = reg.solver.stats(0); t5 = time.perf_counter()
    for k, a, b in (("gather", t0, t1), ("submit", t1, t2), ("sync", t2, t3), ("solution", t3, t4), ("stats", t4, t5)): T[k].append((b - a) * 1e3)
print(json.dumps({"n_corr": int(pairs.shape[0]), "host_ms_p50": {k: round(float(np.median(v)), 4) for k, v in T.items()}, "device_ms": {k: round(v, 4) for k, v in reg.solver.timings_ms().items() if v}, "stats": {k: st[k] for k in ("clique_size", "n_pruned", "heuristic_lb", "fused_tail", "gnc_iterations", "bnb_nodes")}}))
