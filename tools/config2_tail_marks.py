"""Config-2 frames: the `teaser` stage's device time and the fused tail's cycle marks, frame by frame (which part of the
solve a depth frame pays for).  Run on the GPU box: python tools/config2_tail_marks.py"""
import sys
sys.path.insert(0, '/root/repo')
import torch
from poseestimator_b200 import synth, batch as b2batch
from poseestimator_b200.pipeline import FrameRegistration, PipelineConfig

frames = [synth.make_frame(seed=s) for s in (2, 3, 4, 5, 6, 7, 8, 9)]
reg = FrameRegistration(frames[0].cad, config=PipelineConfig(), device=0)
reg.solver = b2batch.BatchSolver(reg.params, device=0, stream=torch.cuda.current_stream().cuda_stream or 1, profiling=True)
for s, f in zip((2, 3, 4, 5, 6, 7, 8, 9), frames):
    for _ in range(3):
        res = reg.register(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale, timing=True)
    st = {k: res.stats.get(k) for k in ("clique_size", "heuristic_lb", "n_pruned", "fused_tail", "gnc_iterations", "bnb_nodes")}
    c = reg.solver.tail_cycles(0)
    marks = c.pop("marks")
    print(s, "n_corr", res.n_corr, st, "teaser ms %.3f" % res.timings_ms.get("teaser", -1), "total ms %.3f" % res.timings_ms.get("total", -1), c, marks)
