"""Where the host time of a config-2 frame goes: cProfile of FrameRegistration.register (C calls vs Python glue).
Run on the GPU box: python tools/config2_host_profile.py"""
import cProfile, pstats, sys, time
sys.path.insert(0, '/root/repo')
import torch
from poseestimator_b200 import synth
from poseestimator_b200.pipeline import FrameRegistration, PipelineConfig

f = synth.make_frame(seed=4)
reg = FrameRegistration(f.cad, config=PipelineConfig(), device=0)
for _ in range(5):
    reg.register(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
N = 200
t0 = time.perf_counter()
for _ in range(N):
    reg.register(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
print("wall ms per frame", (time.perf_counter() - t0) * 1e3 / N)
pr = cProfile.Profile()
pr.enable()
for _ in range(N):
    reg.register(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)
