"""The config-2 leg of bench.py on its own (one GPU): python tools/config2_leg_only.py [n_frames]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
import bench_legs

torch.cuda.set_device(0)
cx = bench_legs.Ctx(torch, None, 0, 1, 0)
out = bench_legs.config2_leg(cx, n_frames=int(sys.argv[1]) if len(sys.argv) > 1 else 256)
print(json.dumps({k: out[k] for k in ("frames_per_s", "frames_per_s_serial", "frame_ms_p50", "frame_ms_p99", "stage_ms", "launches_per_frame",
                                      "nn_fallback_rows_per_frame") if k in out}))
