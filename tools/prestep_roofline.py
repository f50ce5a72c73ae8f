"""Pre-step kernels at sizes that fill a B200 (the config-2 frame is launch-latency bound: 56k points, 2.4 ms for 62 launches).

For each kernel the north-star names -- back-projection, FPFH, the NN GEMM -- one large synthetic input, timed with CUDA events
on the library's stream after warm-up, with the ALGORITHMIC bytes / flops of SURVEY.md 8(d):

  back-projection  2 B depth + 1 B mask read per bbox pixel, 24 B written per valid point
  FPFH             per point: <= max_nn neighbours x (2 x 24 B gathers) for SPFH + <= max_nn x 33 x 8 B SPFH gathers, 264 B written
  NN GEMM          2 * Vs * Vd * 33 flops (the 3xTF32 split issues 3x that; reported against the algorithmic figure)

Prints one JSON object; run it plain for the numbers and under ncu for the dram / tensor-pipe counters (profiles/).
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from poseestimator_b200.presteps import PreSteps


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(reps):
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def main():
    quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
    peaks = {}
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        peaks = json.load(open(pk))
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    ps = PreSteps(0)
    rng = np.random.default_rng(0)
    out = {"hbm_peak_gbs": hbm}

    # ---- back-projection: a 4096 x 4096 depth image, 70 % of the pixels valid ----
    H = W = 2048 if quick else 4096
    depth = rng.integers(400, 2500, (H, W), dtype=np.uint16)
    depth[rng.random((H, W)) < 0.1] = 0
    mask = (rng.random((H, W)) < 0.8).astype(np.uint8) * 255
    d_depth = torch.from_numpy(depth.view(np.int16)).cuda().view(torch.uint16)
    d_mask = torch.from_numpy(mask).cuda()
    res = {}

    def bp():
        res["pts"] = ps.backproject(d_depth, d_mask, None, 615.0, 615.0, W / 2, H / 2, 0.001, depth_trunc=3.0)

    ms = timed(bp, reps=5 if quick else 10)
    m = int(res["pts"].shape[0])
    alg = 3 * H * W + 24 * m
    out["backproject"] = {"pixels": H * W, "valid_points": m, "ms": ms, "algorithmic_bytes": alg, "achieved_gbs": alg / ms / 1e6,
                          "frac_of_hbm_peak": alg / ms / 1e6 / hbm,
                          "note": "includes the host read of the point count (one 4-byte D2H + sync per call)"}
    del res["pts"]

    # ---- FPFH on a large surface cloud (two noisy planes + a sphere, spacing ~ voxel) ----
    n = 60000 if quick else 400000
    vs = 0.005
    side = int(np.sqrt(n // 2))
    g = (np.stack(np.meshgrid(np.arange(side), np.arange(side)), -1).reshape(-1, 2) * vs).astype(np.float64)
    plane = np.concatenate([g, rng.normal(0, vs * 0.05, (g.shape[0], 1))], 1)
    u = rng.normal(size=(n - plane.shape[0], 3)); u /= np.linalg.norm(u, axis=1, keepdims=True)
    sphere = u * (side * vs / 4) + np.array([side * vs / 2, side * vs / 2, side * vs / 2])
    pts = torch.from_numpy(np.concatenate([plane, sphere], 0)).cuda()
    n = int(pts.shape[0])
    nrm = ps.normals(pts, 2 * vs, 30, 1, (0.0, 0.0, 10.0))
    res = {}

    def fp():
        res["f"] = ps.fpfh(pts, nrm, 5 * vs, 100)

    ms = timed(fp, reps=3 if quick else 5, warm=2)
    _, _, cnt = ps.knn(pts, pts, 100, 5 * vs)
    nb = float(cnt.double().mean().item())
    alg = n * (nb * 48 + nb * 33 * 8 + 264)
    out["fpfh"] = {"points": n, "mean_neighbours": nb, "ms_total_call": ms, "algorithmic_bytes": alg, "achieved_gbs_whole_call": alg / ms / 1e6,
                   "frac_of_hbm_peak_whole_call": alg / ms / 1e6 / hbm,
                   "note": "whole b2r_fpfh call = radius kNN (grid) + spfh_kernel + fpfh_kernel + transpose; per-kernel times in the ncu launch list"}

    # ---- NN GEMM: Vs x Vd descriptors of dim 33 ----
    vq = vr = 8192 if quick else 32768
    fq = torch.from_numpy(rng.gamma(2.0, 5.0, (vq, 33))).cuda()
    fr = torch.from_numpy(rng.gamma(2.0, 5.0, (vr, 33))).cuda()
    res = {}

    def nn():
        res["nn"] = ps.nn_gemm(fq, fr)

    ms = timed(nn, reps=3 if quick else 5, warm=2)
    flops = 2.0 * vq * vr * 33
    out["nn_gemm"] = {"queries": vq, "references": vr, "dim": 33, "ms_total_call": ms, "algorithmic_flops": flops,
                      "achieved_tflops_whole_call": flops / ms / 1e9, "fallback_rows": int(ps.last_fallback_rows),
                      "note": "whole b2r_nn_gemm call = prep (split) + nn_gemm_kernel + nn_finish_kernel; per-kernel times in the ncu launch list"}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
