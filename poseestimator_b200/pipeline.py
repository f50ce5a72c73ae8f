"""Depth frame + detection mask -> 6-DoF pose of a CAD template, entirely on the GPU (BASELINE.json config 2).

Mirrors the reference's call order, one call per stage:

    RealSenseCamera.get_pcd_from_rgbd(mask)        RealSenseClass.py:56-69     back-projection + statistical outlier removal
    PoseEstimator.find_best_template_teaser(dst)   PoseEstimator.py:85-116     voxel_down_sample of scene and template
    run_teaser(source, target, voxel_size)         HelpersRealtime.py:132-172  FPFH of both, correspondences, TEASER++
(paths under /root/reference/src/pose_estimator/EstimHelpers/).  The correspondence step is the north-star's mutual
nearest neighbour on FPFH instead of Open3D's stochastic RANSAC.

`source` is the template (CAD) cloud, `target` the scene cloud, as in run_teaser: the solution maps template -> camera.
Template features are computed once and stay resident on the device across frames.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import batch as b2batch
from .presteps import PreSteps, _torch
from .teaser import RegistrationSolution, _Params


@dataclass
class PipelineConfig:
    voxel: float = 0.005                 # BASELINE config 2 (the reference uses 0.05, PoseEstimator.py:36)
    normal_radius_mult: float = 2.0      # reference: 1 x voxel (HelpersRealtime.py:216); sparse at 5 mm, so a parameter
    normal_max_nn: int = 30              # HelpersRealtime.py:218
    fpfh_radius_mult: float = 5.0        # HelpersRealtime.py:221
    fpfh_max_nn: int = 100               # HelpersRealtime.py:223
    outlier_nb: int = 20                 # RealSenseClass.py:68
    outlier_std: float = 1.0
    depth_trunc: float = 3.0             # Open3D RGBDImage default on the camera path
    noise_bound_mult: float = 1.5        # HelpersRealtime.py:139,159
    nn_method: str = "gemm"              # "gemm" (tcgen05 shortlist + exact re-rank) or "exact"
    orient_normals: bool = True          # scene normals towards the camera; template normals are given or towards `template_view`


@dataclass
class PipelineResult:
    solution: RegistrationSolution
    n_raw: int = 0
    n_filtered: int = 0
    n_scene: int = 0
    n_template: int = 0
    n_corr: int = 0
    pairs: "np.ndarray | None" = None       # [n_corr, 2] (template index, scene index)
    stats: dict = field(default_factory=dict)
    timings_ms: dict = field(default_factory=dict)


class FrameRegistration:
    """One template, many frames.  Everything between the host depth image and the 4x4 pose runs in libb2r.so kernels."""

    def __init__(self, template_points, template_normals=None, config: "PipelineConfig | None" = None, device: int = 0,
                 template_view=(0.0, 0.0, 0.0), voxelize_template: bool = True, solver_params: "_Params | None" = None):
        torch = _torch()
        self.cfg = config or PipelineConfig()
        self.device = int(device)
        self._dev = torch.device("cuda", self.device)
        torch.cuda.set_device(self._dev)
        self.ps = PreSteps(self.device)
        cfg = self.cfg
        tp = torch.from_numpy(np.ascontiguousarray(template_points, np.float64)).to(self._dev)
        if voxelize_template:                                   # PoseEstimator.py:93
            if template_normals is not None:
                raise ValueError("template normals are per input point; pass voxelize_template=False to keep them")
            tp, _, _ = self.ps.voxel_down(tp, cfg.voxel)
        if template_normals is None:
            tn = self.ps.normals(tp, cfg.normal_radius_mult * cfg.voxel, cfg.normal_max_nn, 2 if cfg.orient_normals else 0, template_view)
        else:
            tn = torch.from_numpy(np.ascontiguousarray(template_normals, np.float64)).to(self._dev)
        self.template = tp.contiguous()
        self.template_normals = tn.contiguous()
        self.template_fpfh = self.ps.fpfh(self.template, self.template_normals, cfg.fpfh_radius_mult * cfg.voxel, cfg.fpfh_max_nn)
        self.template_desc = self.template_fpfh.T.contiguous()   # one descriptor per row
        p = (solver_params or _Params())._copy()
        if solver_params is None:                                # run_teaser's parameters, HelpersRealtime.py:158-163
            p.cbar2 = 1.0
            p.noise_bound = cfg.noise_bound_mult * cfg.voxel
            p.estimate_scaling = False
            p.rotation_gnc_factor = 1.4
            p.rotation_max_iterations = 100
            p.rotation_cost_threshold = 1e-12
        self.params = p
        self.solver = b2batch.BatchSolver(p, device=self.device, stream=torch.cuda.current_stream(self._dev).cuda_stream or 1)

    # ------------------------------------------------------------------------------------------------------------
    def scene_cloud(self, depth, mask, bbox, fx, fy, cx, cy, depth_scale):
        """RealSenseCamera.get_pcd_from_rgbd: masked back-projection + remove_statistical_outlier(20, 1.0)."""
        cfg = self.cfg
        raw = self.ps.backproject(depth, mask, bbox, fx, fy, cx, cy, depth_scale, cfg.depth_trunc)
        if raw.shape[0] == 0:
            return raw, raw
        kept, _, _, _ = self.ps.statistical_outlier(raw, cfg.outlier_nb, cfg.outlier_std)
        return raw, kept

    def features(self, points, view=(0.0, 0.0, 0.0)):
        """extract_fpfh (HelpersRealtime.py:215-224): normals then FPFH; returns (normals, descriptors [n,33])."""
        cfg = self.cfg
        nrm = self.ps.normals(points, cfg.normal_radius_mult * cfg.voxel, cfg.normal_max_nn, 1 if cfg.orient_normals else 0, view)
        f = self.ps.fpfh(points, nrm, cfg.fpfh_radius_mult * cfg.voxel, cfg.fpfh_max_nn)
        return nrm, f.T.contiguous()

    def register(self, depth, mask, bbox, fx, fy, cx, cy, depth_scale, want_pairs: bool = False, timing: bool = False) -> PipelineResult:
        torch = _torch()
        ev = []

        def mark(name):
            if timing:
                e = torch.cuda.Event(enable_timing=True)
                e.record(torch.cuda.current_stream(self._dev))
                ev.append((name, e))

        mark("start")
        raw, kept = self.scene_cloud(depth, mask, bbox, fx, fy, cx, cy, depth_scale)
        mark("backproject+outliers")
        res = PipelineResult(solution=RegistrationSolution(), n_raw=int(raw.shape[0]), n_filtered=int(kept.shape[0]),
                             n_template=int(self.template.shape[0]))
        if kept.shape[0] < 3:
            return res
        scene, _, _ = self.ps.voxel_down(kept, self.cfg.voxel)          # PoseEstimator.py:87
        mark("voxel")
        res.n_scene = int(scene.shape[0])
        _, desc = self.features(scene)
        mark("normals+fpfh")
        pairs, _, _ = self.ps.mutual_nn(self.template_desc, desc, method=self.cfg.nn_method)
        mark("mutual_nn")
        res.n_corr = int(pairs.shape[0])
        res.stats["nn_fallback_rows"] = getattr(self.ps, "last_fallback_rows", 0)
        if want_pairs:
            res.pairs = pairs.cpu().numpy()
        if res.n_corr < 3:                                              # registration_utils.py:245-246
            return res
        src, dst = self.ps.gather_correspondences(self.template, scene, pairs)
        self.solver.solve_device_async([res.n_corr], [src.data_ptr()], [dst.data_ptr()], keep_alive=(src, dst))
        self.solver.sync()
        mark("teaser")
        res.solution = self.solver.solution(0)
        res.stats.update(self.solver.stats(0))
        if timing:
            torch.cuda.synchronize(self._dev)
            for (_, a), (name, b) in zip(ev[:-1], ev[1:]):
                res.timings_ms[name] = a.elapsed_time(b)
            res.timings_ms["total"] = ev[0][1].elapsed_time(ev[-1][1])
        return res

    def transform(self, sol: RegistrationSolution) -> np.ndarray:
        """4x4 as run_teaser returns it (HelpersRealtime.py:168-172)."""
        T = np.eye(4)
        T[:3, :3] = sol.scale * sol.rotation
        T[:3, 3] = sol.translation
        return T
