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


class TemplateMatcher:
    """`PoseEstimator.find_best_template_teaser` on the GPU (/root/reference/src/pose_estimator/EstimHelpers/PoseEstimator.py:85-116):

        dst_down = dst_cloud.voxel_down_sample(voxel)
        for every template:  src_down = src.voxel_down_sample(voxel);  H = run_teaser(src_down, dst_down, voxel)
                             T = registration_icp(src_down, dst_down, 0.05, H, PointToPoint, max_iteration=1000).transformation
                             score = alignment_score(src_down.transform(T), src_down, dst_down, voxel);  keep the lowest score
        return best H (the TEASER++ pose, as the reference does) and its src_down

    Templates are the batch axis: their voxel clouds, normals and FPFH descriptors are computed once and stay resident.  The
    correspondence step is the north-star's mutual nearest neighbour on FPFH."""

    def __init__(self, templates, config: "PipelineConfig | None" = None, device: int = 0, icp_distance: float = 0.05,
                 icp_max_iteration: int = 1000, template_view=(0.0, 0.0, 0.0)):
        from .refine import Refiner
        self.cfg = config or PipelineConfig()
        self.icp_distance = float(icp_distance)          # PoseEstimator.py:98
        self.icp_max_iteration = int(icp_max_iteration)  # PoseEstimator.py:100
        self.regs = [FrameRegistration(t, config=self.cfg, device=device, template_view=template_view) for t in templates if t is not None]
        self.rf = Refiner(device)

    def run_teaser(self, k: int, dst_down, dst_desc):
        """run_teaser(source=template k, target=dst_down) -> (4x4 H, solution, n_corr)  (HelpersRealtime.py:132-172)."""
        reg = self.regs[k]
        pairs, _, _ = reg.ps.mutual_nn(reg.template_desc, dst_desc, method=self.cfg.nn_method)
        n = int(pairs.shape[0])
        if n < 3:
            return np.eye(4), RegistrationSolution(), n
        src, dst = reg.ps.gather_correspondences(reg.template, dst_down, pairs)
        reg.solver.solve_device_async([n], [src.data_ptr()], [dst.data_ptr()], keep_alive=(src, dst))
        reg.solver.sync()
        sol = reg.solver.solution(0)
        return reg.transform(sol), sol, n

    def find_best_template_teaser(self, dst_cloud, details: bool = False):
        """dst_cloud: [n,3] scene points (numpy or CUDA tensor).  -> (H_best, src_down_best) (+ per-template records)."""
        if not self.regs:
            return (np.eye(4), None, []) if details else (np.eye(4), None)
        r0 = self.regs[0]
        dst_down, _, _ = r0.ps.voxel_down(dst_cloud, self.cfg.voxel)                       # PoseEstimator.py:87
        _, dst_desc = r0.features(dst_down)
        best = dict(T=np.eye(4), score=np.inf, src=None, idx=-1)
        recs = []
        for k, reg in enumerate(self.regs):
            H, sol, n_corr = self.run_teaser(k, dst_down, dst_desc)
            icp = self.rf.registration_icp(reg.template, dst_down, self.icp_distance, H, max_iteration=self.icp_max_iteration)
            src_aligned = self.rf.transform(reg.template, icp.transformation)
            score = self.rf.alignment_score(src_aligned, reg.template, dst_down, self.cfg.voxel)
            recs.append(dict(template=k, n_corr=n_corr, valid=bool(sol.valid), H=H, T_icp=icp.transformation, fitness=icp.fitness,
                             inlier_rmse=icp.inlier_rmse, icp_iterations=icp.iterations, score=score))
            if score < best["score"]:
                best.update(score=score, T=H, src=reg.template, idx=k)
        if details:
            return best["T"], best["src"], recs
        return best["T"], best["src"]
