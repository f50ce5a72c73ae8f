"""BASELINE config 2 end to end: 640x480 depth frame -> bbox/mask back-projection -> outlier removal -> 5 mm voxel ->
normals + FPFH -> mutual NN against the CAD cloud -> TEASER++ solve.  GPU pipeline vs the same stages of the oracle."""
import numpy as np
import pytest

from poseestimator_b200 import synth

pytestmark = pytest.mark.gpu


def _oracle_pipeline(oracle_mod, op, f, cfg, cad):
    raw, _ = op.backproject(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale, cfg.depth_trunc)
    keep, _ = op.statistical_outlier(raw, cfg.outlier_nb, cfg.outlier_std)
    kept = raw[keep]
    scene, _, _ = op.voxel_down(kept, cfg.voxel)
    tmpl, _, _ = op.voxel_down(cad, cfg.voxel)
    rn, rf = cfg.normal_radius_mult * cfg.voxel, cfg.fpfh_radius_mult * cfg.voxel
    n_s = op.normals(scene, rn, cfg.normal_max_nn, 1, (0, 0, 0))
    n_t = op.normals(tmpl, rn, cfg.normal_max_nn, 2, (0, 0, 0))
    f_s = op.fpfh(scene, n_s, rf, cfg.fpfh_max_nn).T.copy()
    f_t = op.fpfh(tmpl, n_t, rf, cfg.fpfh_max_nn).T.copy()
    pairs, _, _ = op.mutual_nn(f_t, f_s)
    src = tmpl[pairs[:, 0]].T
    dst = scene[pairs[:, 1]].T
    ref = oracle_mod.solve(src, dst, noise_bound=cfg.noise_bound_mult * cfg.voxel, estimate_scaling=0, rotation_cost_threshold=1e-12)
    return dict(raw=raw, kept=kept, scene=scene, tmpl=tmpl, pairs=pairs, ref=ref)


_SHAPES = {"box": (0.30, 0.15, 0.10), "slab": (0.22, 0.22, 0.06), "bar": (0.40, 0.10, 0.08)}


@pytest.mark.parametrize("seed,shape", [(s, sh) for sh in _SHAPES for s in range(2, 12)])
def test_config2_pipeline_matches_oracle(oracle_mod, seed, shape):
    from oracle import presteps as op
    from poseestimator_b200.pipeline import FrameRegistration, PipelineConfig
    f = synth.make_frame(seed=seed + 100 * list(_SHAPES).index(shape), size=_SHAPES[shape], n_cad=3000)
    cfg = PipelineConfig()
    o = _oracle_pipeline(oracle_mod, op, f, cfg, f.cad)
    reg = FrameRegistration(f.cad, config=cfg)
    res = reg.register(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale, want_pairs=True, timing=True)
    assert res.n_raw == o["raw"].shape[0] and res.n_filtered == o["kept"].shape[0]
    assert np.array_equal(reg.template.cpu().numpy(), o["tmpl"])
    assert res.n_scene == o["scene"].shape[0]
    # correspondence indices: identical (a histogram bin flipped by a last-ulp libm difference would show up here)
    assert np.array_equal(res.pairs, o["pairs"])
    ref = o["ref"]
    assert res.solution.valid == bool(ref.valid)
    assert res.stats["clique_size"] == len(ref.clique)
    assert reg.solver.max_clique(0).tolist() == ref.clique.tolist()
    if len(ref.clique) >= 3:       # a 2-clique gives collinear TIMs: rank-1 H, the rotation about them is free
        assert synth.rotation_error_rad(res.solution.rotation, ref.rotation) < 1e-5
        scene_scale = float(np.linalg.norm(o["scene"].max(0) - o["scene"].min(0)))
        assert np.linalg.norm(res.solution.translation - ref.translation) < 1e-6 * scene_scale
    assert res.n_corr >= 3 and res.timings_ms["total"] > 0


def test_pipeline_degenerate_frames():
    from poseestimator_b200.pipeline import FrameRegistration
    f = synth.make_frame(seed=2, n_cad=800)
    reg = FrameRegistration(f.cad)
    empty = reg.register(f.depth, np.zeros_like(f.mask), f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    assert not empty.solution.valid and empty.n_raw == 0
    tiny = np.zeros_like(f.mask); tiny[240, 320] = 255; tiny[240, 321] = 255
    r = reg.register(f.depth, tiny, None, f.fx, f.fy, f.cx, f.cy, f.depth_scale)
    assert not r.solution.valid and r.n_raw <= 2


def test_find_best_template_matches_oracle_loop(oracle_mod):
    """PoseEstimator.find_best_template_teaser: three templates (the right box, a wrong box, a half-size box); every stage of
    the loop -- TEASER pose, ICP refinement, alignment score, the arg-min -- against the oracle's."""
    from oracle import presteps as op
    from oracle import refine as orf
    from poseestimator_b200.pipeline import PipelineConfig, TemplateMatcher
    cfg = PipelineConfig()
    f = synth.make_frame(seed=2)
    wrong = synth.make_frame(seed=3, size=(0.22, 0.22, 0.06)).cad
    small = synth.make_frame(seed=4, size=(0.15, 0.075, 0.05)).cad
    raw, _ = op.backproject(f.depth, f.mask, f.bbox, f.fx, f.fy, f.cx, f.cy, f.depth_scale, cfg.depth_trunc)
    keep, _ = op.statistical_outlier(raw, cfg.outlier_nb, cfg.outlier_std)
    dst_cloud = raw[keep]
    tm = TemplateMatcher([wrong, f.cad, small], config=cfg, icp_max_iteration=60)
    H, src_best, recs = tm.find_best_template_teaser(dst_cloud, details=True)
    # oracle loop
    rn, rfp = cfg.normal_radius_mult * cfg.voxel, cfg.fpfh_radius_mult * cfg.voxel
    dst_down, _, _ = op.voxel_down(dst_cloud, cfg.voxel)
    f_d = op.fpfh(dst_down, op.normals(dst_down, rn, cfg.normal_max_nn, 1, (0, 0, 0)), rfp, cfg.fpfh_max_nn).T.copy()
    scores = []
    for k, cad in enumerate([wrong, f.cad, small]):
        src_down, _, _ = op.voxel_down(cad, cfg.voxel)
        f_s = op.fpfh(src_down, op.normals(src_down, rn, cfg.normal_max_nn, 2, (0, 0, 0)), rfp, cfg.fpfh_max_nn).T.copy()
        pairs, _, _ = op.mutual_nn(f_s, f_d)
        assert recs[k]["n_corr"] == pairs.shape[0]
        Ho = np.eye(4)
        determined = False
        if pairs.shape[0] >= 3:
            ref = oracle_mod.solve(src_down[pairs[:, 0]].T, dst_down[pairs[:, 1]].T, noise_bound=cfg.noise_bound_mult * cfg.voxel,
                                   estimate_scaling=0, rotation_cost_threshold=1e-12)
            assert recs[k]["valid"] == bool(ref.valid)
            if ref.valid:
                Ho[:3, :3] = ref.rotation; Ho[:3, 3] = ref.translation
            determined = len(ref.clique) >= 3      # a 2-clique gives collinear TIMs: rank-1 H, the rotation about them is free
        if determined or pairs.shape[0] < 3:
            assert synth.rotation_error_rad(recs[k]["H"][:3, :3], Ho[:3, :3]) < 1e-5
            assert np.linalg.norm(recs[k]["H"][:3, 3] - Ho[:3, 3]) < 1e-6
        # ICP and score from the GPU's own H (the oracle's H differs in the last bits; ICP from each converges to the same pose)
        icp = orf.icp(src_down, dst_down, 0.05, recs[k]["H"], 60)
        assert recs[k]["icp_iterations"] == icp.iterations
        assert np.allclose(recs[k]["T_icp"], icp.transformation, atol=1e-9)
        sc = orf.alignment_score(orf.transform(recs[k]["T_icp"], src_down), src_down, dst_down, cfg.voxel)["score"]
        assert recs[k]["score"] == sc
        scores.append(sc)
    # the winner is whatever the reference's score prefers (a full box seen from one side is penalised by its hidden faces)
    kb = int(np.argmin(scores))
    assert np.array_equal(H, recs[kb]["H"])
    assert np.array_equal(src_best.cpu().numpy(), op.voxel_down([wrong, f.cad, small][kb], cfg.voxel)[0])
    assert recs[1]["fitness"] > 0.3 and recs[1]["inlier_rmse"] < 0.05
