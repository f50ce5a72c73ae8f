"""Two handles on two devices in ONE process, driven from two host threads (include/b2r.h: "thread-safe across handles"; SURVEY.md
8e row 1: one host thread + stream set per GPU).  Every kernel that opts into more than 48 KB of shared memory must have done so on
BOTH devices (the attribute is per device).  Skipped below two devices: `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multidevice.py`."""
import threading

import numpy as np
import pytest

from poseestimator_b200 import synth

pytestmark = pytest.mark.gpu


def _device_count():
    from poseestimator_b200 import _lib
    return _lib.lib().b2r_device_count()


def test_two_devices_two_threads_one_process(oracle_mod):
    if _device_count() < 2:
        pytest.skip("needs two CUDA devices")
    import teaserpp_python as tpp
    from poseestimator_b200 import batch as b2batch
    from poseestimator_b200.presteps import PreSteps
    pbs = [synth.config1(seed=1000 + d) for d in range(2)]
    big = [synth.make_problem(3000, 0.9, seed=77 + d) for d in range(2)]     # tensor-core filter (83 KB of shared memory) + fused tail
    refs = [oracle_mod.solve(p.src, p.dst, noise_bound=p.noise_bound, estimate_scaling=0, rotation_cost_threshold=1e-12) for p in pbs]
    refs_big = [oracle_mod.solve(p.src, p.dst, noise_bound=p.noise_bound, estimate_scaling=0, rotation_cost_threshold=1e-12) for p in big]
    out, errs = {}, []

    def work(dev):
        try:
            res = []
            for rep in range(3):
                for pb in (pbs[dev], big[dev]):
                    prm = tpp.RobustRegistrationSolver.Params()
                    prm.noise_bound = pb.noise_bound; prm.estimate_scaling = False; prm.rotation_cost_threshold = 1e-12
                    s = tpp.RobustRegistrationSolver(prm, device=dev)
                    sol = s.solve(pb.src, pb.dst)
                    res.append((sol, s.getInlierMaxClique(), s._adjacency_bits()))
            # batch path + pre-steps (nn_gemm: 193 KB of shared memory) on the same device
            bs = b2batch.BatchSolver(prm, device=dev)
            bs.solve_packed_async([(np.ascontiguousarray(big[dev].src.T), np.ascontiguousarray(big[dev].dst.T))] * 4); bs.sync()
            ps = PreSteps(dev)
            rng = np.random.default_rng(dev)
            fa, fb = rng.uniform(0, 100, (300, 33)), rng.uniform(0, 100, (500, 33))
            nn, _ = ps.nn_gemm(fa, fb); nn_e, _ = ps.nn_exact(fa, fb)
            assert np.array_equal(nn.cpu().numpy(), nn_e.cpu().numpy())
            out[dev] = (res, bs.solutions_array())
        except Exception as e:  # noqa: BLE001
            errs.append((dev, repr(e)))

    th = [threading.Thread(target=work, args=(d,)) for d in range(2)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
    for dev in range(2):
        res, bsol = out[dev]
        for k, (sol, clique, adj) in enumerate(res):
            ref = (refs if k % 2 == 0 else refs_big)[dev]
            assert sol.valid and np.array_equal(adj, ref.adjacency) and clique == ref.clique.tolist()
            assert synth.rotation_error_rad(sol.rotation, ref.rotation) < 1e-5
        assert bool(bsol["valid"].all())
        for i in range(4):
            assert synth.rotation_error_rad(bsol["rotation"][i], refs_big[dev].rotation) < 1e-5
