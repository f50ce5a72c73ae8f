"""First-contact script for the GPU box: runs the path stage by stage and prints what differs from the oracle."""
import sys, os, time, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle
from poseestimator_b200 import synth, batch, _lib
import teaserpp_python as tpp


def params(nb, **kw):
    p = tpp.RobustRegistrationSolver.Params()
    p.noise_bound = nb; p.estimate_scaling = False; p.rotation_cost_threshold = 1e-12
    for k, v in kw.items(): setattr(p, k, v)
    return p


def run(n, ratio, seed, **kw):
    pb = synth.make_problem(n, ratio, seed)
    print(f"--- n={n} ratio={ratio} seed={seed} {kw}", flush=True)
    bs = batch.BatchSolver(params(pb.noise_bound, **kw), profiling=True)
    t = time.time()
    sols = bs.solve([(pb.src, pb.dst)])
    dt = time.time() - t
    st = bs.stats(0)
    print(" stats", st)
    print(" timings", {k: round(v, 4) for k, v in bs.timings_ms().items()}, "wall_ms", round(dt * 1e3, 3))
    ref = oracle.solve(pb.src, pb.dst, noise_bound=pb.noise_bound, estimate_scaling=0, rotation_cost_threshold=1e-12,
                       materialize_tims=0 if n > 2000 else 1)
    adj = bs.adjacency(0, n)
    diff = np.unpackbits((adj ^ ref.adjacency).view(np.uint8)).sum()
    print(" adjacency differing bits:", int(diff), "edges gpu", st["n_edges"], "ref", ref.n_edges)
    cl = bs.max_clique(0)
    print(" clique equal:", np.array_equal(cl, ref.clique), len(cl), len(ref.clique), "lb", st["heuristic_lb"], ref.heuristic_lb)
    print(" rot err vs oracle:", synth.rotation_error_rad(sols[0].rotation, ref.rotation), "t err:",
          np.linalg.norm(sols[0].translation - ref.translation), "valid", sols[0].valid, ref.valid,
          "iters", st["gnc_iterations"], ref.gnc_iters)
    print(" rot inl equal:", np.array_equal(bs.indices(0, _lib.IDX_ROT_INLIERS), ref.rotation_inliers),
          "trans inl equal:", np.array_equal(bs.indices(0, _lib.IDX_TRANS_INLIERS), ref.translation_inliers), flush=True)


if __name__ == "__main__":
    L = _lib.lib()
    print("devices", L.b2r_device_count(), "version", L.b2r_version(), flush=True)
    for args, kw in [((64, 0.5, 1), dict(graph_exact_only=True)), ((64, 0.5, 1), {}), ((400, 0.6, 7), {}), ((1000, 0.5, 1000), {}),
                     ((1000, 0.9, 3), {}), ((5000, 0.95, 5000), {})]:
        try:
            run(*args, **kw)
        except Exception:
            traceback.print_exc()
            break
