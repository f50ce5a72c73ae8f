"""Scale vote (estimate_scaling=True, upstream's default; the reference sets False): device time and HBM reading.

Per problem of N correspondences: P = N (N - 1) / 2 TIM pairs, E = 2 P interval endpoints, one stable LSD radix sort of E (u64 key, u32 payload)
in 8 passes of 8 bits (each pass: histogram read 8 E + scatter read 12 E + write 12 E bytes), key generation (48 N read, 16 P + 24 P written),
two sweep passes (4 E payload + gathers of X and R).  Algorithmic bytes below count the sort passes and the key / sweep streams."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import teaserpp_python as tpp
from poseestimator_b200 import batch as b2batch, synth

peaks = {}
pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(pk):
    peaks = json.load(open(pk))
hbm = float(peaks.get("hbm_gbs", 6650.0))
out = {"hbm_peak_gbs": hbm, "radix_passes": 8, "radix_bits": 8, "cases": []}
for n, batch in ((1000, 1), (1000, 16), (5000, 1), (5000, 4)):
    pbs = [synth.make_problem(n, 0.5, seed=70 + i, scale=1.3) for i in range(batch)]
    p = tpp.RobustRegistrationSolver.Params()
    p.noise_bound = pbs[0].noise_bound; p.estimate_scaling = True; p.rotation_cost_threshold = 1e-12
    sv = b2batch.BatchSolver(p, profiling=True)
    packed = [(np.ascontiguousarray(q.src.T), np.ascontiguousarray(q.dst.T)) for q in pbs]
    for _ in range(3):
        sv.solve_packed_async(packed); sv.sync()
    acc = []
    l0 = sv.launch_count()
    for _ in range(5):
        sv.solve_packed_async(packed); sv.sync()
        acc.append(sv.timings_ms()["scale_vote"])
    launches = (sv.launch_count() - l0) / 5
    ms = float(np.median(acc))
    P = n * (n - 1) // 2
    E = 2 * P
    sort_bytes = 8 * (8 * E + 12 * E + 12 * E)
    other = 48 * n + 40 * P + 2 * (4 * E + 16 * E)
    alg = batch * (sort_bytes + other)
    sols = sv.solutions_array()
    assert bool(sols["valid"].all()) and np.allclose(sols["scale"], 1.3, atol=0.02), sols["scale"]
    out["cases"].append({"n": n, "batch": batch, "endpoints_per_problem": E, "scale_vote_ms": ms, "ms_per_problem": ms / batch,
                         "algorithmic_bytes": alg, "GBps": alg / ms / 1e6, "frac_of_hbm": alg / ms / 1e6 / hbm, "launches_per_batch": launches,
                         "scale_estimates": [float(x) for x in sols["scale"][:2]]})
print(json.dumps(out))
