"""B200-native registration stage of patricksko/PoseEstimator (YOLO -> TEASER++ path).

Public surface:
  poseestimator_b200.teaser     -- `teaserpp_python`-compatible RobustRegistrationSolver (also importable as the
                                   top-level `teaserpp_python` package shipped in this repo)
  poseestimator_b200.batch      -- batched / multi-GPU registration of independent detections
  poseestimator_b200.synth      -- seeded synthetic problems (BASELINE.json configs)
The arithmetic lives in csrc/ (hand-written sm_100a CUDA behind the C ABI of include/b2r.h); nothing here falls
back to the CPU.
"""
from .teaser import (  # noqa: F401
    InlierGraphFormulation,
    InlierSelectionMode,
    RegistrationSolution,
    RobustRegistrationSolver,
    RotationEstimationAlgorithm,
)

__version__ = "0.1.0"
