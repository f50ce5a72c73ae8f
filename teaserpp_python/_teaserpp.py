"""Stand-in for the compiled `_teaserpp` module of upstream's scikit-build package."""
from poseestimator_b200.teaser import (  # noqa: F401
    InlierGraphFormulation,
    InlierSelectionMode,
    RegistrationSolution,
    RobustRegistrationSolver,
    RotationEstimationAlgorithm,
)

# nested-enum spellings also available at module level, like upstream's bindings
ROTATION_ESTIMATION_ALGORITHM = RotationEstimationAlgorithm
INLIER_SELECTION_MODE = InlierSelectionMode
INLIER_GRAPH_FORMULATION = InlierGraphFormulation

__all__ = [
    "RobustRegistrationSolver", "RegistrationSolution", "RotationEstimationAlgorithm", "InlierSelectionMode",
    "InlierGraphFormulation", "ROTATION_ESTIMATION_ALGORITHM", "INLIER_SELECTION_MODE", "INLIER_GRAPH_FORMULATION",
]
