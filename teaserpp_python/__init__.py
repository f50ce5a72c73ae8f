"""Drop-in `teaserpp_python` package backed by poseestimator_b200 (B200 CUDA kernels, no CPU path).

Both import forms used by the reference work:
    import teaserpp_python                                  (HelpersRealtime.py:3)
    from teaserpp_python import _teaserpp as tpp            (registration_utils.py:5, main_image.py:4)
"""
from . import _teaserpp  # noqa: F401
from ._teaserpp import *  # noqa: F401,F403
