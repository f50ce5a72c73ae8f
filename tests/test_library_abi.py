"""The C-ABI library loads and exports every symbol include/b2r.h declares (no compute calls: no GPU here)."""
import ctypes
import os
import re

from poseestimator_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    txt = open(os.path.join(ROOT, "include", "b2r.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b2r_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_are_exported():
    L = _lib.lib()
    names = _declared_symbols()
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(_lib.EXPORTS) == names, set(names) ^ set(_lib.EXPORTS)


def test_no_device_means_error_not_fallback():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    L = _lib.lib()
    assert L.b2r_device_count() == 0
    p = _lib.B2RParams()
    L.b2r_default_params(ctypes.byref(p))
    assert p.noise_bound == 0.01 and p.estimate_scaling == 1 and p.rotation_gnc_factor == 1.4
    h = ctypes.c_void_p()
    assert L.b2r_create(ctypes.byref(p), 0, None, ctypes.byref(h)) == 2  # B2R_ERR_CUDA
    import teaserpp_python
    with pytest.raises(RuntimeError):
        teaserpp_python.RobustRegistrationSolver()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "poseestimator_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "liboracle" not in src, f
    for f in os.listdir(os.path.join(ROOT, "teaserpp_python")):
        if f.endswith(".py"):
            assert "oracle" not in open(os.path.join(ROOT, "teaserpp_python", f)).read()
