"""The C-ABI library loads and exports every symbol include/b2r.h declares (no compute calls: no GPU here)."""
import ctypes
import os
import re

from poseestimator_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols(header="b2r.h"):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b2r_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_are_exported():
    L = _lib.lib()
    names = _declared_symbols()
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert sorted(_lib.EXPORTS) == names, set(names) ^ set(_lib.EXPORTS)
    # the test hooks live in their own header, outside the drop-in boundary
    dbg = _declared_symbols("b2r_debug.h")
    assert not [n for n in dbg if not hasattr(L, n)]
    assert sorted(_lib.DEBUG_EXPORTS) == dbg, set(dbg) ^ set(_lib.DEBUG_EXPORTS)
    assert not [n for n in names if n.startswith("b2r_debug")]


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


def test_header_is_plain_c_and_struct_layouts_match_the_binding(tmp_path):
    """include/b2r.h compiles as C99 and as C++ with nothing but the standard headers, a C program links against libb2r.so, and
    the ctypes mirrors in _lib.py have the sizes the C compiler gives the structs (the stub INTEGRATION.md shows relies on it)."""
    import ctypes as C
    import subprocess
    src = tmp_path / "abi.c"
    src.write_text(
        '#include "b2r.h"\n#include <stdio.h>\n'
        "int main(void) {\n"
        '  printf("%zu %zu %zu %zu %zu %d\\n", sizeof(b2r_params), sizeof(b2r_solution), sizeof(b2r_stats), sizeof(b2r_camera),\n'
        "         sizeof(b2r_icp_result), b2r_version());\n"
        "  b2r_params p; b2r_default_params(&p);\n"
        '  printf("%.17g %.17g %d\\n", p.noise_bound, p.cbar2, (int)p.estimate_scaling);\n'
        "  return 0;\n}\n")
    inc = os.path.join(ROOT, "include")
    libdir = os.path.dirname(_lib.LIB_PATH)
    exe = tmp_path / "abi"
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-I", inc, str(src), "-o", str(exe), "-L", libdir, "-lb2r",
                    f"-Wl,-rpath,{libdir}"], check=True, capture_output=True)
    subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-x", "c++", "-I", inc, str(src)], check=True, capture_output=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    sizes = [int(x) for x in out[:5]]
    assert sizes == [C.sizeof(_lib.B2RParams), C.sizeof(_lib.B2RSolution), C.sizeof(_lib.B2RStats), C.sizeof(_lib.B2RCamera),
                     C.sizeof(_lib.B2RIcpResult)]
    assert int(out[5]) == _lib.lib().b2r_version()
    assert C.sizeof(_lib.B2RSolution) == _lib.SOLUTION_DTYPE.itemsize
    p = _lib.B2RParams()
    _lib.lib().b2r_default_params(C.byref(p))
    assert float(out[6]) == p.noise_bound and float(out[7]) == p.cbar2 and int(out[8]) == p.estimate_scaling
