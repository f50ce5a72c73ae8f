#!/bin/bash
# The whole GPU test suite with guard zones around every workspace buffer (B2R_CANARY=1, include/b2r_debug.h): an out-of-bounds
# write of any kernel past either end of a buffer fails the call that made it.  compute-sanitizer is closed on the GPU pool
# (profiles/r02_sanitizer_unavailable.md), this is the check that replaces memcheck for the workspace.
D=gpurun_out/${1:-canary}
mkdir -p $D
B2R_CANARY=1 timeout 1500 python -m pytest tests -m gpu -x -q -p no:cacheprovider > $D/pytest_canary.log 2>&1
echo "canary run: exit $? -- $(tail -1 $D/pytest_canary.log)"
