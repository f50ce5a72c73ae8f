#!/bin/bash
# compute-sanitizer over the GPU tests of the solver and the pre-steps: memcheck, then racecheck.
# tools/sanitize.sh <dir under gpurun_out> [pytest -k expression]
D=gpurun_out/${1:-sanitize}
K=${2:-"teaser or presteps"}
mkdir -p $D
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 86 --print-limit 50 --log-file $D/$tool.log \
      python -m pytest tests -m gpu -x -q -k "$K" -p no:cacheprovider > $D/pytest_$tool.log 2>&1
  echo "$tool: exit $? -- $(tail -1 $D/pytest_$tool.log)"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error|hazard" $D/$tool.log | sort | uniq -c | sort -rn | head -12
done
