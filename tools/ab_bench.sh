#!/bin/bash
# tools/ab_bench.sh tag1 tag2 ... : bench.py once per A/B library (poseestimator_b200/csrc/_ab/libb2r_<tag>.so), key numbers per line
for tag in "$@"; do
  for rep in 1 2; do
    B2R_LIB=$PWD/poseestimator_b200/csrc/_ab/libb2r_$tag.so timeout 300 python bench.py --no-cpu-baseline > /tmp/ab_$tag.json 2> /tmp/ab_$tag.err || { echo "$tag FAILED"; tail -3 /tmp/ab_$tag.err; continue; }
    python - "$tag" <<PY
import json, sys
j = json.loads(open("/tmp/ab_%s.json" % sys.argv[1]).read().strip().splitlines()[-1])
s = j["stage_ms_per_step"]
print(sys.argv[1], "value %.0f one %.0f e2e %.0f | graph %.4f trans %.4f canon %.4f total %.4f" % (j["value"], j["value_one_handle"], j["e2e"]["value"], s["graph"], s["translation"], s["canonical"], s["total"]))
PY
  done
done
