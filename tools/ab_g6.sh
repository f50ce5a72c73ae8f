#!/bin/bash
# A/B of the tensor-core graph kernel variants (B2R_G6, see graph_tc.cu) with and without the fused tail: correctness of each
# (teaser GPU tests) and the per-stage times of a config-3 batch.
D=gpurun_out/${1:-g6}
mkdir -p $D
for v in ${2:-1 2 3}; do
  echo "== B2R_G6=$v"
  B2R_G6=$v timeout 600 python -m pytest tests/test_gpu_teaser.py tests/test_gpu_full_size.py -x -q > $D/pytest_g6_$v.log 2>&1; tail -1 $D/pytest_g6_$v.log
  for rep in 1 2; do B2R_G6=$v timeout 300 python tools/stage_times.py 64 5000 0.95 2>&1 | tail -5 | head -3; done
done 2>&1 | tee $D/ab_g6.log
echo "== no fused tail, default G6"
B2R_NO_FUSED_TAIL=1 timeout 600 python -m pytest tests/test_gpu_teaser.py -x -q 2>&1 | tail -1
B2R_NO_FUSED_TAIL=1 timeout 300 python tools/stage_times.py 64 5000 0.95 2>&1 | tail -2
