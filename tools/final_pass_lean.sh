#!/bin/bash
# Round-end evidence, lean form (the kernels whose --set full captures are already under profiles/ and have not changed are skipped):
# tools/final_pass_lean.sh <scratch dir under gpurun_out/>
D=gpurun_out/$1
mkdir -p $D
timeout 1200 python -m pytest tests -m gpu -x -q > $D/pytest_gpu.log 2>&1; tail -2 $D/pytest_gpu.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke(); print("smoke ok")' > $D/smoke.log 2>&1; tail -1 $D/smoke.log
timeout 900 python bench.py > $D/bench.json 2> $D/bench.err || tail -5 $D/bench.err
timeout 600 python bench.py --impl reference --steps 10 > $D/bench_ref.json 2> $D/bench_ref.err || tail -5 $D/bench_ref.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-legs --min-device-ms 0 > $D/bench_short.json 2> $D/bench_short.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $D/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-legs --min-device-ms 0 > $D/ncu_list.log 2>&1
timeout 600 python tools/prestep_roofline.py > $D/prestep_roofline.json 2> $D/prestep_roofline.err
for k in knn_grid_kernel fpfh_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -o $D/$k -f python tools/prestep_roofline.py quick > $D/ncu_$k.log 2>&1
done
timeout 300 python tools/pipeline_times.py > $D/config2_stage_times.json 2>/dev/null
timeout 300 python tools/config2_tail_marks.py > $D/config2_tail_marks.log 2>/dev/null
timeout 300 python tools/config4_times.py > $D/config4_times.json 2>/dev/null
timeout 300 python tools/tail_marks.py > $D/tail_marks.log 2>/dev/null
timeout 500 ncu --metrics gpu__time_duration.sum --clock-control none -s 600 -c 140 --csv --log-file $D/frame_launches.csv python tools/pipeline_times.py > $D/ncu_frame.log 2>&1
ls -la $D | tail -30
python - <<PY
import json
j = json.loads(open("$D/bench.json").read().strip().splitlines()[-1])
print("value", j["value"], "one", j["value_one_handle"], "e2e", j["e2e"]["value"], "pageable", j["e2e_pageable_inputs"]["value"], j["clocks"], j.get("cpu_baseline"))
print("config2", json.dumps(j.get("config2"))[:900])
PY
