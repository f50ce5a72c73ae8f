#!/bin/bash
# Round-end evidence in one gpurun call: tools/final_pass.sh <scratch dir under gpurun_out/>
# GPU tests, bench (both arms), the ncu launch list of the bench command and one --set full capture of the graph kernel.
# Every step under its own timeout; a number printed by a run under ncu is never a bench value.
D=gpurun_out/$1
mkdir -p $D
timeout 900 python -m pytest tests -m gpu -x -q > $D/pytest_gpu.log 2>&1; tail -2 $D/pytest_gpu.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke(); print("smoke ok")' > $D/smoke.log 2>&1; tail -1 $D/smoke.log
timeout 600 python bench.py > $D/bench.json 2> $D/bench.err || tail -5 $D/bench.err
timeout 600 python bench.py --impl reference > $D/bench_ref.json 2> $D/bench_ref.err || tail -5 $D/bench_ref.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $D/bench_short.json 2> $D/bench_short.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $D/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $D/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tim_graph6 -s 3 -c 1 -o $D/graph6_full -f \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $D/ncu_full.log 2>&1
ls -la $D | tail -12
python - <<PY
import json
j = json.loads(open("$D/bench.json").read().strip().splitlines()[-1])
print("value", j["value"], "one", j["value_one_handle"], "e2e", j["e2e"]["value"], "pageable", j["e2e_pageable_inputs"]["value"], j["clocks"], j.get("cpu_baseline"))
PY
