#!/bin/bash
# quick GPU check of the tensor-core graph filter: stage times, per-kernel ncu times, teaser parity tests
mkdir -p gpurun_out/$1
timeout 120 python tools/stage_times.py 2>&1 | tail -2
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"graph" -c 30 --csv --log-file gpurun_out/$1/l.csv python tools/stage_times.py > gpurun_out/$1/l.log 2>&1
python - <<PY
import csv,io
t=open("gpurun_out/$1/l.csv").read(); t=t[t.find('"ID"'):]
for r in list(csv.DictReader(io.StringIO(t)))[-3:]: print(r["Kernel Name"][:40], r["Grid Size"], r["Metric Value"], r["Metric Unit"])
PY
timeout 300 python -m pytest tests/test_gpu_teaser.py -m gpu -x -q 2>&1 | tail -4
if [ "$2" = "ncu" ]; then timeout 600 ncu --set full --clock-control none --import-source on -k regex:tim_graph5 -s 3 -c 1 -o gpurun_out/$1/graph5 -f python tools/stage_times.py > gpurun_out/$1/ncu.log 2>&1; fi
