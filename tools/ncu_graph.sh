#!/bin/bash
# One ncu --set full capture of the tensor-core graph kernel (and optionally the fused tail) on a config-3 batch.
# tools/ncu_graph.sh <dir under gpurun_out> [kernel regex]
D=gpurun_out/${1:-ncu}
K=${2:-tim_graph6}
mkdir -p $D
timeout 300 python tools/stage_times.py 64 5000 0.95 > $D/plain.log 2>&1 || { tail -5 $D/plain.log; exit 1; }
tail -6 $D/plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -o $D/$K -f python tools/stage_times.py 64 5000 0.95 > $D/ncu_$K.log 2>&1
tail -3 $D/ncu_$K.log; ls -la $D
