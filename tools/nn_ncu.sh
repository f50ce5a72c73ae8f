#!/bin/bash
# NN GEMM at a GPU-filling size: plain run (times), launch list, one ncu --set full capture of nn_gemm_kernel.
# tools/nn_ncu.sh <dir under gpurun_out>
D=gpurun_out/${1:-nn}
mkdir -p $D
timeout 600 python tools/prestep_roofline.py > $D/prestep_roofline.json 2> $D/prestep_roofline.err || { tail -5 $D/prestep_roofline.err; exit 1; }
cat $D/prestep_roofline.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $D/prestep_launches.csv python tools/prestep_roofline.py quick > $D/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:nn_gemm_kernel -s 2 -c 1 -o $D/nn_gemm_full -f python tools/prestep_roofline.py > $D/ncu_nn.log 2>&1
tail -2 $D/ncu_nn.log; ls -la $D
