#!/bin/bash
# One 8-GPU box: H2D bandwidth with 1 / 2 / 4 / 8 ranks copying at once, then bench.py at 4 and 8 GPUs.
D=gpurun_out/${1:-scale8}
mkdir -p $D
nvidia-smi topo -m > $D/topo.txt 2>&1; lscpu | head -25 > $D/lscpu.txt
for n in 1 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n tools/h2d_concurrent.py 2> $D/h2d_$n.err | grep "^{" >> $D/h2d_concurrent.jsonl
done
cat $D/h2d_concurrent.jsonl | python -c "
import sys, json
for l in sys.stdin:
    j = json.loads(l); print(j['world'], 'ranks: default', round(j['default_aggregate_GBps'],1), 'WC', round(j['write_combined_aggregate_GBps'],1), 'numa-bound', round(j['numa_bound_aggregate_GBps'],1), 'nodes', j.get('numa_nodes_online'), j['numa_node_of_gpu'])"
for n in 4 8; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n bench.py --gpus $n > $D/bench_n$n.json 2> $D/bench_n$n.err
  grep -v "^\[W\|Warning" $D/bench_n$n.err | tail -3
  python -c "
import json
j=json.loads(open('$D/bench_n$n.json').read().strip().splitlines()[-1])
print('N=$n value', round(j['value']), 'e2e', round(j['e2e']['value']), 'pageable', round(j['e2e_pageable_inputs']['value']), 'c2 f/s', round(j['config2']['frames_per_s']), 'c4 reg/s', round(j['config4']['registrations_per_s']), 'c5', {k: v.get('sharded_solve_ms') for k, v in j['config5']['sharded'].items()}, 'single', j['config5']['single_gpu_solve_ms'])
print(j['per_rank_ms'])"
done
