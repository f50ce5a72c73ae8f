#!/bin/bash
# Why does `value` not scale? bench.py at N = 1 and N = 2 with the suspects removed one at a time.
D=gpurun_out/${1:-probe}
mkdir -p $D
run() { # name, nproc, env...
  name=$1; np=$2; shift 2
  if [ $np = 1 ]; then env "$@" python bench.py --steps 200 --no-legs --no-cpu-baseline > $D/$name.json 2> $D/$name.err
  else env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $np --steps 200 --no-legs --no-cpu-baseline > $D/$name.json 2> $D/$name.err; fi
  python - <<PY
import json
try:
    j=json.loads(open("$D/$name.json").read().strip().splitlines()[-1])
    print("$name", "value", round(j["value"]), "ms/step", round(j["ms_per_step"],4), "one_handle", round(j["value_one_handle"]), "e2e", round(j["e2e"]["value"]), "pageable", round(j["e2e_pageable_inputs"]["value"]), j["clocks"])
except Exception as e:
    print("$name failed", e); print(open("$D/$name.err").read()[-800:])
PY
}
nproc; free -g | head -2
run n2 2 A=1
python -c "
import json; j=json.loads(open('$D/n2.json').read().strip().splitlines()[-1]); print('per rank', j['per_rank_ms'])"
# rank 1 alone on cuda:1 with both devices visible
LOCAL_RANK=1 python bench.py --steps 200 --no-legs --no-cpu-baseline > $D/dev1.json 2> $D/dev1.err
python -c "
import json; j=json.loads(open('$D/dev1.json').read().strip().splitlines()[-1]); print('cuda:1 alone value', round(j['value']), 'e2e', round(j['e2e']['value']))"
# two plain processes, both devices visible to each, LOCAL_RANK picks the device
LOCAL_RANK=0 python bench.py --steps 200 --no-legs --no-cpu-baseline > $D/v0.json 2> $D/v0.err &
LOCAL_RANK=1 python bench.py --steps 200 --no-legs --no-cpu-baseline > $D/v1.json 2> $D/v1.err &
wait
python -c "
import json
for f in ('v0','v1'):
    j=json.loads(open('$D/'+f+'.json').read().strip().splitlines()[-1]); print(f, 'value', round(j['value']), 'e2e', round(j['e2e']['value']))"
