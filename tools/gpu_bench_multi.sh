#!/bin/bash
# the driver's multi-GPU bench command at N = $1
N=${1:-2}
mkdir -p gpurun_out
T0=$(date +%s)
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$? wall=$(( $(date +%s) - T0 ))s"
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_n$N.json') if l.startswith('{')][-1])
print(d['n_gpus'], d['ms_per_step'], d['value'])
for k,v in d.get('secondary',{}).items(): print(k, v['ms_per_step'], v['value'], v.get('collectives_per_step'), v['gpu_launches'])
PY
tail -5 gpurun_out/bench_n$N.err
