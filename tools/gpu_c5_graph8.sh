#!/bin/bash
# config 5 on N GPUs: eager vs the whole T-step loop (NCCL all-gathers included) captured in one CUDA graph
N=${1:-8}
mkdir -p gpurun_out
for mode in "" "--graph"; do
for H in 256 64; do
T0=$(date +%s)
timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --workload c5 --hidden $H $mode --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/c5_n${N}_h${H}${mode}.json 2> gpurun_out/c5_n${N}_h${H}${mode}.err
echo "H=$H mode=[$mode] rc=$? wall=$(( $(date +%s) - T0 ))s"
python -c "
import json
try:
    d=json.loads([l for l in open('gpurun_out/c5_n${N}_h${H}${mode}.json') if l.startswith('{')][-1]); print(d['n_gpus'], d['ms_per_step'], d['value'], d['gpu_launches'])
except Exception as e: print('no line', e)
"
done; done
