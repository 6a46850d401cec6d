#!/bin/bash
# per-rank share of config 5 at 8 GPUs, without the exchange: a 12.5k-particle scene (same density) on one GPU
for H in 256 64; do
timeout 300 python bench.py --workload c5 --batch 12500 --hidden $H --steps 20 --warmup 5 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('c5 12.5k h$H', d['ms_per_step'], d['gpu_launches'], d['config']['edges'])
"; done
CMD="python bench.py --workload c5 --batch 12500 --hidden 256 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-graph"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_c5_12k.csv $CMD > gpurun_out/ncu_c5.log 2>&1; echo "ncu rc=$?"
