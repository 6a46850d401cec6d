#!/bin/bash
# full-size launch list of the config-4 step (2 steps: warm-up + timed), ncu serialised cold-cache times
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_optim_gpu.py -q -m gpu --timeout 300 2>&1 | tail -2
FULL="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-secondary"
$FULL > gpurun_out/plain_full.log 2>&1 || { tail -5 gpurun_out/plain_full.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/launches_c4_full_bf16.csv $FULL > gpurun_out/ncu_fullsize.log 2>&1
echo "launch list rc=$?"
