#!/bin/bash
# the driver's bench command (N = 1): primary line + secondary legs + baselines + parity
mkdir -p gpurun_out
T0=$(date +%s)
timeout 1500 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "rc=$? wall=$(( $(date +%s) - T0 ))s"
tail -c 7000 gpurun_out/bench_full.json; tail -20 gpurun_out/bench_full.err
