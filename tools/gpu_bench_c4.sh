#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_c4_h256.json 2> gpurun_out/bench_c4_h256.err; echo "c4 h256 rc=$?"; tail -c 2500 gpurun_out/bench_c4_h256.json; tail -5 gpurun_out/bench_c4_h256.err
timeout 600 python bench.py --hidden 64 --steps 5 --warmup 3 > gpurun_out/bench_c4_h64.json 2> gpurun_out/bench_c4_h64.err; echo "c4 h64 rc=$?"; tail -c 2500 gpurun_out/bench_c4_h64.json
CMD="python bench.py --batch 64 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_c4_b64.csv $CMD > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu.log
