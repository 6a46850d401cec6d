#!/bin/bash
# whole GPU suite + smoke + the driver's bench command at N = 1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -4 gpurun_out/smoke.log
T0=$(date +%s)
timeout 1500 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err; echo "bench rc=$? wall=$(( $(date +%s) - T0 ))s"
tail -c 5000 gpurun_out/bench_full.json; tail -5 gpurun_out/bench_full.err
T0=$(date +%s)
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$? wall=$(( $(date +%s) - T0 ))s"
tail -c 700 gpurun_out/bench_ref.json
