#!/bin/bash
# Round 2 re-entry: the whole GPU suite, smoke, the chained-kernel micro-benchmark, the default bench line.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -4 gpurun_out/smoke.log
timeout 300 python tools/chain_bench.py > gpurun_out/chain_bench.txt 2>&1; echo "chain bench rc=$?"
cat gpurun_out/chain_bench.txt
H=128 timeout 300 python tools/chain_bench.py > gpurun_out/chain_bench_h128.txt 2>&1; cat gpurun_out/chain_bench_h128.txt
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_c4_h256_bf16.json 2> gpurun_out/bench_c4_h256_bf16.err; echo "c4 h256 bf16 rc=$?"; tail -c 3000 gpurun_out/bench_c4_h256_bf16.json; tail -5 gpurun_out/bench_c4_h256_bf16.err
