#!/bin/bash
# A/B of the experimental AETHER_TC_REGSPLIT build of tc_gemm.cu on the GPU box:
# micro-bench of the default build, then rebuild with the flag, GEMM parity tests, micro-bench, config-4 bench,
# and the default build is restored at the end.  Every step runs under its own timeout.
mkdir -p gpurun_out
HALF=1 timeout 300 python tools/tc_bench.py > gpurun_out/tcbench_default.txt 2>&1
AETHER_NVCC_FLAGS=-DAETHER_TC_REGSPLIT timeout 600 python -m aether_b200.build -f > gpurun_out/build_regsplit.txt 2>&1 || { echo "regsplit build failed"; python -m aether_b200.build -f; exit 1; }
timeout 600 python -m pytest tests/test_tc_gpu.py tests/test_bf16_gpu.py -q -m gpu -x --timeout 300 > gpurun_out/pytest_regsplit.txt 2>&1
echo "pytest rc=$?"
timeout 300 python tools/tc_bench.py > gpurun_out/tcbench_regsplit.txt 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_regsplit.txt 2>&1
# ring-depth sensitivity: 3 A slots instead of 4 at H = 256 (what a 16 KB operand buffer would cost)
AETHER_NVCC_FLAGS=-DAETHER_TC_MAX_SLOTS=3 timeout 600 python -m aether_b200.build -f > gpurun_out/build_slots3.txt 2>&1 && \
  timeout 300 python tools/tc_bench.py > gpurun_out/tcbench_slots3.txt 2>&1
timeout 600 python -m aether_b200.build -f > /dev/null 2>&1
tail -3 gpurun_out/pytest_regsplit.txt
tail -12 gpurun_out/tcbench_slots3.txt
paste gpurun_out/tcbench_default.txt gpurun_out/tcbench_regsplit.txt | tail -20
tail -1 gpurun_out/bench_regsplit.txt
