#!/bin/bash
# chained edge-MLP kernel: parity (bit-identical to the two GEMMs), then the micro-benchmark
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_chain_gpu.py -q -m gpu -x --timeout 120 > gpurun_out/pytest_chain.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_chain.log
tail -25 gpurun_out/pytest_chain.log
timeout 300 python tools/chain_bench.py > gpurun_out/chain_bench.txt 2>&1; echo "bench rc=$?"
cat gpurun_out/chain_bench.txt
