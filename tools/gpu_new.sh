#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_chain_gpu.py -q -m gpu -x --timeout 300 2>&1 | tail -6
timeout 300 python tools/chain_bench.py > gpurun_out/chain_bench.txt 2>&1; cat gpurun_out/chain_bench.txt
