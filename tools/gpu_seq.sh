#!/bin/bash
# seq2seq decoder path: GPU tests + timing of the captured rollout at the reference's experiment shape
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_seq2seq_gpu.py -q -m gpu --timeout 300 2>&1 | tail -40 | tee gpurun_out/seq_pytest.log
timeout 300 python tools/seq_bench.py 2>&1 | tee gpurun_out/seq_bench.txt
