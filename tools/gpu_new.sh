#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_chain_bwd_gpu.py tests/test_optim_gpu.py -q -m gpu -x --timeout 300 2>&1 | tail -15
