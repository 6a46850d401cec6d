#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_optim_gpu.py tests/test_dynfield_gpu.py -q -m gpu --timeout 300 2>&1 | tail -25
