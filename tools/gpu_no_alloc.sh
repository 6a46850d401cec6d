#!/bin/bash
# the no-allocation entry point's test alone in a fresh process, then the whole GPU suite (it is collected last), then smoke
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -x --timeout 300 > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "alone rc=$?"
tail -5 gpurun_out/pytest_no_alloc_alone.log
timeout 1800 python -m pytest tests -q -m gpu -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
tail -4 gpurun_out/smoke.log
