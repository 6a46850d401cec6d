#!/bin/bash
# A/B of the AETHER_TC_REGSPLIT build of tc_gemm.cu (16 warps, setmaxnreg 200/56, gathered rows + silu' rows deeper ahead)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_optim_gpu.py -q -m gpu --timeout 300 2>&1 | tail -2
run() { timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-secondary --no-e2e 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$1', d['ms_per_step'], d['clocks']['sm_mhz'], {k: round(v*d['ms_per_step'],2) for k,v in d['time_share'].items()})
"; }
run default
AETHER_NVCC_FLAGS=-DAETHER_TC_REGSPLIT timeout 900 python -m aether_b200.build -f > gpurun_out/build_regsplit.txt 2>&1 || { echo "regsplit build failed"; tail -5 gpurun_out/build_regsplit.txt; }
timeout 900 python -m pytest tests/test_tc_gpu.py tests/test_bf16_gpu.py -q -m gpu -x --timeout 300 2>&1 | tail -3
run regsplit
run regsplit
timeout 900 python -m aether_b200.build -f > /dev/null 2>&1
run default
