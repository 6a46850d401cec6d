#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_chain_gpu.py -q -m gpu -x --timeout 120 2>&1 | tail -3
timeout 300 python tools/chain_bench.py > gpurun_out/chain_bench.txt 2>&1; cat gpurun_out/chain_bench.txt
timeout 1500 python -m pytest tests -q -m gpu -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/bench_chain.json 2> gpurun_out/bench_chain.err; echo "chain rc=$?"
python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_chain.json') if l.startswith('{')][-1]); print(d['ms_per_step'], d['time_share'], d['roofline']['avg_ms_per_launch_group'], d['clocks'])
"
