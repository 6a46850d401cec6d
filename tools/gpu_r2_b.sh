#!/bin/bash
# chain kernel integrated in the step: GPU suite, then the config-4 bench line with and without it
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_chain.json 2> gpurun_out/bench_chain.err; echo "chain rc=$?"; tail -c 1800 gpurun_out/bench_chain.json; tail -5 gpurun_out/bench_chain.err
AETHER_B200_NO_CHAIN=1 timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_nochain.json 2> gpurun_out/bench_nochain.err; echo "nochain rc=$?"; python -c "
import json,sys
for f in ['gpurun_out/bench_chain.json','gpurun_out/bench_nochain.json']:
    d=json.loads([l for l in open(f) if l.startswith('{')][-1]); print(f, d['ms_per_step'], d['time_share'], d['roofline']['avg_ms_per_launch_group'])
"
