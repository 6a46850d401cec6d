#!/bin/bash
# single-launch rollout (csrc/small.cu): its tests, the kNN / golden rollout tests that now route through it, then the
# config-1 / config-3 bench lines (single launch, and --multi-launch for the path it replaces)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_small_gpu.py tests/test_abi.py -q -m gpu -x --timeout 300 > gpurun_out/pytest_small.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_small.log
tail -15 gpurun_out/pytest_small.log
timeout 600 python -m pytest tests/test_properties_gpu.py tests/test_step_gpu.py -q -m gpu -x --timeout 300 -k "rollout or knn" > gpurun_out/pytest_small2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_small2.log
tail -4 gpurun_out/pytest_small2.log
if [ "$1" = "memcheck" ]; then
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python -m pytest tests/test_small_gpu.py -q -m gpu -x -k "golden or edge_cases" > gpurun_out/memcheck_small.log 2>&1; echo "memcheck rc=$?"
grep -E "ERROR SUMMARY|Invalid|passed|failed" gpurun_out/memcheck_small.log | head -20
fi
for w in c1 c3; do for f in "" "--precision fp32" "--multi-launch"; do timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline --no-e2e $f 2> gpurun_out/bench_small.err | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$w $f', d['ms_per_step'], d['value'], d['gpu_launches'], d['roofline'].get('flops_view'), d['roofline'].get('multi_launch_graph_ms_per_step'))
        open('gpurun_out/bench_${w}$(echo $f | tr -d " -").json','w').write(l)
"; tail -2 gpurun_out/bench_small.err; done; done
