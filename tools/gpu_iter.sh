#!/bin/bash
# one iteration on the GPU box: the whole GPU suite, then the config-4 line (no baselines / secondary legs)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_iter.err
python -c "
import json
d=json.loads([l for l in open('gpurun_out/bench_iter.json') if l.startswith('{')][-1])
print(d['ms_per_step'], d['gpu_launches'], d['clocks']['sm_mhz'], d['roofline']['avg_ms_per_launch_group'])
print({k: round(v*d['ms_per_step'],2) for k,v in d['time_share'].items()})
"
for w in c1 c2 c3; do timeout 300 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$w', d['ms_per_step'], d['value'], d['gpu_launches'])
"; done
