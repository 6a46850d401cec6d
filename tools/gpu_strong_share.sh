#!/bin/bash
# per-GPU share of config 4 under strong scaling at 8 GPUs (128 graphs), on one GPU: where the step time goes
mkdir -p gpurun_out
for b in 128 256; do
timeout 300 python bench.py --batch $b --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-secondary 2>gpurun_out/strong.err | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('c4 batch $b', round(d['ms_per_step'],3), d['gpu_launches'], d['clocks']['sm_mhz'])
        print({k: round(v*d['ms_per_step'],3) for k,v in d['time_share'].items()})
"; tail -2 gpurun_out/strong.err; done
timeout 300 python bench.py --batch 128 --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-secondary --no-graph 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('c4 batch 128 eager', round(d['ms_per_step'],3), d['gpu_launches'])
"
timeout 300 python bench.py --workload c2 --steps 20 --warmup 5 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('c2', round(d['ms_per_step'],3), d['gpu_launches'], d['value'])
        print({k: round(v*d['ms_per_step'],3) for k,v in (d['time_share'] or {}).items()})
"
