#!/bin/bash
for mode in "--no-graph" "--graph"; do
timeout 300 python bench.py --batch 128 $mode --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-secondary 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('b128 $mode', d['ms_per_step'], d['gpu_launches'], d['config']['mode'][:40])
"; done
timeout 300 python bench.py --workload c5 --hidden 256 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('c5 h256 1gpu', d['ms_per_step'], d['gpu_launches'])
"
timeout 300 python bench.py --workload c5 --hidden 64 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('c5 h64 1gpu', d['ms_per_step'], d['gpu_launches'])
"
