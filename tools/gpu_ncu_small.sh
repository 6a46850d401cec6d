#!/bin/bash
# --set full captures of the single-launch rollout kernel at config 1 and config 3 sizes
mkdir -p gpurun_out
for w in c1 c3; do
CMD="python bench.py --workload $w --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:small_rollout_kernel --launch-skip 1 --launch-count 1 -f -o gpurun_out/full_small_$w $CMD > gpurun_out/ncu_full_small_$w.log 2>&1
echo "$w rc=$?"; ls -la gpurun_out/full_small_$w.ncu-rep
done
