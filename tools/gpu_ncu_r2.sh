#!/bin/bash
# Round 2: --set full captures of the chained forward kernel at full config-4 size (second step's launches:
# per step the chain runs layer 1 (K1 = 64, no gather) then layers 2-4 (gather)).
mkdir -p gpurun_out
FULL="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-secondary"
cap() { # name skip
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_chain_kernel --launch-skip $2 --launch-count 1 -f -o gpurun_out/full_$1 $FULL > gpurun_out/ncu_full_$1.log 2>&1
  echo "ncu full $1 rc=$?"
}
cap chain_gather 5
cap chain_layer1 4
ls -la gpurun_out | grep ncu-rep
