#!/bin/bash
# Round 2 (final state): full-size launch list of the config-4 step and --set full captures of the three chained kernels
mkdir -p gpurun_out
FULL="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-secondary"
$FULL > gpurun_out/plain_full.log 2>&1 || { tail -5 gpurun_out/plain_full.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/launches_c4_full_bf16.csv $FULL > gpurun_out/ncu_fullsize.log 2>&1
echo "launch list rc=$?"
cap() { # name regex skip
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$2 --launch-skip $3 --launch-count 1 -f -o gpurun_out/full_$1 $FULL > gpurun_out/ncu_full_$1.log 2>&1
  echo "ncu full $1 rc=$?"
}
cap chain_bwd tc_chain_bwd_kernel 5
cap chain_fwd_gather 'tc_chain_kernel' 13
cap wgrad 'tc_wgrad_kernel' 40
ls -la gpurun_out | grep ncu-rep
