#!/bin/bash
# ncu captures of the edge-level (big) kernels + a full-size launch list.
mkdir -p gpurun_out
CMD="python bench.py --batch 256 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain256.log 2>&1 || { tail -5 gpurun_out/plain256.log; exit 1; }
cap() { # name regex skip
  timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$2" --launch-skip $3 --launch-count 1 -f -o gpurun_out/full_$1 $CMD > gpurun_out/ncu_full_$1.log 2>&1
  echo "ncu full $1 rc=$?"
}
cap gemm_fwd_silu 'tc_gemm_kernel<256, 256, 1, true>' 3
cap gemm_bwd_dsilu 'tc_gemm_kernel<256, 256, 0, true>' 9
cap wgrad_silu 'tc_wgrad_kernel<256, 256, 1>' 2
cap segmean 'segreduce_bf16_kernel' 1
cap feat 'edge_features_hilo_kernel' 0
cap dpre2 'edge_dpre2_bf16_kernel' 0
cap featbwd 'edge_features_bwd_kernel' 0
FULL="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/launches_c4_full_bf16.csv $FULL > gpurun_out/ncu_fullsize.log 2>&1
echo "launch list rc=$?"
ls -la gpurun_out | grep -c ncu-rep
