#!/bin/bash
# ncu --set full of the two forward edge-MLP GEMMs of one layer at FULL config-4 size (roofline.traffic), after
# the plain run exited 0.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_full.log 2>&1 || { tail -5 gpurun_out/plain_full.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:tc_gemm_kernel<\(int\)256, \(int\)256, \(int\)1, \(bool\)1>' --launch-skip 2 --launch-count 2 -f -o gpurun_out/full_c4_fwd_layer $CMD > gpurun_out/ncu_full_c4_fwd_layer.log 2>&1
echo "ncu rc=$?"
timeout 900 ncu --set full --clock-control none --kernel-name-base demangled -k 'regex:segreduce_bf16_kernel|edge_features_hilo_kernel' --launch-skip 0 --launch-count 3 -f -o gpurun_out/full_c4_hbm_kernels $CMD > gpurun_out/ncu_full_c4_hbm.log 2>&1
echo "ncu rc=$?"
