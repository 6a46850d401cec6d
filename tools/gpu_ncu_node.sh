#!/bin/bash
# --set full capture of one fp32-output node GEMM (tc_gemm_kernel<256,256,0,0>) at config-4 size
mkdir -p gpurun_out
FULL="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-secondary"
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:tc_gemm_kernel<\(int\)256, \(int\)256, \(int\)0, \(bool\)0>" --launch-skip 12 --launch-count 1 -f -o gpurun_out/full_node_gemm $FULL > gpurun_out/ncu_full_node_gemm.log 2>&1
echo "rc=$?"; ls -la gpurun_out/full_node_gemm.ncu-rep
