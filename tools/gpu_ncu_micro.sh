#!/bin/bash
mkdir -p gpurun_out
for spec in plain:1 b1:13; do
  n=${spec%%:*}; sk=${spec##*:}
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:tc_gemm_kernel --launch-skip $sk --launch-count 1 -f -o gpurun_out/micro_$n python tools/tc_bench.py > gpurun_out/ncu_micro_$n.log 2>&1
  echo "ncu micro $n rc=$?"
done
