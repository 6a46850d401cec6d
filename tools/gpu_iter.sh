#!/bin/bash
# Quick iteration: tc unit tests + bf16 parity + bench (+ optional ncu of the big GEMM kernels).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x --timeout 600 > gpurun_out/pytest_iter.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/pytest_iter.log; cat gpurun_out/tc_rates.txt 2>/dev/null
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench rc=$?"; tail -c 2200 gpurun_out/bench_iter.json; tail -3 gpurun_out/bench_iter.err
if [ "$1" = "ncu" ]; then
CMD="python bench.py --batch 256 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
cap() {
  timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$2" --launch-skip $3 --launch-count 1 -f -o gpurun_out/full_$1 $CMD > gpurun_out/ncu_full_$1.log 2>&1
  echo "ncu full $1 rc=$?"
}
cap gemm_fwd_silu 'tc_gemm_kernel<\(int\)256, \(int\)256, \(int\)1, \(bool\)1>' 3
cap gemm_bwd_dsilu 'tc_gemm_kernel<\(int\)256, \(int\)256, \(int\)0, \(bool\)1>' 9
cap wgrad_silu 'tc_wgrad_kernel<\(int\)256, \(int\)256, \(int\)1>' 2
fi
