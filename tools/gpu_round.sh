#!/bin/bash
# One GPU-box call: parity tests, smoke, the default bench line, ncu launch list and --set full captures.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1200 python -m pytest tests -q -m gpu --timeout 600 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_c4_h256_bf16.json 2> gpurun_out/bench_c4_h256_bf16.err; echo "c4 h256 bf16 rc=$?"; tail -c 3000 gpurun_out/bench_c4_h256_bf16.json; tail -5 gpurun_out/bench_c4_h256_bf16.err
timeout 600 python bench.py --hidden 64 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_h64_bf16.json 2> gpurun_out/bench_c4_h64_bf16.err; echo "c4 h64 bf16 rc=$?"; tail -c 3000 gpurun_out/bench_c4_h64_bf16.json
CMD="python bench.py --batch 64 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_c4_b64_bf16.csv $CMD > gpurun_out/ncu.log 2>&1
echo "ncu list rc=$?"; tail -3 gpurun_out/ncu.log
cap() { # name regex skip   (demangled names: template arguments print as (int)256 / (bool)1)
  timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$2" --launch-skip $3 --launch-count 1 -f -o gpurun_out/full_$1 $CMD > gpurun_out/ncu_full_$1.log 2>&1
  echo "ncu full $1 rc=$?"
}
if [ "$1" = "full" ]; then
CMD="python bench.py --batch 256 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
cap gemm_fwd_silu 'tc_gemm_kernel<\(int\)256, \(int\)256, \(int\)1, \(bool\)1>' 3
cap gemm_bwd_dsilu 'tc_gemm_kernel<\(int\)256, \(int\)256, \(int\)0, \(bool\)1>' 9
cap wgrad_silu 'tc_wgrad_kernel<\(int\)256, \(int\)256, \(int\)1>' 2
cap segmean 'segreduce_bf16_kernel' 1
cap feat 'edge_features_hilo_kernel' 0
fi
ls -la gpurun_out
