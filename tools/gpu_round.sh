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
for k in tc_gemm_kernel tc_wgrad_kernel segreduce_bf16_kernel edge_features_hilo_kernel edge_dpre2_bf16_kernel; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 3 --launch-count 2 -f -o gpurun_out/full_$k $CMD > gpurun_out/ncu_full_$k.log 2>&1
  echo "ncu full $k rc=$?"
done
ls -la gpurun_out
