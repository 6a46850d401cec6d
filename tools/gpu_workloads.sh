#!/bin/bash
# The non-default bench workloads on one GPU (parity-test shapes; not the headline line).
mkdir -p gpurun_out
for w in c1 c2 c2b c3 c5; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --cpu-seconds 6 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "bench $w rc=$?"; tail -c 1800 gpurun_out/bench_$w.json; tail -3 gpurun_out/bench_$w.err
done
timeout 600 python bench.py --workload c5 --hidden 256 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_h256.json 2> gpurun_out/bench_c5_h256.err; echo "bench c5 h256 rc=$?"; tail -c 1500 gpurun_out/bench_c5_h256.json; tail -3 gpurun_out/bench_c5_h256.err
