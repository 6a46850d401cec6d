#!/bin/bash
# usage: tools/gpu_multi.sh N  -- data-parallel c4 and receiver-partitioned c5 on N GPUs of one box
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_n$N.json 2> gpurun_out/bench_c4_n$N.err; echo "c4 n$N rc=$?"; tail -c 1600 gpurun_out/bench_c4_n$N.json; tail -3 gpurun_out/bench_c4_n$N.err
timeout 600 $TR bench.py --gpus $N --workload c5 --hidden 256 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_h256_n$N.json 2> gpurun_out/bench_c5_h256_n$N.err; echo "c5 h256 n$N rc=$?"; tail -c 1500 gpurun_out/bench_c5_h256_n$N.json; tail -3 gpurun_out/bench_c5_h256_n$N.err
timeout 600 $TR bench.py --gpus $N --workload c5 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err; echo "c5 n$N rc=$?"; tail -c 600 gpurun_out/bench_c5_n$N.json
timeout 600 $TR bench.py --gpus $N --workload c3 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3_n$N.json 2> gpurun_out/bench_c3_n$N.err; echo "c3 n$N rc=$?"; tail -c 600 gpurun_out/bench_c3_n$N.json
timeout 600 $TR bench.py --gpus $N --impl reference --steps 1 --warmup 1 | cut -c1-300
