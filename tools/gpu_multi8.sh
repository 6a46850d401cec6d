#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512"
timeout 600 $TR bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_n$N.json 2> gpurun_out/bench_c4_n$N.err; echo "c4 n$N rc=$?"; tail -c 700 gpurun_out/bench_c4_n$N.json | head -c 400; tail -2 gpurun_out/bench_c4_n$N.err
timeout 600 $TR bench.py --gpus $N --workload c5 --hidden 256 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_h256_n$N.json 2> gpurun_out/bench_c5_h256_n$N.err; echo "c5 h256 n$N rc=$?"; tail -2 gpurun_out/bench_c5_h256_n$N.err
