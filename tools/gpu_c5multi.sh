#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"
for h in 256 64; do
timeout 600 $TR bench.py --gpus $N --workload c5 --hidden $h --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c5_h${h}_n$N.json 2> gpurun_out/bench_c5_h${h}_n$N.err; echo "c5 h$h n$N rc=$?"; tail -2 gpurun_out/bench_c5_h${h}_n$N.err | cut -c1-300
done
