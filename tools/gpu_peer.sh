#!/bin/bash
# usage: tools/gpu_peer.sh N   -- peer-memory exchange vs NCCL on N GPUs
N=${1:-2}
mkdir -p gpurun_out
for H in 256 64; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/peer_check.py $H 2>&1 | grep -v "^W\|^\*\*\*\|OMP_NUM" | tail -15 | tee -a gpurun_out/peer_check_$N.log
done
