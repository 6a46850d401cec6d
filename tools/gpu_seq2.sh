#!/bin/bash
mkdir -p gpurun_out
bash tools/gpu_seq.sh
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -8 | tee gpurun_out/seq_smoke.log
bash tools/gpu_seq_ncu.sh
