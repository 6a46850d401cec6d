#!/bin/bash
# ncu --set full of the chained kernel (one launch each of the training+gather and the inference variants)
mkdir -p gpurun_out
timeout 300 python tools/chain_bench.py > gpurun_out/chain_bench.txt 2>&1; cat gpurun_out/chain_bench.txt
export M=$((128*148*16))
MODES=train_g timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_chain_kernel --launch-skip 2 --launch-count 1 -f -o gpurun_out/chain_train_g python tools/chain_bench.py > gpurun_out/ncu_chain1.log 2>&1
echo "ncu1 rc=$?"
MODES=infer timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_chain_kernel --launch-skip 2 --launch-count 1 -f -o gpurun_out/chain_infer python tools/chain_bench.py > gpurun_out/ncu_chain2.log 2>&1
echo "ncu2 rc=$?"
tail -3 gpurun_out/ncu_chain1.log
