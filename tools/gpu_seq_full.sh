#!/bin/bash
# --set full capture of the seq2seq GEMM launches of one eager rollout step (electrostatic shape)
mkdir -p gpurun_out
python - <<'PY'
import re
s = open('tools/gpu_seq_ncu.sh').read()
body = s[s.index("cat > /tmp/seq_one.py"):s.index("PY\n", s.index("cat > /tmp/seq_one.py")) + 3]
open('/tmp/mk.sh', 'w').write(body)
PY
bash /tmp/mk.sh
# launches of the last step: skip 2 steps x 7 gemm launches = 14, capture 7
timeout 600 ncu --set full --clock-control none --import-source on -k regex:seq_gemm_kernel --launch-skip 14 --launch-count 7 -f -o gpurun_out/full_seq_gemm python /tmp/seq_one.py > gpurun_out/ncu_full_seq.log 2>&1
echo rc=$?; ls -la gpurun_out/full_seq_gemm.ncu-rep
