#!/bin/bash
# launch list (gpu__time_duration) of an eager seq2seq rollout at the electrostatic shape
mkdir -p gpurun_out
cat > /tmp/seq_one.py <<'PY'
import sys, os, torch
sys.path.insert(0, os.getcwd())
from aether_b200.seq2seq import DecoderRollout, Seq2SeqAether
torch.manual_seed(0)
D, N, B, H, K, Hf, T = 2, 5, 64, 512, 2, 512, 3
p = {"num_vars": N, "input_size": 2 * D, "use_3d": False, "decoder_hidden": H, "num_edge_types": K, "skip_first": False,
     "decoder_dropout": 0.0, "encoder_hidden": Hf, "gumbel_temp": 0.5}
m = Seq2SeqAether(p, precision=os.environ.get("SEQ_PREC", "fp32")).cuda().eval()
E = N * (N - 1)
x = torch.cat([torch.randn(B, N, D), 0.5 * torch.nn.functional.normalize(torch.randn(B, N, D), dim=-1)], -1).cuda()
edges = torch.nn.functional.one_hot(torch.randint(0, K, (T, B, E)), K).float().cuda()
r = DecoderRollout(m, B, T, cuda_graph=False)
r.run(x, edges); torch.cuda.synchronize()
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 200 --csv --log-file gpurun_out/seq_launches.csv python /tmp/seq_one.py > gpurun_out/seq_ncu.log 2>&1
tail -3 gpurun_out/seq_ncu.log
