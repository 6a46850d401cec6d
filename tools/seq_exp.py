"""Diagnostic: decoder rollout timing with experiment builds of the library (AETHER_EXP_SO=path)."""
import os, sys, json, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aether_b200 import _lib
if os.environ.get("AETHER_EXP_SO"):
    _lib.SO_PATH = os.environ["AETHER_EXP_SO"]
from aether_b200.seq2seq import DecoderRollout, Seq2SeqAether
for precision in ("fp32", "bf16"):
    torch.manual_seed(0)
    D, N, B, H, K, Hf, T = 2, 5, 64, 512, 2, 512, 20
    p = {"num_vars": N, "input_size": 2 * D, "use_3d": False, "decoder_hidden": H, "num_edge_types": K, "skip_first": False,
         "decoder_dropout": 0.0, "encoder_hidden": Hf, "gumbel_temp": 0.5}
    m = Seq2SeqAether(p, precision=precision).cuda().eval()
    x = torch.cat([torch.randn(B, N, D), 0.5 * torch.nn.functional.normalize(torch.randn(B, N, D), dim=-1)], -1).cuda()
    edges = torch.nn.functional.one_hot(torch.randint(0, K, (T, B, N * (N - 1))), K).float().cuda()
    r = DecoderRollout(m, B, T, cuda_graph=True)
    for _ in range(3):
        r.run(x, edges)
    ts = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); r.run(x, edges); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    print(json.dumps({"so": os.path.basename(_lib.SO_PATH), "precision": precision, "us_per_step": 1e3 * ts[5] / T}))
