"""Timing of the seq2seq decoder rollout (csrc/seqdec.cu) at the reference's experiment shapes: CUDA-graph replay of
T prediction steps, CUDA events, L2 flushed between replays.  Prints one JSON line per shape."""
import json
import sys
import os

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aether_b200.seq2seq import DecoderRollout, Seq2SeqAether  # noqa: E402


def run(D, N, B, H, K, Hf, T, reps=20, precision="fp32"):
    torch.manual_seed(0)
    p = {"num_vars": N, "input_size": 2 * D, "use_3d": D == 3, "decoder_hidden": H, "num_edge_types": K,
         "skip_first": False, "decoder_dropout": 0.0, "encoder_hidden": Hf, "gumbel_temp": 0.5}
    m = Seq2SeqAether(p, precision=precision).cuda().eval()
    E = N * (N - 1)
    x = torch.cat([torch.randn(B, N, D), 0.5 * torch.nn.functional.normalize(torch.randn(B, N, D), dim=-1)], -1).cuda()
    edges = torch.nn.functional.one_hot(torch.randint(0, K, (T, B, E)), K).float().cuda()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    out = {}
    for graph in (True, False):
        r = DecoderRollout(m, B, T, cuda_graph=graph)
        for _ in range(3):
            r.run(x, edges)
        ts = []
        for _ in range(reps):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            r.run(x, edges)
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ts.sort()
        out["graph" if graph else "eager"] = ts[len(ts) // 2]
    nparam = sum(q.numel() for q in m.decoder.parameters())
    flops = 2 * B * (E * (K * (2 * H * H + H * H + (2 * (4 * D + D * (D - 1) // 2) + 3 * D) * H + H * H)) +
                     N * (6 * H * H + 3 * (7 * D + D * (D - 1) // 2) * H + 2 * H * H + 2 * D * H + 2 * Hf * Hf + D * Hf))
    print(json.dumps({"shape": dict(D=D, N=N, B=B, H=H, K=K, Hf=Hf, T=T), "precision": precision, "ms_graph": out["graph"], "ms_eager": out["eager"],
                      "us_per_step_graph": 1e3 * out["graph"] / T, "particle_steps_per_s": B * N * T / (out["graph"] * 1e-3),
                      "decoder_params": nparam, "reference_gflop_per_step": flops / 1e9,
                      "tflops_on_reference_count": flops * T / (out["graph"] * 1e-3) / 1e12}))


if __name__ == "__main__":
    run(2, 5, 64, 512, 2, 512, 20)     # scripts/electrostatic_field_aether.sh, --test_pred_steps 20
    run(3, 5, 32, 512, 2, 512, 5)      # scripts/gravitational_field_3d_aether.sh, --test_pred_steps 5
    run(2, 5, 1024, 256, 2, 256, 20)   # a large evaluation batch
    run(2, 5, 64, 512, 2, 512, 20, precision="bf16")
    run(2, 5, 64, 512, 2, 512, 20, precision="fp32x3")
    run(2, 5, 1024, 256, 2, 256, 20, precision="fp32x3")
    run(2, 5, 1024, 256, 2, 256, 20, precision="bf16")


def run_full(D, N, B, H, K, He, R, T, reps=10, precision="fp32"):
    """The whole prediction loop (field -> encoder -> gumbel sample -> decoder) as one CUDA graph."""
    torch.manual_seed(0)
    p = {"num_vars": N, "input_size": 2 * D, "use_3d": D == 3, "decoder_hidden": H, "num_edge_types": K, "skip_first": False,
         "decoder_dropout": 0.0, "encoder_hidden": He, "gumbel_temp": 0.5, "encoder_rnn_hidden": R, "encoder_rnn_type": "lstm",
         "encoder_dropout": 0.0, "encoder_mlp_num_layers": 3, "encoder_mlp_hidden": 256, "prior_num_layers": 3,
         "prior_hidden_size": 256, "pos_representation": "polar"}
    m = Seq2SeqAether.with_encoder(p, precision=precision).cuda().eval()
    x = torch.cat([torch.randn(B, N, D), 0.5 * torch.nn.functional.normalize(torch.randn(B, N, D), dim=-1)], -1).cuda()
    r = DecoderRollout(m, B, T, cuda_graph=True)
    state = m.encoder.initial_state(B, x.device)
    hid = torch.zeros(B, N, H, device="cuda")
    for _ in range(3):
        r.run_sampled(x, hid, state)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); r.run_sampled(x, hid, state); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    ms = ts[len(ts) // 2]
    print(json.dumps({"whole_model_rollout": dict(D=D, N=N, B=B, H=H, K=K, He=He, R=R, T=T), "precision": precision, "ms_graph": ms,
                      "us_per_step": 1e3 * ms / T, "library_launches_per_rollout": r.launches_per_replay_sampled}))


if __name__ == "__main__":
    run_full(2, 5, 64, 512, 2, 512, 128, 20)    # electrostatic script: encoder_hidden 512, encoder_rnn_hidden 128
    run_full(2, 5, 64, 512, 2, 512, 128, 20, precision="bf16")
    run_full(2, 5, 64, 512, 2, 512, 128, 20, precision="fp32x3")
