"""Golden vectors of the seq2seq decoder path from the UNMODIFIED reference (nn/seq2seq/aether.py).

Run (only where /root/reference exists):
    python tests/golden/make_golden_seq2seq.py

Recorded, per case (``seq2seq*.npz``):
  * ``RecurrentDecoder.forward`` (nn/seq2seq/aether.py:589-654) on random states with SOFT edge-type weights: inputs,
    hidden state, edge weights, field, every decoder parameter -> outputs, new hidden state, and the three
    ``AugmentedLocalizer`` outputs the step consumes (nn/utils/augmented_global_to_local.py:51-68);
  * ``Aether.predict_future`` (nn/seq2seq/aether.py:155-191) of the whole reference model (real ``Encoder``, real
    gumbel sampling under a fixed seed) with ``return_everything``: the burn-in inputs, the one-hot edge types the
    encoder + sampler produced at every step (they are INPUTS of the decoder path: the encoder is not part of the
    path built here), the field network's parameters and Fourier matrix, and every prediction.

Stand-ins: ``torch_scatter.scatter`` as in make_golden.py, and ``Tensor.cuda()`` made the identity for the run (the
decoder calls ``self.recv_edges.cuda()``, nn/seq2seq/aether.py:617,636; this container has no GPU).  No reference
source is copied; only inputs and outputs are recorded.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import make_golden as mg  # noqa: E402


def params_dict(D, N, n_hid, enc_hid, n_types, skip_first, pos_rep="polar"):
    return {
        "num_vars": N, "input_size": 2 * D, "gpu": False, "use_3d": D == 3,
        "decoder_hidden": n_hid, "num_edge_types": n_types, "skip_first": skip_first, "decoder_dropout": 0.0,
        "decoder_type": "recurrent",
        "encoder_hidden": enc_hid, "encoder_rnn_hidden": 16, "encoder_rnn_type": "lstm", "encoder_dropout": 0.0,
        "encoder_mlp_num_layers": 3, "encoder_mlp_hidden": 16, "prior_num_layers": 3, "prior_hidden_size": 16,
        "gumbel_temp": 0.5, "train_hard_sample": False, "teacher_forcing_steps": -1, "normalize_kl": True,
        "normalize_nll": True, "kl_coef": 1.0, "nll_loss_type": "gaussian", "prior_variance": 5e-5,
        "add_uniform_prior": True, "no_edge_prior": 0.5, "rff_std": 1.0, "pos_representation": pos_rep,
    }


def synth_states(B, N, D, gen, zero_vel_first=False):
    pos = torch.randn(B, N, D, generator=gen)
    d = torch.randn(B, N, D, generator=gen)
    vel = 0.5 * d / d.norm(dim=-1, keepdim=True)
    if zero_vel_first:
        vel[0, 0] = 0.0
    return torch.cat([pos, vel], -1)


def run_case(name, D, N, B, n_hid, n_types, skip_first, seed, zero_vel_first=False, burn_in=4, pred_steps=3,
             pos_rep="polar"):
    from nn.seq2seq.aether import Aether
    torch.manual_seed(seed)
    np.random.seed(seed)
    p = params_dict(D, N, n_hid, 16, n_types, skip_first, pos_rep)
    model = Aether(p)
    model.eval()
    dec = model.decoder
    bn_gen = torch.Generator().manual_seed(seed + 50)
    for m in model.encoder.modules():        # fresh BatchNorm statistics are (0, 1): give them values so that the
        if isinstance(m, torch.nn.BatchNorm1d):   # recorded eval-mode outputs depend on them (state only, no code changed)
            m.running_mean.copy_(0.2 * torch.randn(m.num_features, generator=bn_gen))
            m.running_var.copy_(0.5 + torch.rand(m.num_features, generator=bn_gen))
            m.weight.data.copy_(1.0 + 0.1 * torch.randn(m.num_features, generator=bn_gen))
            m.bias.data.copy_(0.1 * torch.randn(m.num_features, generator=bn_gen))
    gen = torch.Generator().manual_seed(seed + 100)
    out = {"D": D, "N": N, "B": B, "n_hid": n_hid, "n_types": n_types, "skip_first": int(skip_first), "seed": seed}
    for k, v in dec.state_dict().items():
        out["dec." + k] = v.clone()
    for k, v in model.field_net.state_dict().items():
        out["field_net." + k] = v.clone()
    for k, v in model.encoder.state_dict().items():
        out["enc." + k] = v.clone()
    out["enc_pos_representation"] = p["pos_representation"]
    out["enc_rnn_hidden"], out["enc_mlp_hidden"], out["enc_prior_hidden"] = 16, 16, 16
    out["rff_B"] = model.coordinate_embedding.B.clone()

    # --- one decoder step with soft edge weights -----------------------------------------------
    E = N * (N - 1)
    inputs = synth_states(B, N, D, gen, zero_vel_first)
    hidden = 0.5 * torch.randn(B, N, n_hid, generator=gen)
    edges = torch.softmax(torch.randn(B, E, n_types, generator=gen), -1)
    field = 0.3 * torch.randn(B, N, D, generator=gen)
    with torch.no_grad():
        outputs, hidden_new = dec(inputs, hidden, edges, field)
        rel_feat, Rinv, edge_attr, edge_pos = dec.localizer(torch.cat([inputs, field], -1))
        field_pred, _ = model.predict_field(inputs)
    out.update(step_inputs=inputs, step_hidden=hidden, step_edges=edges, step_field=field, step_outputs=outputs,
               step_hidden_new=hidden_new, loc_rel_feat=rel_feat, loc_Rinv=Rinv, loc_edge_attr=edge_attr,
               field_pred=field_pred, send_edges=dec.send_edges.clone(), recv_edges=dec.recv_edges.clone())

    # --- predict_future of the whole model -----------------------------------------------------
    traj = [synth_states(B, N, D, gen)]
    for _ in range(burn_in - 1):   # a smooth-ish burn-in trajectory: x += v
        prev = traj[-1]
        nxt = prev.clone()
        nxt[..., :D] = prev[..., :D] + 0.1 * prev[..., D:]
        traj.append(nxt)
    burn = torch.stack(traj, 1)    # [B, T, N, 2D]
    with torch.no_grad():          # the encoder alone: burn-in pass (prior logits + LSTM state) and one more step
        x_t = burn[:, :-1].transpose(2, 1).contiguous()
        pfield, _ = model.predict_field(x_t)
        prior_logits, _, prior_state = model.encoder(burn[:, :-1], pfield)
        f_last, _ = model.predict_field(burn[:, -1])
        step_logits, step_state = model.encoder.single_step_forward(burn[:, -1], prior_state, f_last)
    out.update(enc_prior_logits=prior_logits, enc_state_h=prior_state[0], enc_state_c=prior_state[1],
               enc_step_logits=step_logits, enc_step_state_h=step_state[0], enc_step_state_c=step_state[1])
    torch.manual_seed(seed + 7)    # the gumbel noise
    with torch.no_grad():
        preds, all_edges = model.predict_future(burn, pred_steps, return_edges=True, return_everything=True)
    out.update(pf_inputs=burn, pf_edges=all_edges, pf_predictions=preds, pf_steps=pred_steps)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **mg._np(out))
    print(name, "step out", tuple(outputs.shape), "predict_future", tuple(preds.shape), "edges", tuple(all_edges.shape))


def main():
    mg._install_shims()
    torch.Tensor.cuda = lambda self, *a, **k: self   # see the module docstring
    run_case("seq2seq2d_h32", D=2, N=5, B=4, n_hid=32, n_types=2, skip_first=False, seed=1, zero_vel_first=True)
    run_case("seq2seq3d_h64", D=3, N=5, B=3, n_hid=64, n_types=3, skip_first=True, seed=2, pos_rep="cart")
    run_case("seq2seq2d_h48_n3", D=2, N=3, B=2, n_hid=48, n_types=2, skip_first=True, seed=3, burn_in=3, pred_steps=5)


if __name__ == "__main__":
    main()
