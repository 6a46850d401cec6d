"""seq2seq decoder path on the GPU (aether_b200/seq2seq.py -> csrc/seqdec.cu through the C ABI) against fixtures recorded
from the unmodified reference model (tests/golden/make_golden_seq2seq.py) and against the oracle at the reference's own
experiment shape (5 particles, 2 edge types, decoder_hidden up to 512).  fp32: 1e-5 relative to each tensor's scale for
single steps; the multi-step rollouts compound step errors through the recurrent state, bar 1e-4."""
import os

import pytest
import torch

from conftest import load_golden, rel_err

pytestmark = pytest.mark.gpu
CASES = ["seq2seq2d_h32", "seq2seq3d_h64", "seq2seq2d_h48_n3"]


def params_of(g, enc_hidden=16):
    D = int(g["D"])
    return {"num_vars": int(g["N"]), "input_size": 2 * D, "use_3d": D == 3, "decoder_hidden": int(g["n_hid"]),
            "num_edge_types": int(g["n_types"]), "skip_first": bool(g["skip_first"]), "decoder_dropout": 0.0,
            "encoder_hidden": enc_hidden, "gumbel_temp": 0.5, "rff_std": 1.0}


def build(g):
    from aether_b200.seq2seq import Seq2SeqAether
    m = Seq2SeqAether(params_of(g))
    sd = {"decoder." + k[4:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("dec.")}
    sd.update({k: torch.from_numpy(v) for k, v in g.items() if k.startswith("field_net.")})
    sd["coordinate_embedding.B"] = torch.from_numpy(g["rff_B"])
    m.load_state_dict(sd)          # strict: every reference key of the decoder side exists here and vice versa
    return m.cuda().eval()


def dev(g, k):
    return torch.from_numpy(g[k]).cuda()


@pytest.mark.parametrize("name", CASES)
def test_localizer_matches_reference(name):
    g = load_golden(name)
    m = build(g)
    rel, Rinv, ea, pos = m.decoder.localizer(torch.cat([dev(g, "step_inputs"), dev(g, "step_field")], -1))
    assert (rel.cpu() - torch.from_numpy(g["loc_rel_feat"])).abs().max() <= 1e-5
    assert (Rinv.cpu() - torch.from_numpy(g["loc_Rinv"])).abs().max() <= 1e-6
    assert (ea.cpu() - torch.from_numpy(g["loc_edge_attr"])).abs().max() <= 1e-5


@pytest.mark.parametrize("name", CASES)
def test_decoder_step_matches_reference(name):
    g = load_golden(name)
    m = build(g)
    out, hid = m.decoder(dev(g, "step_inputs"), dev(g, "step_hidden"), dev(g, "step_edges"), dev(g, "step_field"))
    assert rel_err(out.cpu(), g["step_outputs"]) <= 1e-5
    assert rel_err(hid.cpu(), g["step_hidden_new"]) <= 1e-5
    f, coords = m.predict_field(dev(g, "step_inputs"))
    assert rel_err(f.cpu(), g["field_pred"]) <= 2e-5
    assert torch.equal(coords.cpu(), torch.from_numpy(g["step_inputs"])[..., :int(g["D"])])


@pytest.mark.parametrize("name", CASES)
def test_predict_future_matches_reference(name):
    """The whole predict_future of the reference model (burn-in + prediction steps); the encoder's sampled edge types
    are the recorded ones."""
    g = load_golden(name)
    m = build(g)
    preds, edges = m.predict_future(dev(g, "pf_inputs"), int(g["pf_steps"]), return_edges=True, return_everything=True,
                                    edges=dev(g, "pf_edges"))
    assert preds.shape == g["pf_predictions"].shape and torch.equal(edges.cpu(), torch.from_numpy(g["pf_edges"]))
    assert rel_err(preds.cpu(), g["pf_predictions"]) <= 1e-4
    only_future = m.predict_future(dev(g, "pf_inputs"), int(g["pf_steps"]), edges=dev(g, "pf_edges"))
    assert torch.equal(only_future, preds[:, -int(g["pf_steps"]):])


def _oracle_weights(m):
    w = {k[len("decoder."):]: v.detach().cpu() for k, v in m.state_dict().items() if k.startswith("decoder.")}
    w.update({k: v.detach().cpu() for k, v in m.state_dict().items() if k.startswith("field_net.")})
    w["rff_B"] = m.coordinate_embedding.B.detach().cpu()
    return w


@pytest.mark.parametrize("D,N,B,H,K,skip,Hf", [(2, 5, 64, 512, 2, False, 512), (3, 5, 32, 256, 2, True, 128),
                                               (2, 2, 3, 36, 1, False, 20), (3, 7, 5, 64, 4, True, 64),
                                               (2, 4, 6, 32, 1, True, 16), (2, 3, 4, 32, 4, False, 16)])
def test_step_and_rollout_against_oracle(D, N, B, H, K, skip, Hf):
    """Reference experiment shapes (scripts/electrostatic_field_aether.sh: 5 particles, batch 64, decoder_hidden 512,
    2 edge types; gravitational 3D: batch 32) and edge cases: 2 particles, one edge type, no active type at all
    (skip_first with one type: only the GRU inputs drive the state), widths that are not multiples of the tile."""
    from aether_b200.seq2seq import DecoderRollout, Seq2SeqAether
    from oracle import seq2seq_oracle as so
    torch.manual_seed(11)
    p = {"num_vars": N, "input_size": 2 * D, "use_3d": D == 3, "decoder_hidden": H, "num_edge_types": K,
         "skip_first": skip, "decoder_dropout": 0.0, "encoder_hidden": Hf, "gumbel_temp": 0.5}
    m = Seq2SeqAether(p).cuda().eval()
    w = _oracle_weights(m)
    E, T = N * (N - 1), 4
    x = torch.cat([torch.randn(B, N, D), 0.5 * torch.nn.functional.normalize(torch.randn(B, N, D), dim=-1)], -1)
    hid = 0.5 * torch.randn(B, N, H)
    soft = torch.softmax(torch.randn(B, E, K), -1)
    field = 0.3 * torch.randn(B, N, D)
    out, h2 = m.decoder(x.cuda(), hid.cuda(), soft.cuda(), field.cuda())
    o_ref, h_ref = so.decoder_step(w, x, hid, soft, field, D, K, skip)
    assert rel_err(out.cpu(), o_ref) <= 1e-5 and rel_err(h2.cpu(), h_ref) <= 1e-5
    # prediction steps with one-hot edge types, captured as one CUDA graph
    onehot = torch.nn.functional.one_hot(torch.randint(0, K, (T, B, E)), K).float()
    r = DecoderRollout(m, B, T, cuda_graph=True)
    traj = r.run(x.cuda(), onehot.cuda()).clone()
    traj2 = r.run(x.cuda(), onehot.cuda()).clone()                 # replay
    eager = DecoderRollout(m, B, T, cuda_graph=False).run(x.cuda(), onehot.cuda())
    assert torch.equal(traj, traj2) and torch.equal(traj, eager)
    # (four active edge types: batch A goes out in two launches)
    assert r.launches_per_replay == (11 if K - int(skip) == 0 else (15 if K - int(skip) == 4 else 14)) * T
    burn = x.unsqueeze(1)
    ref = so.predict_future(w, burn, onehot.transpose(0, 1), T, D, K, skip, return_everything=False)
    assert rel_err(traj.transpose(0, 1).cpu(), ref) <= 1e-4


def test_module_surface_and_errors():
    from aether_b200.seq2seq import RecurrentDecoder, Seq2SeqAether
    g = load_golden("seq2seq2d_h32")
    m = build(g)
    dec = m.decoder
    x = dev(g, "step_inputs")
    assert dec.get_initial_hidden(dev(g, "pf_inputs")).shape == (4, 5, 32)
    with pytest.raises(RuntimeError):
        dec(x.cpu(), dev(g, "step_hidden").cpu(), dev(g, "step_edges").cpu(), dev(g, "step_field").cpu())
    with pytest.raises(RuntimeError):
        dec(x, dev(g, "step_hidden")[:, :, :16], dev(g, "step_edges"), dev(g, "step_field"))
    with pytest.raises(RuntimeError):
        m.predict_future(dev(g, "pf_inputs"), 2)          # no encoder, no edges

    class Enc(torch.nn.Module):                            # the reference Encoder's call surface, logits from positions
        def forward(self, inputs, predicted_field):
            B, T, N, _ = inputs.shape
            lg = inputs[:, :, :, :1].repeat_interleave(N - 1, 2).repeat(1, 1, 1, 2) * torch.tensor([1.0, -1.0], device=inputs.device)
            return lg, lg, (torch.zeros(1), torch.zeros(1))

        def single_step_forward(self, inputs, prior_state, predicted_field):
            N = inputs.shape[1]
            lg = inputs[:, :, :1].repeat_interleave(N - 1, 1).repeat(1, 1, 2) * torch.tensor([1.0, -1.0], device=inputs.device)
            return lg, prior_state

    m.encoder = Enc()
    torch.manual_seed(3)
    preds, edges = m.predict_future(dev(g, "pf_inputs"), 3, return_edges=True)
    assert preds.shape == (4, 3, 5, 4) and edges.shape == (4, 3, 20, 2)
    assert bool(((edges == 0) | (edges == 1)).all()) and torch.equal(edges.sum(-1), torch.ones(4, 3, 20, device="cuda"))
    again = m.predict_future(dev(g, "pf_inputs"), 3, edges=torch.cat([torch.zeros(4, 3, 20, 2, device="cuda"), edges], 1))
    assert torch.isfinite(preds).all() and again.shape == preds.shape


# ---- edge-type encoder, prior branch ------------------------------------------------------------------------------------
def enc_params_of(g):
    p = params_of(g)
    p.update({"encoder_rnn_hidden": int(g["enc_rnn_hidden"]), "encoder_rnn_type": "lstm", "encoder_dropout": 0.0,
              "encoder_mlp_num_layers": 3, "encoder_mlp_hidden": int(g["enc_mlp_hidden"]), "prior_num_layers": 3,
              "prior_hidden_size": int(g["enc_prior_hidden"]), "pos_representation": str(g["enc_pos_representation"])})
    return p


def build_full(g):
    from aether_b200.seq2seq import Seq2SeqAether
    m = Seq2SeqAether.with_encoder(enc_params_of(g))
    sd = {"decoder." + k[4:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("dec.")}
    sd.update({"encoder." + k[4:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("enc.")})
    sd.update({k: torch.from_numpy(v) for k, v in g.items() if k.startswith("field_net.")})
    sd["coordinate_embedding.B"] = torch.from_numpy(g["rff_B"])
    m.load_state_dict(sd)          # strict: the reference model's whole state_dict
    return m.cuda().eval()


@pytest.mark.parametrize("name", CASES)
def test_encoder_matches_reference(name):
    """Burn-in pass of the reference Encoder (prior logits of every step, final LSTM state) and one single_step_forward,
    with non-trivial BatchNorm running statistics."""
    g = load_golden(name)
    m = build_full(g)
    burn = dev(g, "pf_inputs")
    field, _ = m.predict_field(burn[:, :-1].transpose(2, 1).contiguous())
    logits, post, (h, c) = m.encoder(burn[:, :-1], field)
    assert post is None and logits.shape == g["enc_prior_logits"].shape and h.shape == g["enc_state_h"].shape
    assert rel_err(logits.cpu(), g["enc_prior_logits"]) <= 1e-5
    assert rel_err(h.cpu(), g["enc_state_h"]) <= 1e-5 and rel_err(c.cpu(), g["enc_state_c"]) <= 1e-5
    f_last, _ = m.predict_field(burn[:, -1])
    lg, (h2, c2) = m.encoder.single_step_forward(burn[:, -1], (h, c), f_last)
    assert rel_err(lg.cpu(), g["enc_step_logits"]) <= 1e-5
    assert rel_err(h2.cpu(), g["enc_step_state_h"]) <= 1e-5 and rel_err(c2.cpu(), g["enc_step_state_c"]) <= 1e-5


def test_predict_future_whole_model_on_device():
    """predict_future as the reference calls it (no edges given): encoder, gumbel hard samples, field, decoder all on the
    GPU.  The random stream differs from the reference's CPU torch.rand, so the check is structural + the decoder side
    re-run on the sampled edges."""
    g = load_golden("seq2seq2d_h32")
    m = build_full(g)
    burn = dev(g, "pf_inputs")
    torch.manual_seed(5)
    preds, edges = m.predict_future(burn, 4, return_edges=True, return_everything=True)
    assert preds.shape == (4, 3 + 4, 5, 4) and edges.shape == (4, 3 + 4, 20, 2)
    assert bool(((edges == 0) | (edges == 1)).all()) and bool((edges.sum(-1) == 1).all())
    torch.manual_seed(5)
    again = m.predict_future(burn, 4, return_everything=True)
    assert torch.equal(again, preds)
    replay = m.predict_future(burn, 4, return_everything=True, edges=edges)
    assert torch.equal(replay, preds)
    # with the noise switched off (eps so large that every sample is the arg-max of the logits), the edges follow the
    # encoder's logits: compare with the oracle's arg-max on the burn-in steps
    from oracle import seq2seq_oracle as so
    w = {k[4:]: torch.from_numpy(v) for k, v in g.items() if k.startswith("enc.")}
    w.update({k: torch.from_numpy(v) for k, v in g.items() if k.startswith("field_net.")})
    w["rff_B"] = torch.from_numpy(g["rff_B"])
    lg, _, _ = so.encoder_prior(w, torch.from_numpy(g["pf_inputs"])[:, :-1], 2, str(g["enc_pos_representation"]))
    field, _ = m.predict_field(burn[:, :-1].transpose(2, 1).contiguous())
    mine, _, _ = m.encoder(burn[:, :-1], field)
    margin = (lg[..., 0] - lg[..., 1]).abs() > 1e-4
    assert torch.equal(mine.argmax(-1).cpu()[margin], lg.argmax(-1)[margin])


@pytest.mark.parametrize("D,N,B,He,R,PH,layers,rep", [(2, 5, 16, 128, 64, 32, 3, "polar"), (3, 4, 5, 64, 32, 0, 1, "polar"),
                                                      (3, 5, 3, 32, 16, 24, 4, "cart"), (2, 2, 3, 36, 20, 12, 2, "cart")])
def test_encoder_against_oracle(D, N, B, He, R, PH, layers, rep):
    from aether_b200.seq2seq import Encoder
    from oracle import seq2seq_oracle as so
    torch.manual_seed(21)
    p = {"num_vars": N, "input_size": 2 * D, "use_3d": D == 3, "num_edge_types": 3, "encoder_hidden": He,
         "encoder_rnn_hidden": R, "encoder_rnn_type": "lstm", "encoder_dropout": 0.0, "encoder_mlp_num_layers": 2,
         "encoder_mlp_hidden": 16, "prior_num_layers": layers, "prior_hidden_size": PH, "pos_representation": rep}
    enc = Encoder(p)
    for mod in enc.modules():
        if isinstance(mod, torch.nn.BatchNorm1d):
            mod.running_mean.normal_(0, 0.2)
            mod.running_var.uniform_(0.5, 1.5)
            mod.weight.data.normal_(1.0, 0.1)
            mod.bias.data.normal_(0, 0.1)
    enc = enc.cuda().eval()
    w = {k: v.detach().cpu() for k, v in enc.state_dict().items()}
    E = N * (N - 1)
    x = torch.cat([torch.randn(B, N, D), 0.5 * torch.nn.functional.normalize(torch.randn(B, N, D), dim=-1)], -1)
    field = 0.3 * torch.randn(B, N, D)
    h, c = 0.3 * torch.randn(B * E, R), 0.3 * torch.randn(B * E, R)
    lg, (h2, c2) = enc.single_step_forward(x.cuda(), (h.cuda().unsqueeze(0), c.cuda().unsqueeze(0)), field.cuda())
    lo, ho, co = so.encoder_step(w, x, field, h, c, D, rep)
    assert rel_err(lg.cpu(), lo) <= 2e-5
    assert rel_err(h2.cpu()[0], ho) <= 1e-5 and rel_err(c2.cpu()[0], co) <= 1e-5


def test_sampled_rollout_graph_matches_stepwise():
    """DecoderRollout.run_sampled: field -> encoder -> sample -> decoder for T steps as ONE CUDA graph; replaying the
    decoder side on the edges it sampled reproduces its trajectory bit for bit, eager and captured."""
    from aether_b200.seq2seq import DecoderRollout
    g = load_golden("seq2seq3d_h64")
    m = build_full(g)
    burn = dev(g, "pf_inputs")
    B, T = burn.shape[0], 4
    field, _ = m.predict_field(burn[:, :-1].transpose(2, 1).contiguous())
    _, _, state = m.encoder(burn[:, :-1], field)
    hid = 0.1 * torch.randn(B, 5, 64, device="cuda")
    for graph in (False, True):
        r = DecoderRollout(m, B, T, cuda_graph=graph)
        for _ in range(2):                                  # second call = replay of the captured graph
            traj = r.run_sampled(burn[:, -1], hid, state).clone()
            edges = r.edges.clone()
            assert bool((edges.sum(-1) == 1).all()) and torch.isfinite(traj).all()
            again = DecoderRollout(m, B, T, cuda_graph=False).run(burn[:, -1], edges, hidden0=hid)
            assert torch.equal(again, traj)


# ---- tensor-core mode (precision="bf16"): the bf16 tolerance of the other paths, 2e-2 relative to each tensor's scale -----
@pytest.mark.parametrize("name", CASES)
def test_bf16_mode_matches_reference(name):
    from aether_b200.seq2seq import Seq2SeqAether
    g = load_golden(name)
    m = Seq2SeqAether.with_encoder(enc_params_of(g), precision="bf16")
    m.load_state_dict(build_full(g).state_dict())
    m = m.cuda().eval()
    out, hid = m.decoder(dev(g, "step_inputs"), dev(g, "step_hidden"), dev(g, "step_edges"), dev(g, "step_field"))
    disp = out.cpu() - torch.from_numpy(g["step_inputs"])          # what the step adds to its input
    ref_disp = torch.from_numpy(g["step_outputs"] - g["step_inputs"])
    assert rel_err(disp, ref_disp) <= 2e-2 and rel_err(hid.cpu(), g["step_hidden_new"]) <= 2e-2
    f, _ = m.predict_field(dev(g, "step_inputs"))
    assert rel_err(f.cpu(), g["field_pred"]) <= 2e-2
    burn = dev(g, "pf_inputs")
    field, _ = m.predict_field(burn[:, :-1].transpose(2, 1).contiguous())
    logits, _, (h, c) = m.encoder(burn[:, :-1], field)
    assert rel_err(logits.cpu(), g["enc_prior_logits"]) <= 2e-2
    assert rel_err(h.cpu(), g["enc_state_h"]) <= 2e-2 and rel_err(c.cpu(), g["enc_state_c"]) <= 2e-2
    preds = m.predict_future(burn, int(g["pf_steps"]), return_everything=True, edges=dev(g, "pf_edges"))
    assert rel_err(preds.cpu(), g["pf_predictions"]) <= 2e-2


def test_bf16_mode_at_experiment_shape_against_oracle():
    from aether_b200.seq2seq import DecoderRollout, Seq2SeqAether
    from oracle import seq2seq_oracle as so
    torch.manual_seed(11)
    D, N, B, H, K, T = 2, 5, 64, 512, 2, 4
    p = {"num_vars": N, "input_size": 2 * D, "use_3d": False, "decoder_hidden": H, "num_edge_types": K,
         "skip_first": False, "decoder_dropout": 0.0, "encoder_hidden": H, "gumbel_temp": 0.5}
    m = Seq2SeqAether(p, precision="bf16").cuda().eval()
    w = _oracle_weights(m)
    E = N * (N - 1)
    x = torch.cat([torch.randn(B, N, D), 0.5 * torch.nn.functional.normalize(torch.randn(B, N, D), dim=-1)], -1)
    onehot = torch.nn.functional.one_hot(torch.randint(0, K, (T, B, E)), K).float()
    r = DecoderRollout(m, B, T, cuda_graph=True)
    traj = r.run(x.cuda(), onehot.cuda()).clone()
    assert torch.equal(traj, r.run(x.cuda(), onehot.cuda()))
    ref = so.predict_future(w, x.unsqueeze(1), onehot.transpose(0, 1), T, D, K, False, return_everything=False)
    step = traj.transpose(0, 1).cpu() - torch.cat([x.unsqueeze(1), ref[:, :-1]], 1)      # per-step displacement vs oracle's
    ref_step = ref - torch.cat([x.unsqueeze(1), ref[:, :-1]], 1)
    assert rel_err(traj.transpose(0, 1).cpu(), ref) <= 2e-2
    assert rel_err(step[:, 0], ref_step[:, 0]) <= 2e-2


# ---- fp32x3 mode: tensor cores at the fp32 tolerance (each operand split into three bf16, six products) -----------------
@pytest.mark.parametrize("name", CASES)
def test_fp32x3_mode_meets_the_fp32_bars(name):
    from aether_b200.seq2seq import Seq2SeqAether
    g = load_golden(name)
    m = Seq2SeqAether.with_encoder(enc_params_of(g), precision="fp32x3")
    m.load_state_dict(build_full(g).state_dict())
    m = m.cuda().eval()
    out, hid = m.decoder(dev(g, "step_inputs"), dev(g, "step_hidden"), dev(g, "step_edges"), dev(g, "step_field"))
    assert rel_err(out.cpu(), g["step_outputs"]) <= 1e-5 and rel_err(hid.cpu(), g["step_hidden_new"]) <= 1e-5
    f, _ = m.predict_field(dev(g, "step_inputs"))
    assert rel_err(f.cpu(), g["field_pred"]) <= 2e-5
    burn = dev(g, "pf_inputs")
    field, _ = m.predict_field(burn[:, :-1].transpose(2, 1).contiguous())
    logits, _, (h, c) = m.encoder(burn[:, :-1], field)
    assert rel_err(logits.cpu(), g["enc_prior_logits"]) <= 1e-5
    assert rel_err(h.cpu(), g["enc_state_h"]) <= 1e-5 and rel_err(c.cpu(), g["enc_state_c"]) <= 1e-5
    preds = m.predict_future(burn, int(g["pf_steps"]), return_everything=True, edges=dev(g, "pf_edges"))
    assert rel_err(preds.cpu(), g["pf_predictions"]) <= 1e-4


def test_fp32x3_mode_at_experiment_shape_against_oracle():
    from aether_b200.seq2seq import DecoderRollout, Seq2SeqAether
    from oracle import seq2seq_oracle as so
    torch.manual_seed(11)
    D, N, B, H, K, T = 2, 5, 64, 512, 2, 4
    p = {"num_vars": N, "input_size": 2 * D, "use_3d": False, "decoder_hidden": H, "num_edge_types": K,
         "skip_first": False, "decoder_dropout": 0.0, "encoder_hidden": H, "gumbel_temp": 0.5}
    m = Seq2SeqAether(p, precision="fp32x3").cuda().eval()
    w = _oracle_weights(m)
    E = N * (N - 1)
    x = torch.cat([torch.randn(B, N, D), 0.5 * torch.nn.functional.normalize(torch.randn(B, N, D), dim=-1)], -1)
    hid = 0.5 * torch.randn(B, N, H)
    soft = torch.softmax(torch.randn(B, E, K), -1)
    field = 0.3 * torch.randn(B, N, D)
    out, h2 = m.decoder(x.cuda(), hid.cuda(), soft.cuda(), field.cuda())
    o_ref, h_ref = so.decoder_step(w, x, hid, soft, field, D, K, False)
    assert rel_err(out.cpu(), o_ref) <= 1e-5 and rel_err(h2.cpu(), h_ref) <= 1e-5
    onehot = torch.nn.functional.one_hot(torch.randint(0, K, (T, B, E)), K).float()
    traj = DecoderRollout(m, B, T, cuda_graph=True).run(x.cuda(), onehot.cuda())
    ref = so.predict_future(w, x.unsqueeze(1), onehot.transpose(0, 1), T, D, K, False, return_everything=False)
    assert rel_err(traj.transpose(0, 1).cpu(), ref) <= 1e-4


def test_hard_sample_kernel_matches_torch():
    from aether_b200.seq2seq import sample_hard
    torch.manual_seed(9)
    for K in (1, 2, 3, 4):
        lg = torch.randn(1000, K, device="cuda")
        U = torch.rand(1000, K, device="cuda")
        got = sample_hard(lg, U)
        y = lg - torch.log(1e-10 - torch.log(U + 1e-10))
        want = torch.zeros_like(lg).scatter_(-1, y.argmax(-1, keepdim=True), 1.0)
        top2 = y.topk(min(2, K), dim=-1).values
        clear = torch.ones(1000, dtype=torch.bool, device="cuda") if K == 1 else (top2[:, 0] - top2[:, 1]) > 1e-5
        assert bool((got.sum(-1) == 1).all()) and torch.equal(got[clear], want[clear])
    tie = torch.zeros(4, 3, device="cuda")
    assert torch.equal(sample_hard(tie, torch.full((4, 3), 0.5, device="cuda")).argmax(-1).cpu(), torch.zeros(4, dtype=torch.long))
