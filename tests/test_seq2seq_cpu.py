"""Host-side checks of the seq2seq decoder mirror that need no GPU: parameter names / shapes / creation order against
the reference's recorded state_dict, constructor errors."""
import numpy as np
import pytest
import torch

from conftest import load_golden


@pytest.mark.parametrize("name", ["seq2seq2d_h32", "seq2seq3d_h64", "seq2seq2d_h48_n3"])
def test_state_dict_matches_reference(name):
    from aether_b200.seq2seq import Seq2SeqAether
    g = load_golden(name)
    D = int(g["D"])
    m = Seq2SeqAether({"num_vars": int(g["N"]), "input_size": 2 * D, "use_3d": D == 3, "decoder_hidden": int(g["n_hid"]),
                       "num_edge_types": int(g["n_types"]), "skip_first": bool(g["skip_first"]), "decoder_dropout": 0.0,
                       "encoder_hidden": 16, "gumbel_temp": 0.5, "rff_std": 1.0})
    ref_dec = [k[4:] for k in g if k.startswith("dec.")]
    mine = [k for k in m.decoder.state_dict()]
    assert mine == ref_dec                                       # same names in the same (creation) order
    for k in ref_dec:
        assert tuple(m.decoder.state_dict()[k].shape) == g["dec." + k].shape, k
    assert [k for k in m.field_net.state_dict()] == [k[len("field_net."):] for k in g if k.startswith("field_net.")]
    assert np.array_equal(m.coordinate_embedding.B.numpy(), g["rff_B"])   # default_rng(42) like the reference
    assert np.array_equal(m.decoder.send_edges.numpy(), g["send_edges"])
    assert np.array_equal(m.decoder.recv_edges.numpy(), g["recv_edges"])


def test_constructor_errors_and_cpu_refusal():
    from aether_b200.seq2seq import RecurrentDecoder
    base = {"num_vars": 5, "input_size": 4, "decoder_hidden": 32, "num_edge_types": 2, "skip_first": False,
            "decoder_dropout": 0.0}
    with pytest.raises(ValueError):
        RecurrentDecoder(dict(base, num_edge_types=5))
    with pytest.raises(ValueError):
        RecurrentDecoder(dict(base, decoder_hidden=30))
    with pytest.raises(ValueError):
        RecurrentDecoder(dict(base, input_size=6))
    dec = RecurrentDecoder(base).eval()
    with pytest.raises(RuntimeError):                            # no CPU fallback
        dec(torch.zeros(1, 5, 4), torch.zeros(1, 5, 32), torch.zeros(1, 20, 2), torch.zeros(1, 5, 2))


def test_seeded_init_reproduces_the_reference():
    """torch.manual_seed(s); Aether(params) in the reference creates encoder, decoder, field_net in that order with default /
    xavier inits: the mirror consumes the generator identically, so the initial weights are the reference's bit for bit."""
    from aether_b200.seq2seq import Seq2SeqAether
    g = load_golden("seq2seq2d_h32")
    torch.manual_seed(int(g["seed"]))
    np.random.seed(int(g["seed"]))
    m = Seq2SeqAether.with_encoder({
        "num_vars": 5, "input_size": 4, "use_3d": False, "decoder_hidden": 32, "num_edge_types": 2, "skip_first": False,
        "decoder_dropout": 0.0, "encoder_hidden": 16, "encoder_rnn_hidden": 16, "encoder_rnn_type": "lstm",
        "encoder_dropout": 0.0, "encoder_mlp_num_layers": 3, "encoder_mlp_hidden": 16, "prior_num_layers": 3,
        "prior_hidden_size": 16, "gumbel_temp": 0.5, "rff_std": 1.0, "pos_representation": "polar"})
    sd = m.state_dict()
    for k in g:
        if k.startswith("dec."):
            assert np.array_equal(sd["decoder." + k[4:]].numpy(), g[k]), k
        elif k.startswith("field_net."):
            assert np.array_equal(sd[k].numpy(), g[k]), k
        elif k.startswith("enc.") and ".bn." not in k:
            assert np.array_equal(sd["encoder." + k[4:]].numpy(), g[k]), k


def test_pickle_drops_workspace():
    import io
    from aether_b200.seq2seq import Seq2SeqAether
    m = Seq2SeqAether.with_encoder({
        "num_vars": 3, "input_size": 4, "use_3d": False, "decoder_hidden": 8, "num_edge_types": 2, "skip_first": False,
        "decoder_dropout": 0.0, "encoder_hidden": 8, "encoder_rnn_hidden": 4, "encoder_rnn_type": "lstm",
        "encoder_dropout": 0.0, "encoder_mlp_num_layers": 1, "prior_num_layers": 1, "gumbel_temp": 0.5})
    m.decoder._ws = torch.zeros(1 << 20, dtype=torch.uint8)
    m.encoder._ws = torch.zeros(1 << 20, dtype=torch.uint8)
    buf = io.BytesIO()
    torch.save(m, buf)                                      # the reference saves whole modules (experiments/*/main.py)
    assert buf.tell() < (1 << 19)
    buf.seek(0)
    m2 = torch.load(buf, weights_only=False)
    assert m2.decoder._ws is None and m2.encoder._ws is None
    assert all(torch.equal(a, b) for a, b in zip(m.state_dict().values(), m2.state_dict().values()))


def test_training_entry_point_says_what_is_missing():
    from aether_b200.seq2seq import Seq2SeqAether
    m = Seq2SeqAether({"num_vars": 3, "input_size": 4, "use_3d": False, "decoder_hidden": 8, "num_edge_types": 2,
                       "skip_first": False, "decoder_dropout": 0.0, "encoder_hidden": 8, "gumbel_temp": 0.5})
    with pytest.raises(RuntimeError, match="training is not built"):
        m.calculate_loss(torch.zeros(1, 3, 3, 4))
