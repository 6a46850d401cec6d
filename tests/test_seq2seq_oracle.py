"""The seq2seq decoder oracle against vectors of the unmodified reference (tests/golden/make_golden_seq2seq.py)."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import seq2seq_oracle as so

HERE = os.path.dirname(os.path.abspath(__file__))
CASES = sorted(glob.glob(os.path.join(HERE, "golden", "seq2seq*.npz")))


def load(path):
    g = np.load(path)
    w = {k[4:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("dec.")}
    w.update({k: torch.from_numpy(g[k]) for k in g.files if k.startswith("field_net.") or k == "rff_B"})
    return g, w


def test_cases_present():
    assert len(CASES) == 3


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[:-4] for p in CASES])
def test_oracle_matches_reference(path):
    g, w = load(path)
    D, K, skip = int(g["D"]), int(g["n_types"]), bool(g["skip_first"])
    t = lambda k: torch.from_numpy(g[k])
    s, r = so.fc_edges_sender_major(int(g["N"]))
    assert np.array_equal(s.numpy(), g["send_edges"]) and np.array_equal(r.numpy(), g["recv_edges"])
    rel, Rinv, ea = so.augmented_localizer(torch.cat([t("step_inputs"), t("step_field")], -1), D)
    assert torch.equal(rel, t("loc_rel_feat")) or (rel - t("loc_rel_feat")).abs().max() <= 1e-6
    assert (Rinv - t("loc_Rinv")).abs().max() <= 1e-6
    assert (ea - t("loc_edge_attr")).abs().max() <= 2e-6
    out, hid = so.decoder_step(w, t("step_inputs"), t("step_hidden"), t("step_edges"), t("step_field"), D, K, skip)
    assert (out - t("step_outputs")).abs().max() <= 2e-6
    assert (hid - t("step_hidden_new")).abs().max() <= 2e-6
    assert (so.predict_field(w, t("step_inputs"), D) - t("field_pred")).abs().max() <= 2e-6
    pf = so.predict_future(w, t("pf_inputs"), t("pf_edges"), int(g["pf_steps"]), D, K, skip)
    assert pf.shape == t("pf_predictions").shape
    assert (pf - t("pf_predictions")).abs().max() <= 1e-5


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[:-4] for p in CASES])
def test_encoder_oracle_matches_reference(path):
    """Prior branch of the reference Encoder (BatchNorm with non-trivial running statistics, LSTM over the burn-in,
    then one single_step_forward)."""
    g, w = load(path)
    w.update({k[4:]: torch.from_numpy(g[k]) for k in g.files if k.startswith("enc.")})
    D, rep = int(g["D"]), str(g["enc_pos_representation"])
    t = lambda k: torch.from_numpy(g[k])
    burn = t("pf_inputs")
    logits, h, c = so.encoder_prior(w, burn[:, :-1], D, rep)
    assert (logits - t("enc_prior_logits")).abs().max() <= 2e-6
    assert (h - t("enc_state_h")[0]).abs().max() <= 2e-6 and (c - t("enc_state_c")[0]).abs().max() <= 2e-6
    lg, h2, c2 = so.encoder_step(w, burn[:, -1], so.predict_field(w, burn[:, -1], D), h, c, D, rep)
    assert (lg - t("enc_step_logits")).abs().max() <= 2e-6
    assert (h2 - t("enc_step_state_h")[0]).abs().max() <= 2e-6 and (c2 - t("enc_step_state_c")[0]).abs().max() <= 2e-6
