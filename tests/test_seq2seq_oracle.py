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
