"""The reference arm of bench.py (CPU oracle, no GPU needed) keeps the JSON contract the driver reads."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(300)
@pytest.mark.parametrize("workload", ["c1", "c2"])
def test_reference_arm_json_contract(workload):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", workload,
                        "--steps", "1", "--warmup", "1"], capture_output=True, text=True, timeout=280, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "edge_messages_per_sec" and d["unit"] == "edge-messages/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["steps"] == 1 and d["warmup"] == 1
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["sample"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert abs(d["e2e"]["value"] - d["value"]) < 1e-6 * d["value"]


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--workload",
                        "c1", "--steps", "1", "--warmup", "1"], capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert r.returncode == 0 and not [l for l in r.stdout.splitlines() if l.startswith("{")]


def test_roofline_formulas_match_survey():
    """per-edge MACs Fe*H + 13 H^2 = 54 912 (3D, H=64) / 858 624 (3D, H=256); the node-factorised
    form executes Fe*H + 7 H^2; HBM bytes per launch of the segmented mean and the feature kernel as stated in DESIGN.md 3."""
    sys.path.insert(0, ROOT)
    import bench
    assert bench.algorithmic_flops_edge_mlp(3, 64, 1) == 2 * 54912
    assert bench.algorithmic_flops_edge_mlp(3, 256, 1) == 2 * 858624
    assert bench.executed_flops_edge_mlp(3, 256, 1) == 2 * (26 * 256 + 7 * 256 * 256)
    prof = {"ms": [0, 0, 0, 0, 1.0, 0, 2.0], "spans": [0, 0, 0, 0, 1, 0, 2]}
    E, Nn, H = 1000, 100, 256
    r = bench.hbm_rooflines(prof, 2, H, E, Nn, True, {"hbm_gbs": 1.0})
    seg = r["segmented_mean (receiver-sorted, no atomics)"]
    assert seg["algorithmic_bytes_per_launch"] == E * H * 2 + Nn * H * 8 + (Nn + 1) * 4 and seg["launches"] == 2
    feat = r["edge_features (gather + frame transform)"]
    assert feat["algorithmic_bytes_per_launch"] == E * (12 + 8 + 128)
    assert abs(feat["achieved"] - E * 148 / 1e-3 / 1e9) < 1e-9


def test_product_arm_generator_draws_the_oracles_numbers():
    """bench.py's product arm takes its inputs from aether_b200.synth (nothing from oracle/); both generators must
    produce the same particles so that the CPU baseline and the GPU arm time the same workload."""
    import torch
    sys.path.insert(0, ROOT)
    from aether_b200 import synth
    from oracle import aether_oracle as O
    for (B, N, D, seed) in [(3, 5, 2, 1), (2, 100, 2, 4), (4, [3, 7, 5, 2], 3, 9)]:
        a = synth.synth_state(B, N, D, seed=seed)
        b = O.synth_state(B, N, D, seed=seed)
        assert all(torch.equal(u, w) for u, w in zip(a, b))
