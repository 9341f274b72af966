"""Pin the CPU oracle of the nearest-neighbour matcher (oracle/nn_matcher.py) against vectors produced by the
unmodified reference (tests/golden/make_golden_nn.py).  CPU-only."""
import os

import numpy as np
import pytest
import torch

from oracle import nn_matcher as onn


def load(golden_dir, name):
    with np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


@pytest.mark.parametrize("name", ["nn_a", "nn_b", "nn_c"])
def test_nn_oracle_matches_reference(golden_dir, name):
    g = load(golden_dir, name)
    d0, d1 = torch.from_numpy(g["descriptors0"]), torch.from_numpy(g["descriptors1"])
    pred = onn.forward(d0, d1, float(g["ratio_thresh"]) or None, float(g["distance_thresh"]) or None,
                       bool(g["mutual_check"]))
    np.testing.assert_allclose(pred["similarity"].numpy(), g["out_similarity"], rtol=0, atol=1e-6)
    np.testing.assert_allclose(pred["log_assignment"].numpy(), g["out_log_assignment"], rtol=0, atol=2e-5)
    for k in ("matches0", "matches1"):
        assert pred[k].dtype == torch.int64
        np.testing.assert_array_equal(pred[k].numpy(), g["out_" + k])
    for k in ("matching_scores0", "matching_scores1"):
        np.testing.assert_array_equal(pred[k].numpy(), g["out_" + k])
    # the fixtures exercise the tests they are meant to: some rows rejected, some kept
    kept = (g["out_matches0"] > -1).mean()
    assert 0.2 < kept < 0.95, kept


def test_nn_oracle_teacher_forced_is_exact(golden_dir):
    """Fed the reference's own similarity, the decisions are bit-exact (same fp32 comparisons)."""
    g = load(golden_dir, "nn_b")
    pred = onn.forward(None, None, float(g["ratio_thresh"]), float(g["distance_thresh"]), True,
                       sim=torch.from_numpy(g["out_similarity"]))
    np.testing.assert_array_equal(pred["matches0"].numpy(), g["out_matches0"])
    np.testing.assert_array_equal(pred["matches1"].numpy(), g["out_matches1"])


def test_nn_oracle_edge_cases():
    sim = torch.tensor([[[0.9, 0.9, 0.1]]])  # tie: lowest index; single row
    out = onn.forward(None, None, sim=sim)
    assert out["matches0"].tolist() == [[0]]
    assert out["matches1"].tolist() == [[0, -1, -1]]
    assert out["log_assignment"].shape == (1, 2, 4)
    assert out["log_assignment"][0, -1].abs().sum() == 0 and out["log_assignment"][0, :, -1].abs().sum() == 0
    out = onn.forward(None, None, distance_thresh=0.3, sim=sim)  # 2 (1 - 0.9) = 0.2 > 0.09
    assert out["matches0"].tolist() == [[-1]]
