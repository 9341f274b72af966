"""Pin the CPU oracle of the SuperGlue configuration (oracle/sp_nonfree.py, oracle/superglue.py) against
vectors produced by the unmodified reference (tests/golden/make_golden_nonfree.py).  CPU-only."""
import ast
import os

import numpy as np
import pytest
import torch

from glue_factory_b200 import synth
from oracle import sp_nonfree as ospn
from oracle import superglue as osg


def load(golden_dir, name):
    with np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


@pytest.fixture(scope="module")
def spn_weights():
    return synth.superpoint_nonfree_state_dict()


@pytest.fixture(scope="module")
def sg_weights():
    return synth.superglue_state_dict()


def test_weight_recipe_is_reproducible(golden_dir, spn_weights, sg_weights):
    want = dict(line.split() for line in open(os.path.join(golden_dir, "weights_nonfree.sha256")))
    assert synth.state_dict_sha256(spn_weights) == want["superpoint_nonfree"]
    assert synth.state_dict_sha256(sg_weights) == want["superglue"]


@pytest.mark.parametrize("name", ["spn_a", "spn_b", "spn_c", "spn_d"])
def test_superpoint_nonfree_matches_reference(golden_dir, spn_weights, name):
    g = load(golden_dir, name)
    conf = ast.literal_eval(str(g["conf"]))
    image = torch.from_numpy(g["image_u8"])[None].float() / 255.0
    size = torch.from_numpy(g["image_size"])[None] if g["image_size"][0] > 0 else None
    with torch.no_grad():
        pred = ospn.forward(spn_weights, image, conf, image_size=size)
    # keypoints: integer positions, identical order (the fixtures have no ties at the k-th boundary)
    np.testing.assert_array_equal(pred["keypoints"].numpy(), g["keypoints"])
    np.testing.assert_allclose(pred["keypoint_scores"].numpy(), g["keypoint_scores"], rtol=1e-4, atol=1e-7)
    np.testing.assert_allclose(pred["descriptors"].numpy(), g["descriptors"], rtol=0, atol=2e-5)


@pytest.mark.parametrize("name", ["sg_a", "sg_b"])
def test_superglue_matches_reference(golden_dir, sg_weights, name):
    g = load(golden_dir, name)
    conf = ast.literal_eval(str(g["conf"]))
    t = {k: torch.from_numpy(v) for k, v in g.items() if k != "conf"}
    with torch.no_grad():
        pred = osg.forward(sg_weights, t["keypoints0"], t["keypoints1"], t["descriptors0"], t["descriptors1"],
                           t["scores0"], t["scores1"], t["image_size"], t["image_size"], conf)
    np.testing.assert_allclose(pred["sinkhorn_cost"].numpy(), g["sinkhorn_cost"], rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(pred["log_assignment"].numpy(), g["log_assignment"], rtol=1e-4, atol=2e-4)
    np.testing.assert_array_equal(pred["matches0"].numpy(), g["matches0"])
    np.testing.assert_array_equal(pred["matches1"].numpy(), g["matches1"])
    np.testing.assert_allclose(pred["matching_scores0"].numpy(), g["matching_scores0"], rtol=1e-3, atol=1e-6)
    np.testing.assert_allclose(pred["matching_scores1"].numpy(), g["matching_scores1"], rtol=1e-3, atol=1e-6)
    assert (g["matches0"] >= 0).sum() >= 10  # the fixture is not degenerate


def test_superglue_pipeline_matches_reference(golden_dir, spn_weights, sg_weights):
    g = load(golden_dir, "sg_pipeline")
    size = torch.tensor([[256.0, 192.0]])
    conf_sp = dict(max_num_keypoints=256, nms_radius=3, detection_threshold=0.0)
    with torch.no_grad():
        p0 = ospn.forward(spn_weights, torch.from_numpy(g["image0_u8"])[None].float() / 255.0, conf_sp, image_size=size)
        p1 = ospn.forward(spn_weights, torch.from_numpy(g["image1_u8"])[None].float() / 255.0, conf_sp, image_size=size)
        pred = osg.forward(sg_weights, p0["keypoints"], p1["keypoints"], p0["descriptors"], p1["descriptors"],
                           p0["keypoint_scores"], p1["keypoint_scores"], size, size,
                           dict(num_sinkhorn_iterations=100, filter_threshold=0.2))
    np.testing.assert_array_equal(p0["keypoints"].numpy(), g["keypoints0"])
    np.testing.assert_array_equal(p1["keypoints"].numpy(), g["keypoints1"])
    np.testing.assert_allclose(pred["log_assignment"][:, ::4, ::4].numpy(), g["log_assignment_sub"], rtol=1e-3, atol=2e-3)
    agree = (pred["matches0"].numpy() == g["matches0"]).mean()
    assert agree >= 0.99, agree
