"""Pin the CPU oracle against vectors produced by the unmodified reference
(tests/golden/make_golden.py).  CPU-only."""
import ast
import os

import numpy as np
import pytest
import torch

from glue_factory_b200 import synth
from oracle import lightglue as olg
from oracle import sp_open as osp


def load(golden_dir, name):
    with np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def test_weight_recipe_is_reproducible(golden_dir, sp_weights, lg_weights):
    want = dict(line.split() for line in open(os.path.join(golden_dir, "weights.sha256")))
    assert synth.state_dict_sha256(sp_weights) == want["superpoint_open"]
    assert synth.state_dict_sha256(lg_weights) == want["lightglue"]


@pytest.mark.parametrize("name", ["sp_a", "sp_b", "sp_c"])
def test_superpoint_open_matches_reference(golden_dir, sp_weights, name):
    g = load(golden_dir, name)
    conf = ast.literal_eval(str(g["conf"]))
    image = torch.from_numpy(g["image_u8"])[None].float() / 255.0
    with torch.no_grad():
        pred = osp.forward(sp_weights, image, conf)
    np.testing.assert_allclose(pred["heatmap"].numpy(), g["heatmap"], rtol=1e-4, atol=1e-7)
    # keypoints: integer positions, identical order (the fixtures have no ties at the k-th boundary)
    np.testing.assert_array_equal(pred["keypoints"].numpy(), g["keypoints"])
    np.testing.assert_allclose(pred["keypoint_scores"].numpy(), g["keypoint_scores"], rtol=1e-4, atol=1e-7)
    np.testing.assert_allclose(pred["descriptors"].numpy(), g["descriptors"], rtol=0, atol=2e-5)


def test_superpoint_open_rgb(golden_dir, sp_weights):
    g = load(golden_dir, "sp_rgb")
    image = torch.from_numpy(g["image_u8"])[None].float() / 255.0
    with torch.no_grad():
        pred = osp.forward(sp_weights, image, dict(max_num_keypoints=32))
    np.testing.assert_array_equal(pred["keypoints"].numpy(), g["keypoints"])
    np.testing.assert_allclose(pred["descriptors"].numpy(), g["descriptors"], rtol=0, atol=2e-5)


def _run_lg(lg_weights, g, conf, with_size=True):
    size = torch.from_numpy(g["image_size"]) if with_size else None
    with torch.no_grad():
        return olg.forward(
            lg_weights,
            torch.from_numpy(g["keypoints0"]),
            torch.from_numpy(g["keypoints1"]),
            torch.from_numpy(g["descriptors0"]),
            torch.from_numpy(g["descriptors1"]),
            size,
            size,
            conf,
        )


def _check_lg(pred, g, keys):
    for k in keys:
        got = pred[k].numpy()
        if got.dtype.kind == "i":
            np.testing.assert_array_equal(got, g[k], err_msg=k)
        else:
            np.testing.assert_allclose(got, g[k], rtol=2e-4, atol=2e-4, err_msg=k)


ALL_KEYS = [
    "matches0", "matches1", "matching_scores0", "matching_scores1",
    "ref_descriptors0", "ref_descriptors1", "log_assignment", "prune0", "prune1",
]


def test_lightglue_matches_reference(golden_dir, lg_weights):
    g = load(golden_dir, "lg_a")
    pred = _run_lg(lg_weights, g, dict(filter_threshold=0.1))
    _check_lg(pred, g, ALL_KEYS)
    assert (g["matches0"] >= 0).sum() > 20  # the fixture is not vacuous


def test_lightglue_without_image_size(golden_dir, lg_weights):
    g = {**load(golden_dir, "lg_a"), **load(golden_dir, "lg_nosize")}
    pred = _run_lg(lg_weights, g, dict(filter_threshold=0.1), with_size=False)
    _check_lg(pred, g, ["matches0", "matches1", "log_assignment"])


@pytest.mark.parametrize("name", ["lg_adaptive", "lg_prune_only", "lg_depth_only"])
def test_lightglue_adaptive_matches_patched_reference(golden_dir, lg_weights, name):
    g = load(golden_dir, name)
    conf = dict(
        filter_threshold=0.1,
        depth_confidence=float(g["depth_confidence"]),
        width_confidence=float(g["width_confidence"]),
    )
    pred = _run_lg(lg_weights, g, conf)
    assert pred["log_assignment"].shape == g["log_assignment"].shape
    _check_lg(pred, g, ALL_KEYS)


def test_boat_pipeline_matches_reference(golden_dir, sp_weights, lg_weights):
    import cv2

    g = load(golden_dir, "boat_pipeline")
    preds, sizes = [], []
    for name in ("boat1.png", "boat2.png"):
        im = cv2.imread(os.path.join(golden_dir, "assets", name), cv2.IMREAD_GRAYSCALE)
        t = torch.from_numpy(im)[None, None].float() / 255.0
        sizes.append(torch.tensor([[im.shape[1], im.shape[0]]]).float())
        with torch.no_grad():
            preds.append(
                osp.forward(sp_weights, t, dict(max_num_keypoints=1024, nms_radius=4, detection_threshold=0.0))
            )
    for i in (0, 1):
        np.testing.assert_array_equal(preds[i]["keypoints"].numpy(), g[f"keypoints{i}"])
        np.testing.assert_allclose(
            preds[i]["descriptors"][:, :48].numpy(), g[f"descriptors{i}_head"], rtol=0, atol=2e-5
        )
    with torch.no_grad():
        out = olg.forward(
            lg_weights, preds[0]["keypoints"], preds[1]["keypoints"], preds[0]["descriptors"],
            preds[1]["descriptors"], sizes[0], sizes[1], dict(filter_threshold=0.1),
        )
    agree = (out["matches0"].numpy() == g["matches0"]).mean()
    assert agree >= 0.995, agree
    np.testing.assert_allclose(out["log_assignment"][:, ::8, ::8].numpy(), g["log_assignment_sub"], rtol=1e-3, atol=1e-3)
