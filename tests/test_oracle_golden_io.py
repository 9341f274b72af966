"""CPU: the oracle restatements of the rows either side of the path against vectors produced by the UNMODIFIED
reference (tests/golden/make_golden_io.py): ImagePreprocessor, export_predictions, match metrics."""
import ast
import os

import numpy as np
import pytest
import torch

from oracle import eval_utils as oev
from oracle import export as oex
from oracle import preprocess as opre

PREP = ["prep_a", "prep_b", "prep_c", "prep_d", "prep_e", "prep_f", "prep_g"]


def _conf(g):
    return {k: ast.literal_eval(v) for k, v in zip(g["conf_keys"].tolist(), g["conf_vals"].tolist())}


@pytest.mark.parametrize("name", PREP)
def test_preprocess_oracle_matches_reference(golden_dir, name):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    out = opre.preprocess(opre.numpy_image_to_torch(g["input_u8"]), _conf(g))
    assert out["image"].shape == g["image"].shape
    assert torch.equal(out["image"], torch.from_numpy(g["image"]))
    assert np.array_equal(out["scales"].numpy(), g["scales"])
    assert np.array_equal(out["image_size"], g["image_size"])
    assert np.array_equal(out["original_image_size"], g["original_image_size"])
    assert np.allclose(np.asarray(out["transform"], dtype=np.float64), g["transform"])
    if "padding_mask" in g.files:
        assert np.array_equal(out["padding_mask"].numpy(), g["padding_mask"])


def test_blur_plan_follows_kornia_rules():
    assert opre.blur_plan((240, 320), (96, 128), True) == (3, 0.75, 3, 0.75)
    ky, sy, kx, sx = opre.blur_plan((330, 440), (60, 80), True)
    assert (ky, kx) == (9, 9) and abs(sy - 2.25) < 1e-12
    assert opre.blur_plan((90, 150), (120, 200), True) is None      # up-scaling
    assert opre.blur_plan((240, 320), (96, 128), False) is None     # antialias off
    ky, sy, kx, sx = opre.blur_plan((200, 300), (75, 400), True)    # only one axis shrinks: the other gets sigma 0.001
    assert sx == 0.001 and kx == 3 and ky == 3
    t = opre.gaussian_taps(3, 0.001)
    assert t.tolist() == [0.0, 1.0, 0.0]


@pytest.mark.parametrize("name,half", [("export_f32", False), ("export_f16", True)])
def test_export_oracle_matches_reference(golden_dir, name, half):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    keys = g["keys"].tolist()
    for i in range(len(g["names"])):
        pred = {k: torch.from_numpy(g[f"in{i}/{k}"]) for k in keys + ["log_assignment"]}
        data = {"view0": {"scales": torch.from_numpy(g[f"in{i}/scales0"])},
                "view1": {"scales": torch.from_numpy(g[f"in{i}/scales1"])}}
        rec = oex.export_records(pred, data, keys=keys, as_half=half)
        assert sorted(rec) == sorted(keys)
        for k in keys:
            want = g[f"out{i}/{k}"]
            assert rec[k].dtype == want.dtype and np.array_equal(rec[k], want), k


def test_eval_oracle_matches_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, "eval_metrics.npz"))
    for i in range(int(g["h_count"])):
        t = lambda k: torch.from_numpy(g[f"h{i}/{k}"])  # noqa: E731
        res = oev.eval_matches_homography(t("H_0to1"), t("keypoints0"), t("keypoints1"), t("matches0"),
                                          t("matching_scores0"))
        for k, v in res.items():
            assert v == float(g[f"h{i}/res/{k}"]), (i, k)
        err, _ = oev.eval_homography_dlt(t("H_0to1"), torch.tensor([640.0, 480.0]), t("keypoints0"), t("keypoints1"),
                                         t("matches0"), t("matching_scores0"))
        want = float(g[f"h{i}/res/H_error_dlt"])
        assert (np.isnan(err) and np.isnan(want)) or err == want, (i, err, want)
    for i in range(int(g["e_count"])):
        t = lambda k: torch.from_numpy(g[f"e{i}/{k}"])  # noqa: E731
        res = oev.eval_matches_epipolar(t("cam0")[2:6], t("cam1")[2:6], t("R"), t("t"), t("keypoints0"), t("keypoints1"),
                                        t("matches0"), t("matching_scores0"))
        for k, v in res.items():
            assert abs(v - float(g[f"e{i}/res/{k}"])) < 1e-7, (i, k, v)


def test_record_file_round_trip(tmp_path):
    """The zip-of-npy container behind export_predictions when h5py is absent, read with h5py's access pattern."""
    from glue_factory_b200.export import RecordFile, _Writer

    w = _Writer(tmp_path / "pred.h5")
    a = {"keypoints0": np.random.rand(5, 2).astype(np.float16), "matches0": np.arange(5, dtype=np.int64) - 1}
    w.write("seq/1.ppm/2.ppm", a)
    w.write("seq/1.ppm/2.ppm", {"keypoints0": np.zeros((1, 2), np.float32)})  # duplicate names are skipped
    w.write("other", {"x": np.zeros((0, 2), np.float32)})
    w.close()
    with RecordFile(tmp_path / "pred.h5") as f:
        assert sorted(f.keys()) == ["other", "seq/1.ppm/2.ppm"]
        grp = f["seq/1.ppm/2.ppm"]
        assert sorted(grp.keys()) == ["keypoints0", "matches0"]
        assert np.array_equal(np.asarray(grp["keypoints0"]), a["keypoints0"]) and grp["keypoints0"].dtype == np.float16
        assert np.array_equal(grp["matches0"][1:3], a["matches0"][1:3])
        assert f["other"]["x"].shape == (0, 2)
