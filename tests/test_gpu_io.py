"""GPU parity of the rows either side of the forward path through the C ABI:
ImagePreprocessor (resize / blur / pad), export packing + record file, match metrics + weighted DLT.

Tolerances: pre-processing is fp32 arithmetic in a different summation order than torch's convolution /
interpolation: |d| <= 2e-6 absolute on images in [0, 1] (stated, achieved printed).  Export records are bit-exact.
The transfer errors behind prec@{1,3}px agree with the reference to ~1e-2 px only: the reference inverts H with an
fp32 torch.pinverse (SVD), which by itself is off by up to 1.3e-2 px against the exact inverse on the fixtures, the
kernel inverts in fp64.  The precision counters are therefore compared up to the matches whose reference error lies
within 2e-2 px of a threshold (1e-6 in normalised coordinates for the epipolar thresholds); the DLT
corner error is compared at 1e-3 relative on the well-posed fixtures (the reference solves the 9 x 9 system by an
fp32 SVD, the kernel by Jacobi rotations in fp64) and at 2 % where 20-40 % outliers make the smallest eigenvector
ill-conditioned.
"""
import ast
import os

import numpy as np
import pytest
import torch

from oracle import eval_utils as oev
from oracle import export as oex
from oracle import preprocess as opre

pytestmark = pytest.mark.gpu

PREP = ["prep_a", "prep_b", "prep_c", "prep_d", "prep_e", "prep_f", "prep_g"]


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _conf(g):
    return {k: ast.literal_eval(v) for k, v in zip(g["conf_keys"].tolist(), g["conf_vals"].tolist())}


@pytest.mark.parametrize("name", PREP)
@pytest.mark.parametrize("as_u8", [False, True])
def test_preprocessor_vs_reference_golden(dev, golden_dir, name, as_u8):
    from gluefactory.utils.image import ImagePreprocessor

    g = np.load(os.path.join(golden_dir, name + ".npz"))
    u8 = g["input_u8"]
    img = torch.from_numpy(u8).to(dev) if as_u8 else opre.numpy_image_to_torch(u8).to(dev)
    out = ImagePreprocessor(_conf(g))(img)
    assert out["image"].shape == g["image"].shape and out["image"].dtype == torch.float32
    err = (out["image"].cpu() - torch.from_numpy(g["image"])).abs().max().item()
    print(f"{name} u8={as_u8}: max |d| = {err:.2e}")
    assert err <= 2e-6
    assert np.array_equal(out["scales"].cpu().numpy(), g["scales"])
    assert np.array_equal(out["image_size"], g["image_size"])
    assert np.array_equal(out["original_image_size"], g["original_image_size"])
    assert np.allclose(np.asarray(out["transform"], dtype=np.float64), g["transform"])
    if "padding_mask" in g.files:
        assert np.array_equal(out["padding_mask"].cpu().numpy(), g["padding_mask"])


@pytest.mark.parametrize("hw,conf", [((1200, 1600), dict(resize=1024, side="long", square_pad=True)),
                                     ((2000, 1500), dict(resize=480, side="short")),
                                     ((480, 640), dict(resize=1600, side="long"))])
def test_preprocessor_full_size_vs_oracle(dev, hw, conf):
    """MegaDepth / HPatches eval sizes (eval/megadepth1500.py resize 1600 long, eval/hpatches.py resize 480 short),
    a batch of two RGB images; oracle = the torch restatement of kornia's resize."""
    from gluefactory.utils.image import ImagePreprocessor

    g = torch.Generator().manual_seed(hw[0])
    u8 = torch.randint(0, 256, (2, hw[0], hw[1], 3), generator=g, dtype=torch.uint8)
    out = ImagePreprocessor(conf)(u8.to(dev))
    for b in range(2):
        ref = opre.preprocess(opre.numpy_image_to_torch(u8[b].numpy()), conf)
        err = (out["image"][b].cpu() - ref["image"]).abs().max().item()
        assert out["image"][b].shape == ref["image"].shape and err <= 2e-6, err
    assert np.array_equal(out["image_size"], ref["image_size"])


def test_uint8_conversion_is_exact_for_all_values(dev):
    """value / 255 of every uint8 value equals numpy_image_to_torch's float32(value / 255.0) bit for bit (the same
    device function converts the pixels in the SuperPoint stem kernels)."""
    from gluefactory.utils.image import ImagePreprocessor

    u8 = torch.arange(256, dtype=torch.uint8).view(16, 16)
    out = ImagePreprocessor(dict(resize=None))(u8.to(dev))["image"].cpu()
    assert torch.equal(out, opre.numpy_image_to_torch(u8.numpy()))


def test_preprocessor_rejects_unsupported(dev):
    from gluefactory.utils.image import ImagePreprocessor

    img = torch.rand(1, 32, 32, device=dev)
    with pytest.raises(NotImplementedError):
        ImagePreprocessor(dict(resize=16, interpolation="bicubic"))(img)
    with pytest.raises(RuntimeError):
        ImagePreprocessor(dict(resize=16))(img.cpu())
    with pytest.raises(ValueError):
        ImagePreprocessor(dict(resize=16, side="diagonal"))(img)


@pytest.mark.parametrize("name,half", [("export_f32", False), ("export_f16", True)])
def test_export_pack_vs_reference_golden(dev, golden_dir, name, half):
    """The three recorded pairs packed as ONE batch of 3: every record equals what the unmodified reference wrote."""
    from glue_factory_b200.export import RecordPacker

    g = np.load(os.path.join(golden_dir, name + ".npz"))
    keys = g["keys"].tolist()
    n = len(g["names"])
    pred = {k: torch.cat([torch.from_numpy(g[f"in{i}/{k}"]) for i in range(n)]).to(dev) for k in keys}
    data = {"view0": {"scales": torch.cat([torch.from_numpy(g[f"in{i}/scales0"]) for i in range(n)]).to(dev)},
            "view1": {"scales": torch.cat([torch.from_numpy(g[f"in{i}/scales1"]) for i in range(n)]).to(dev)}}
    packer = RecordPacker(as_half=half)
    host = packer.pack(pred, data).cpu().numpy()
    for i in range(n):
        rec = packer.unpack(host, i)
        for k in keys:
            want = g[f"out{i}/{k}"]
            assert rec[k].dtype == want.dtype and np.array_equal(rec[k], want), (i, k)


def test_export_predictions_round_trip(dev, tmp_path, sp_weights, lg_weights):
    """export_predictions over a loader of two batches of two pairs through TwoViewPipeline: the record file read
    back equals the oracle's export of the very same predictions (bit-exact), pair by pair."""
    from glue_factory_b200 import synth
    from gluefactory.models import get_model
    from gluefactory.utils.export_predictions import RecordFile, export_predictions

    torch.save(sp_weights, tmp_path / "sp.pth")
    torch.save(lg_weights, tmp_path / "lg.pth")
    pipe = get_model("two_view_pipeline")({
        "extractor": {"name": "extractors.superpoint_open", "weights": str(tmp_path / "sp.pth"), "max_num_keypoints": 256,
                      "detection_threshold": -1, "nms_radius": 3},
        "matcher": {"name": "matchers.lightglue", "weights": str(tmp_path / "lg.pth"), "filter_threshold": 0.1}}).to(dev).eval()
    keys = ["keypoints0", "keypoints1", "keypoint_scores0", "keypoint_scores1", "matches0", "matches1",
            "matching_scores0", "matching_scores1"]
    loader = []
    for i in range(2):
        im0, im1, _ = synth.synthetic_batch(4 * i, 2, 160, 192)
        sc = torch.tensor([[0.4 + 0.1 * i, 0.5], [0.75, 0.3 + 0.2 * i]])
        size = torch.tensor([[192.0, 160.0]] * 2)
        loader.append({"view0": {"image": im0, "image_size": size, "scales": sc},
                       "view1": {"image": im1, "image_size": size, "scales": sc.flip(0)},
                       "name": [f"scene{i}/pair{j}" for j in range(2)]})
    captured = []

    def grab(pred, data):
        captured.append(({k: v.clone() for k, v in pred.items() if k in keys},
                         {v: {"scales": data[v]["scales"].clone()} for v in ("view0", "view1")}))
        return {}

    for half in (False, True):
        captured.clear()
        path = export_predictions(loader, pipe, tmp_path / f"pred_{half}.h5", as_half=half, keys=keys, callback_fn=grab)
        with RecordFile(path) as f:
            assert sorted(f.keys()) == sorted(n for d in loader for n in d["name"])
            for (pred, data), batch in zip(captured, loader):
                want = oex.export_batch({k: v.cpu() for k, v in pred.items()},
                                        {k: {"scales": v["scales"].cpu()} for k, v in data.items()}, as_half=half)
                for b, name in enumerate(batch["name"]):
                    grp = f[name]
                    assert sorted(grp.keys()) == sorted(keys)
                    for k in keys:
                        got = np.asarray(grp[k])
                        assert got.dtype == want[b][k].dtype and np.array_equal(got, want[b][k]), (name, k)


def test_match_metrics_vs_reference_golden(dev, golden_dir):
    from gluefactory.eval.utils import eval_homography_dlt, eval_matches_epipolar, eval_matches_homography

    g = np.load(os.path.join(golden_dir, "eval_metrics.npz"))
    for i in range(int(g["h_count"])):
        t = lambda k: torch.from_numpy(g[f"h{i}/{k}"]).to(dev)  # noqa: E731
        data = {"H_0to1": t("H_0to1"), "view0": {"image_size": torch.tensor([640.0, 480.0], device=dev)}}
        pred = {k: t(k) for k in ("keypoints0", "keypoints1", "matches0", "matching_scores0")}
        res = eval_matches_homography(data, pred)
        p0, p1, _ = oev.get_matches_scores(*(pred[k].cpu() for k in ("keypoints0", "keypoints1", "matches0", "matching_scores0")))
        ref_err = oev.sym_homography_error(p0, p1, data["H_0to1"].cpu())
        n = max(len(p0), 1)
        for k, thr in (("prec@1px", 1.0), ("prec@3px", 3.0)):
            borderline = int(((ref_err - thr).abs() < 2e-2).sum())
            assert abs(res[k] - float(g[f"h{i}/res/{k}"])) <= borderline / n + 1e-6, (i, k, res[k], borderline)
        for k in ("num_matches", "num_keypoints"):
            assert res[k] == float(g[f"h{i}/res/{k}"]), (i, k, res[k])
        assert isinstance(res["num_matches"], int)
        err = eval_homography_dlt(data, pred)["H_error_dlt"]
        want = float(g[f"h{i}/res/H_error_dlt"])
        print(f"homography {i}: H_error_dlt {err:.6f} (reference {want:.6f})")
        if np.isnan(want):
            assert np.isnan(err)
        else:
            tol = 1e-3 if want < 1.0 else 2e-2
            assert abs(err - want) <= tol * max(want, 1e-2), (i, err, want)
    for i in range(int(g["e_count"])):
        t = lambda k: torch.from_numpy(g[f"e{i}/{k}"]).to(dev)  # noqa: E731
        data = {"view0": {"camera": t("cam0")[None]}, "view1": {"camera": t("cam1")[None]},
                "T_0to1": torch.cat([t("R").reshape(-1), t("t")])[None]}
        pred = {k: t(k) for k in ("keypoints0", "keypoints1", "matches0", "matching_scores0")}
        res = eval_matches_epipolar(data, pred)
        p0, p1, _ = oev.get_matches_scores(*(pred[k].cpu() for k in ("keypoints0", "keypoints1", "matches0", "matching_scores0")))
        c0, c1 = t("cam0").cpu(), t("cam1").cpu()
        ref_err = oev.sym_epipolar_distance(((p0 - c0[4:6]) / c0[2:4])[None], ((p1 - c1[4:6]) / c1[2:4])[None],
                                            oev.T_to_E(t("R").cpu(), t("t").cpu()), squared=False)[0]
        n = max(len(p0), 1)
        for k, thr in (("epi_prec@1e-4", 1e-4), ("epi_prec@5e-4", 5e-4), ("epi_prec@1e-3", 1e-3)):
            borderline = int(((ref_err - thr).abs() < 1e-6).sum())
            assert abs(res[k] - float(g[f"e{i}/res/{k}"])) <= borderline / n + 1e-6, (i, k, res[k], borderline)
        for k in ("num_matches", "num_keypoints"):
            assert res[k] == float(g[f"e{i}/res/{k}"]), (i, k, res[k])


def test_match_metrics_batched_vs_oracle(dev):
    """A batch of 8 pairs at 2048 keypoints (one launch per metric) against the oracle run pair by pair, including a
    pair without matches (nan_to_num -> 0 for the homography metric, NaN for the epipolar one, as the reference)."""
    from gluefactory.eval.utils import eval_homography_dlt, eval_matches_epipolar, eval_matches_homography

    g = torch.Generator().manual_seed(3)
    B, M, N = 8, 2048, 2048
    H = torch.eye(3).repeat(B, 1, 1) + torch.randn(B, 3, 3, generator=g) * torch.tensor([[0.05, 0.05, 8.0], [0.05, 0.05, 8.0], [1e-5, 1e-5, 0.0]])
    kp0 = torch.rand(B, M, 2, generator=g) * 1024
    kp1 = torch.rand(B, N, 2, generator=g) * 1024
    m0 = torch.full((B, M), -1, dtype=torch.int64)
    for b in range(B - 1):
        n = 200 + 150 * b
        src, dst = torch.randperm(M, generator=g)[:n], torch.randperm(N, generator=g)[:n]
        q = torch.cat([kp0[b, src], torch.ones(n, 1)], -1) @ H[b].T
        kp1[b, dst] = q[:, :2] / q[:, 2:] + torch.randn(n, 2, generator=g) * (0.3 + 0.4 * b)
        m0[b, src] = dst
    ms0 = torch.where(m0 > -1, torch.rand(B, M, generator=g), torch.zeros(B, M))
    size = torch.tensor([[1024.0, 1024.0]] * B)
    cam = torch.tensor([[1024.0, 1024.0, 800.0, 820.0, 512.0, 500.0]] * B)
    R = torch.linalg.qr(torch.randn(B, 3, 3, generator=g)).Q
    R = R * torch.linalg.det(R)[:, None, None]
    tvec = torch.randn(B, 3, generator=g)
    data = {"H_0to1": H.to(dev), "view0": {"image_size": size.to(dev), "camera": cam.to(dev)},
            "view1": {"camera": cam.to(dev)}, "T_0to1": torch.cat([R.reshape(B, 9), tvec], -1).to(dev)}
    pred = {"keypoints0": kp0.to(dev), "keypoints1": kp1.to(dev), "matches0": m0.to(dev), "matching_scores0": ms0.to(dev)}
    rh, re, rd = eval_matches_homography(data, pred), eval_matches_epipolar(data, pred), eval_homography_dlt(data, pred)
    for b in range(B):
        wh = oev.eval_matches_homography(H[b], kp0[b], kp1[b], m0[b], ms0[b])
        n = wh["num_matches"]
        assert rh["num_matches"][b] == n
        p0, p1, _ = oev.get_matches_scores(kp0[b], kp1[b], m0[b], ms0[b])
        ref_err = oev.sym_homography_error(p0, p1, H[b])
        for k, thr in (("prec@1px", 1.0), ("prec@3px", 3.0)):
            borderline = int(((ref_err - thr).abs() < 2e-2).sum())
            assert abs(rh[k][b] - wh[k]) <= borderline / max(n, 1) + 1e-6, (b, k, borderline)
        we = oev.eval_matches_epipolar(cam[b, 2:6], cam[b, 2:6], R[b], tvec[b], kp0[b], kp1[b], m0[b], ms0[b])
        for k in ("epi_prec@1e-4", "epi_prec@5e-4", "epi_prec@1e-3"):
            if n == 0:
                assert np.isnan(re[k][b]) and np.isnan(we[k])
            else:
                assert abs(re[k][b] - we[k]) <= 2.0 / n + 1e-6, (b, k)  # random geometry: few matches are that close at all
        want, _ = oev.eval_homography_dlt(H[b], size[b], kp0[b], kp1[b], m0[b], ms0[b])
        if n < 4:
            assert np.isnan(rd["H_error_dlt"][b]) and np.isnan(want)
        else:
            assert abs(rd["H_error_dlt"][b] - want) <= 2e-2 * max(want, 0.05), (b, rd["H_error_dlt"][b], want)
