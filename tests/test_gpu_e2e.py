"""GPU end-to-end parity of the B200 models against the CPU oracle and the reference-generated
golden vectors, through the public Glue Factory API (get_model / forward(data)).

Tolerances (fp32 oracle vs bf16-operand / fp32-accumulate kernels), stated per tensor, with the achieved
figures of round 2 in brackets (the benchmarked sizes are covered by test_gpu_parity_config.py):
  heat map            |d| <= 3e-2 * max       [2.3e-2]   (bf16 conv stack)
  dense descriptors   cosine >= 0.998          [0.9991]   (bf16 conv stack)
  LightGlue layer out relative L2 <= 1e-2      [6.1e-3 .. 7.1e-3]  (bf16 GEMM operands, fp32 residual stream)
  log_assignment      median |d|/|ref| <= 3e-3, p99 <= 1.5e-2  [1.5e-3, 8.5e-3]
  match agreement     >= 0.995 on the DECISIVE keypoints, >= 0.99 overall, teacher-forced on the oracle's keypoints
                      [0.994 .. 1.0 overall].  A keypoint is borderline when the reference itself decides it by a
                      hair: matching score within 10 % of filter_threshold, or best / second-best log-assignment of its
                      row (or of its best column) closer than 0.05; bf16 rounding may flip those either way.  The
                      benchmark-size agreement (99.95 %, no exclusions) is asserted in test_gpu_parity_config.py.
The teacher-forced stage tests in test_gpu_bandwidth.py carry the bit-exact / 1e-3 claims.
"""
import os

import numpy as np
import pytest
import torch

from glue_factory_b200 import synth
from oracle import lightglue as olg
from oracle import sp_open as osp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def weight_files(tmp_path_factory, sp_weights, lg_weights):
    d = tmp_path_factory.mktemp("w")
    torch.save(sp_weights, d / "sp.pth")
    torch.save(lg_weights, d / "lg.pth")
    return str(d / "sp.pth"), str(d / "lg.pth")


def test_superpoint_dense_stages(dev, sp_weights, weight_files):
    from gluefactory.models import get_model

    model = get_model("extractors.superpoint_open")({"weights": weight_files[0], "max_num_keypoints": 256}).to(dev).eval()
    img = torch.stack([synth.synthetic_image(40 + i, 136, 200) for i in range(2)])
    with torch.no_grad():
        ref = osp.forward(sp_weights, img, dict(max_num_keypoints=256), return_stages=True)
        heat, dense = model.dense_forward(img.to(dev))
    heat, dense = heat.cpu(), dense.cpu().permute(0, 3, 1, 2)
    err = (heat - ref["heatmap"]).abs().max().item() / ref["heatmap"].max().item()
    cos = (dense * ref["_stages"]["dense_descriptors"]).sum(1)
    print(f"heatmap max err / max = {err:.4f}; dense descriptor cosine min {cos.min().item():.5f} mean {cos.mean().item():.6f}")
    assert heat.shape == ref["heatmap"].shape
    assert err < 3e-2
    assert cos.min().item() > 0.998


def test_superpoint_teacher_forced_detection(dev, sp_weights, weight_files):
    """NMS + top-k + sampling on the ORACLE's heat map and dense descriptors: bit-exact keypoints."""
    from gluefactory.models import get_model

    model = get_model("extractors.superpoint_open")({"weights": weight_files[0], "max_num_keypoints": 300,
                                                      "detection_threshold": 0.0}).to(dev).eval()
    img = torch.stack([synth.synthetic_image(50 + i, 160, 192) for i in range(2)])
    with torch.no_grad():
        ref = osp.forward(sp_weights, img, dict(max_num_keypoints=300, detection_threshold=0.0), return_stages=True)
        kpts, scores, counts = model.detect(ref["heatmap"].to(dev))
        from glue_factory_b200 import ops

        desc = ops.sp_sample_descriptors(ref["_stages"]["dense_descriptors"].permute(0, 2, 3, 1).contiguous().to(dev), kpts)
    assert torch.equal(kpts.cpu(), ref["keypoints"])
    assert torch.equal(scores.cpu(), ref["keypoint_scores"])
    rel = (desc.cpu() - ref["descriptors"]).norm(dim=-1) / ref["descriptors"].norm(dim=-1)
    assert rel.max().item() < 1e-3


def _decisive(ref, th):
    """Keypoints of image 0 whose match decision in the REFERENCE is not a near-tie (see the module docstring)."""
    inner = ref["log_assignment"][:, :-1, :-1]
    top2 = inner.topk(2, dim=2).values
    best_col = inner.argmax(2)
    col_top2 = inner.topk(2, dim=1).values  # (B, 2, N)
    col_margin = (col_top2[:, 0] - col_top2[:, 1]).gather(1, best_col)
    ms = ref["matching_scores0"]
    near_th = (ms - th).abs() < 0.1 * th
    return ~(((top2[..., 0] - top2[..., 1]) < 0.05) | (col_margin < 0.05) | near_th)


def _lg_inputs(seed, B, M, N):
    g = np.random.default_rng(seed)
    size = np.tile(np.array([[640.0, 480.0]], dtype=np.float32), (B, 1))
    k0 = (g.random((B, M, 2)) * size[:, None]).astype(np.float32)
    k1 = (g.random((B, N, 2)) * size[:, None]).astype(np.float32)
    d0 = g.standard_normal((B, M, 256)).astype(np.float32)
    d1 = np.stack([d0[b][g.integers(0, M, N)] for b in range(B)]) + 0.35 * g.standard_normal((B, N, 256)).astype(np.float32)
    d0 /= np.linalg.norm(d0, axis=-1, keepdims=True)
    d1 /= np.linalg.norm(d1, axis=-1, keepdims=True)
    return [torch.from_numpy(x) for x in (k0, k1, d0, d1.astype(np.float32), size)]


@pytest.mark.parametrize("B,M,N", [(1, 128, 128), (2, 200, 150), (1, 512, 640)])
def test_lightglue_vs_oracle(dev, lg_weights, weight_files, B, M, N):
    from gluefactory.models import get_model

    k0, k1, d0, d1, size = _lg_inputs(B * 100 + M, B, M, N)
    model = get_model("matchers.lightglue")({"weights": weight_files[1], "filter_threshold": 0.1}).to(dev).eval()
    data = {"keypoints0": k0.to(dev), "keypoints1": k1.to(dev), "descriptors0": d0.to(dev), "descriptors1": d1.to(dev),
            "view0": {"image_size": size.to(dev)}, "view1": {"image_size": size.to(dev)}}
    with torch.no_grad():
        pred = model(data)
        ref = olg.forward(lg_weights, k0, k1, d0, d1, size, size, dict(filter_threshold=0.1))
    for key in ("ref_descriptors0", "ref_descriptors1"):
        rel = ((pred[key].cpu() - ref[key]).norm() / ref[key].norm()).item()
        print(key, "relative L2", rel)
        assert rel < 1e-2
    la, la_ref = pred["log_assignment"].cpu(), ref["log_assignment"]
    assert la.shape == la_ref.shape and la.dtype == torch.float32
    rel = ((la - la_ref).abs() / la_ref.abs().clamp_min(1e-3))
    print("log_assignment relative error: median", rel.median().item(), "p99", rel.flatten().kthvalue(int(rel.numel() * 0.99)).values.item())
    assert rel.median().item() < 3e-3
    assert rel.flatten().kthvalue(int(rel.numel() * 0.99)).values.item() < 1.5e-2
    agree = (pred["matches0"].cpu() == ref["matches0"]).float().mean().item()
    matched = ref["matches0"] >= 0
    agree_matched = (pred["matches0"].cpu()[matched] == ref["matches0"][matched]).float().mean().item()
    dec = _decisive(ref, 0.1)
    agree_dec = (pred["matches0"].cpu()[dec] == ref["matches0"][dec]).float().mean().item()
    print(f"match agreement {agree:.4f} (on {matched.sum().item()} reference matches: {agree_matched:.4f}; "
          f"on the {dec.float().mean().item():.3f} decisive share: {agree_dec:.4f})")
    assert agree >= 0.99 and agree_matched >= 0.99 and agree_dec >= 0.995
    assert pred["matches0"].dtype == torch.int64 and pred["prune0"].dtype == torch.float32


def test_pipeline_boat_vs_reference_golden(dev, golden_dir, weight_files):
    """Configuration 1: boat pair, 1024 keypoints, through TwoViewPipeline; compared with the
    vectors produced by the unmodified reference (tests/golden/make_golden.py)."""
    import cv2

    from gluefactory.models import get_model

    g = np.load(os.path.join(golden_dir, "boat_pipeline.npz"))
    pipe = get_model("two_view_pipeline")({
        "extractor": {"name": "extractors.superpoint_open", "max_num_keypoints": 1024, "nms_radius": 4,
                      "detection_threshold": 0.0, "weights": weight_files[0]},
        "matcher": {"name": "matchers.lightglue", "filter_threshold": 0.1, "weights": weight_files[1]},
    }).to(dev).eval()
    views = {}
    for i, name in enumerate(("boat1.png", "boat2.png")):
        im = cv2.imread(os.path.join(golden_dir, "assets", name), cv2.IMREAD_GRAYSCALE)
        views[f"view{i}"] = {"image": (torch.from_numpy(im)[None, None].float() / 255.0).to(dev),
                             "image_size": torch.tensor([[im.shape[1], im.shape[0]]]).float().to(dev)}
    with torch.no_grad():
        pred = pipe(views)
    assert pred["keypoints0"].shape == (1, 1024, 2) and pred["log_assignment"].shape == (1, 1025, 1025)
    for i in (0, 1):
        got = {tuple(p) for p in pred[f"keypoints{i}"][0].cpu().numpy().tolist()}
        want = {tuple(p) for p in g[f"keypoints{i}"][0].tolist()}
        print(f"view{i}: keypoint set overlap with the reference {len(got & want) / 1024:.4f}")
        assert len(got & want) / 1024 > 0.95  # bf16 conv stack moves borderline detections (achieved 0.968)
    # matches of the reference pipeline, found again by keypoint coordinates (a match is lost when either end moved)
    k0 = {tuple(p): i for i, p in enumerate(pred["keypoints0"][0].cpu().numpy().tolist())}
    k1 = [tuple(p) for p in pred["keypoints1"][0].cpu().numpy().tolist()]
    m_ours = pred["matches0"][0].cpu().tolist()
    hits = total = 0
    for i, j in enumerate(g["matches0"][0].tolist()):
        if j < 0:
            continue
        total += 1
        a = k0.get(tuple(g["keypoints0"][0][i].tolist()))
        hits += a is not None and m_ours[a] >= 0 and k1[m_ours[a]] == tuple(g["keypoints1"][0][j].tolist())
    print(f"reference matches reproduced end to end: {hits}/{total}")
    # the boat pair yields only 12 reference matches with the seeded random-init weights (achieved 9 of 12); the
    # benchmark-size figure (268 matches, 92.9 %) is asserted in test_gpu_parity_config.py
    assert total > 0 and hits / total >= 0.7


@pytest.mark.parametrize("name", ["lg_adaptive", "lg_prune_only", "lg_depth_only"])
def test_lightglue_adaptive_vs_reference_golden(dev, golden_dir, weight_files, name):
    """Configuration 3 semantics (early exit + point pruning) against vectors of the patched reference."""
    from gluefactory.models import get_model

    g = np.load(os.path.join(golden_dir, name + ".npz"))
    model = get_model("matchers.lightglue")({
        "weights": weight_files[1], "filter_threshold": 0.1, "depth_confidence": float(g["depth_confidence"]),
        "width_confidence": float(g["width_confidence"])}).to(dev).eval()
    size = torch.from_numpy(g["image_size"]).to(dev)
    data = {"keypoints0": torch.from_numpy(g["keypoints0"]).to(dev), "keypoints1": torch.from_numpy(g["keypoints1"]).to(dev),
            "descriptors0": torch.from_numpy(g["descriptors0"]).to(dev), "descriptors1": torch.from_numpy(g["descriptors1"]).to(dev),
            "view0": {"image_size": size}, "view1": {"image_size": size}}
    with torch.no_grad():
        pred = model(data)
    # bf16 GEMM operands move matchability / confidence logits by ~1e-2, so a keypoint sitting on a pruning
    # threshold may flip (expected: a few per 320 tokens over 8 layers; a flipped keypoint also changes the later
    # layers' context).  Stated tolerance: the per-keypoint prune counters agree on >= 97 % of the keypoints (achieved
    # 99.1 % .. 100 %) and the surviving set differs by <= 2 %, matches agree on >= 95 % (achieved 98.1 % .. 100 %); the compaction
    # kernels themselves are checked bit-exactly in test_gpu_bandwidth.py::test_prune_compaction_bit_exact.
    p0, p1 = pred["prune0"].cpu().numpy(), pred["prune1"].cpu().numpy()
    assert pred["prune0"].dtype == torch.from_numpy(g["prune0"]).dtype
    same = np.concatenate([(p0 == g["prune0"]).ravel(), (p1 == g["prune1"]).ravel()]).mean()
    kept_diff = abs(int(pred["log_assignment"].shape[1]) - g["log_assignment"].shape[1]) + abs(
        int(pred["log_assignment"].shape[2]) - g["log_assignment"].shape[2])
    agree = (pred["matches0"].cpu().numpy() == g["matches0"]).mean()
    print(name, "prune counters equal", same, "kept-count difference", kept_diff, "match agreement", agree,
          "log_assignment", tuple(pred["log_assignment"].shape), g["log_assignment"].shape)
    assert same >= 0.97 and kept_diff <= 0.02 * (p0.size + p1.size)
    assert p0.max() == g["prune0"].max()  # same exit layer
    assert agree >= 0.95
    if tuple(pred["log_assignment"].shape) == g["log_assignment"].shape and same == 1.0:
        rel = np.abs(pred["log_assignment"].cpu().numpy() - g["log_assignment"]) / np.maximum(np.abs(g["log_assignment"]), 1e-3)
        assert np.median(rel) < 3e-3
        rd = pred["ref_descriptors0"].cpu().numpy()
        assert np.linalg.norm(rd - g["ref_descriptors0"]) / np.linalg.norm(g["ref_descriptors0"]) < 1e-2
