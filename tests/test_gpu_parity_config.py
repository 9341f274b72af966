"""Floating-point parity AT THE BENCHMARKED CONFIGURATION (BASELINE.json configs[1]): 1024x1024 synthetic pairs,
2048 keypoints, nms_radius 3, detection_threshold -1, filter_threshold 0.1, batch 2, through TwoViewPipeline,
against the fp32 CPU oracle on the same pairs and weights.

Tolerances are stated per tensor next to each assertion; the achieved figures (max / p99 / median) are written to
gpurun_out/parity_config.json and copied into DESIGN.md section 2.1.  Three regimes:

  * final stage teacher-forced on the oracle's last-layer descriptors: fp32-grade arithmetic (two-term bf16 split
    GEMMs) => north_star's 1e-3 relative on log_assignment and >= 99.5 % match agreement are ASSERTED;
  * matcher teacher-forced on the oracle's keypoints / descriptors: bf16 GEMM operands through 9 layers with an fp32
    residual stream; bf16 cannot meet 1e-3 on the layer outputs (operand rounding 2^-9 per product), what it does
    meet is asserted;
  * fully end to end from the images (bf16 convolution stack moves borderline detections): reported and bounded.
"""
import json
import os

import pytest
import torch

from glue_factory_b200 import synth
from oracle import lightglue as olg
from oracle import sp_open as osp

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
H = W = 1024
K = 2048
PAIRS = 2
SP_CONF = dict(max_num_keypoints=K, detection_threshold=-1, nms_radius=3)


def _pct(x, q):
    x = x.flatten().float()
    k = max(1, min(x.numel(), int(round(q * x.numel()))))
    return x.kthvalue(k).values.item()


def _stats(x):
    return {"median": x.median().item(), "p99": _pct(x, 0.99), "max": x.max().item()}


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def setup(dev, tmp_path_factory, sp_weights, lg_weights):
    from gluefactory.models import get_model

    d = tmp_path_factory.mktemp("w")
    torch.save(sp_weights, d / "sp.pth")
    torch.save(lg_weights, d / "lg.pth")
    pipe = get_model("two_view_pipeline")({
        "extractor": {"name": "extractors.superpoint_open", "weights": str(d / "sp.pth"), **SP_CONF},
        "matcher": {"name": "matchers.lightglue", "weights": str(d / "lg.pth"), "filter_threshold": 0.1},
    }).to(dev).eval()
    im0, im1, _ = synth.synthetic_batch(0, PAIRS, H, W, 1234)  # the first pairs of bench.py's workload
    size = torch.tensor([[float(W), float(H)]] * PAIRS)
    torch.set_num_threads(os.cpu_count() or 1)
    with torch.no_grad():
        r0 = osp.forward(sp_weights, im0, SP_CONF, return_stages=True)
        r1 = osp.forward(sp_weights, im1, SP_CONF, return_stages=True)
        rl = olg.forward(lg_weights, r0["keypoints"], r1["keypoints"], r0["descriptors"], r1["descriptors"], size, size,
                         dict(filter_threshold=0.1), return_stages=True)
    report = {"config": "configs[1]: 2 pairs of 1024x1024, 2048 kpts, nms_radius 3, detection_threshold -1"}
    yield dict(pipe=pipe, im0=im0, im1=im1, size=size, r0=r0, r1=r1, rl=rl, report=report)
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        json.dump(report, open(os.path.join(out, "parity_config.json"), "w"), indent=1)
    print("PARITY_REPORT " + json.dumps(report))


def test_extractor_dense_stage(dev, setup):
    """A1-A4 at 1024x1024: heat map and dense descriptors of the bf16 convolution stack vs fp32."""
    s = setup
    with torch.no_grad():
        heat, dense = s["pipe"].extractor.dense_forward(s["im0"].to(dev))
    heat, dense = heat.cpu(), dense.cpu().permute(0, 3, 1, 2)
    ref_heat, ref_dense = s["r0"]["heatmap"], s["r0"]["_stages"]["dense_descriptors"]
    err = (heat - ref_heat).abs() / ref_heat.max()
    cos = (dense * ref_dense).sum(1)
    rel = (dense - ref_dense).norm(dim=1)  # both unit norm => relative L2
    s["report"]["heatmap_abs_err_over_max"] = _stats(err)
    s["report"]["dense_descriptor_rel_l2"] = _stats(rel)
    s["report"]["dense_descriptor_cosine_min"] = cos.min().item()
    # bf16 operands through ten convolutions (2^-9 rounding per product, detector logits scaled x8 by the weight
    # recipe): stated tolerance 6e-2 of the heat-map maximum at the worst pixel, 6e-3 at the 99th percentile, 5e-4
    # at the median (achieved 4.7e-2 / 4.3e-3 / 2.2e-4); dense descriptors cosine >= 0.998, relative L2 <= 6e-2
    # (achieved 0.9989 / 4.7e-2 worst cell, 2.2e-2 median).  bf16 cannot meet 1e-3 here.
    assert err.max().item() < 6e-2 and _pct(err, 0.99) < 6e-3 and err.median().item() < 5e-4
    assert cos.min().item() > 0.998 and rel.max().item() < 6e-2


def test_extractor_teacher_forced_detection_and_sampling(dev, setup):
    """A5-A9 at 1024x1024 / 2048 keypoints on the ORACLE's heat map and dense descriptors: keypoints and scores
    bit-exact (order included), sampled descriptors within 1e-3 relative (achieved ~1e-6)."""
    from glue_factory_b200 import ops

    s = setup
    ext = s["pipe"].extractor
    with torch.no_grad():
        kpts, scores, _ = ext.detect(s["r0"]["heatmap"].to(dev))
        dense = s["r0"]["_stages"]["dense_descriptors"].permute(0, 2, 3, 1).contiguous().to(dev)
        desc = ops.sp_sample_descriptors(dense, kpts)
    assert torch.equal(kpts.cpu(), s["r0"]["keypoints"])
    assert torch.equal(scores.cpu(), s["r0"]["keypoint_scores"])
    rel = (desc.cpu() - s["r0"]["descriptors"]).norm(dim=-1) / s["r0"]["descriptors"].norm(dim=-1)
    s["report"]["sampled_descriptor_rel_l2_teacher_forced"] = _stats(rel)
    assert rel.max().item() < 1e-3


@pytest.mark.parametrize("layout", ["narrow", "wide"])
def test_matcher_teacher_forced(dev, setup, layout, monkeypatch):
    """L1-L12 at 2 x (2048 + 2048) tokens on the oracle's keypoints and descriptors, with either stream layout of the
    attention kernel (two pairs alone would take the narrow one; the benchmark's 32 pairs take the wide one)."""
    monkeypatch.setenv("B200GLUE_ATT_LAYOUT", layout)
    s = setup
    r0, r1, rl = s["r0"], s["r1"], s["rl"]
    matcher = s["pipe"].matcher
    size = s["size"].to(dev)
    matcher.layer_tap = []
    try:
        with torch.no_grad():
            pred = matcher({"keypoints0": r0["keypoints"].to(dev), "keypoints1": r1["keypoints"].to(dev),
                            "descriptors0": r0["descriptors"].to(dev), "descriptors1": r1["descriptors"].to(dev),
                            "view0": {"image_size": size}, "view1": {"image_size": size}})
        taps = [t.cpu() for t in matcher.layer_tap]
    finally:
        matcher.layer_tap = None
    per_layer = []
    for i, x in enumerate(taps):
        d0 = x[: PAIRS * K].view(PAIRS, K, 256)
        ref0 = rl["_stages"]["layers"][i][0]
        per_layer.append(((d0 - ref0).norm(dim=-1) / ref0.norm(dim=-1)).max().item())
    suffix = "" if layout == "wide" else "_narrow_attention"
    s["report"]["matcher_layer_descriptor_rel_l2_max" + suffix] = per_layer
    la, la_ref = pred["log_assignment"].cpu(), rl["log_assignment"]
    rel = (la - la_ref).abs() / la_ref.abs().clamp_min(1e-3)
    agree = (pred["matches0"].cpu() == rl["matches0"]).float().mean().item()
    s["report"]["matcher_teacher_forced" + suffix] = {"log_assignment_rel_err": _stats(rel), "match_agreement": agree,
                                             "reference_matches": int((rl["matches0"] >= 0).sum())}
    # bf16 operands, fp32 residual stream: relative L2 of the WORST token after every layer <= 1e-2 (SURVEY 8d's
    # figure "by layer 9"; achieved 2.8e-3 after layer 1 growing to 7.1e-3 after layer 9).  bf16 cannot meet 1e-3
    # here: one product already carries 2^-9 = 2e-3 of operand rounding.
    assert max(per_layer) < 1e-2
    # log_assignment through 9 bf16 layers: median <= 2e-3, p99 <= 1e-2 relative (achieved 9e-4 / 4.2e-3; the
    # maximum, 4e-2, sits on entries near zero); north_star's match-set agreement >= 99.5 % holds (achieved 99.95 %)
    assert rel.median().item() < 2e-3 and _pct(rel, 0.99) < 1e-2
    assert agree >= 0.995


def test_final_stage_teacher_forced_meets_north_star(dev, setup):
    """L10/L11 on the oracle's layer-9 descriptors: the fp32-grade projection + similarity (two-term bf16 split
    GEMMs) must meet north_star: log_assignment within 1e-3 relative, match agreement >= 99.5 %."""
    s = setup
    rl = s["rl"]
    matcher = s["pipe"].matcher
    with torch.no_grad():
        la, m0, m1, ms0, _ = matcher.assign_from_descriptors(rl["ref_descriptors0"][:, 0].to(dev),
                                                             rl["ref_descriptors1"][:, 0].to(dev))
    rel = (la.cpu() - rl["log_assignment"]).abs() / rl["log_assignment"].abs().clamp_min(1e-3)
    agree0 = (m0.cpu() == rl["matches0"]).float().mean().item()
    agree1 = (m1.cpu() == rl["matches1"]).float().mean().item()
    s["report"]["final_stage_teacher_forced"] = {"log_assignment_rel_err": _stats(rel), "match_agreement0": agree0,
                                                 "match_agreement1": agree1}
    assert rel.max().item() < 1e-3
    assert agree0 >= 0.995 and agree1 >= 0.995
    matched = rl["matches0"] >= 0
    got, want = ms0.cpu()[matched], rl["matching_scores0"][matched]
    assert ((got - want).abs() / want).max().item() < 1e-3


def test_end_to_end_from_images(dev, setup):
    """Whole pipeline from the images: keypoints move where bf16 heat-map errors reorder near-ties, so matches are
    compared by coordinates.  Reported; bounded loosely (this is the regime bf16 cannot pin)."""
    s = setup
    r0, r1, rl = s["r0"], s["r1"], s["rl"]
    size = s["size"].to(dev)
    with torch.no_grad():
        pred = s["pipe"]({"view0": {"image": s["im0"].to(dev), "image_size": size},
                          "view1": {"image": s["im1"].to(dev), "image_size": size}})
    assert pred["keypoints0"].shape == (PAIRS, K, 2) and pred["log_assignment"].shape == (PAIRS, K + 1, K + 1)
    overlap, hits, total, ours_total = [], 0, 0, 0
    for b in range(PAIRS):
        ours0 = {tuple(k): i for i, k in enumerate(pred["keypoints0"][b].cpu().tolist())}
        ours1 = [tuple(k) for k in pred["keypoints1"][b].cpu().tolist()]
        ref0 = [tuple(k) for k in r0["keypoints"][b].tolist()]
        ref1 = [tuple(k) for k in r1["keypoints"][b].tolist()]
        overlap.append(len(set(ours0) & set(ref0)) / K)
        m_ours, m_ref = pred["matches0"][b].cpu().tolist(), rl["matches0"][b].tolist()
        ours_total += sum(1 for j in m_ours if j >= 0)
        for i, j in enumerate(m_ref):
            if j < 0:
                continue
            total += 1
            k = ours0.get(ref0[i])
            if k is not None and m_ours[k] >= 0 and ours1[m_ours[k]] == ref1[j]:
                hits += 1
    s["report"]["end_to_end"] = {"keypoint_set_overlap": sum(overlap) / PAIRS, "reference_matches": total,
                                 "reference_matches_reproduced": hits / max(1, total), "our_matches": ours_total}
    # achieved: 97.0 % of the reference's keypoints detected, 92.9 % of its matches reproduced with both end points
    # identical (a reference match is lost as soon as either of its keypoints moved)
    assert min(overlap) > 0.95
    assert hits / max(1, total) > 0.88
