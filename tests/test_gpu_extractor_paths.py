"""SuperPoint-open output assembly paths that the benchmark configuration does not
exercise: data-dependent keypoint counts, `force_num_keypoints` padding, `dense_outputs`, `max_num_keypoints=None`.
(reference: gluefactory/models/extractors/superpoint_open.py:154-217, models/utils/misc.py:19-70)

Real keypoints are checked bit-exactly against the oracle's NMS / border / selection run on the kernels' own heat
map; padded entries follow the reference's contract (score 0, coordinates inside the per-axis extent of the real
keypoints; the values come from the torch RNG and are not reproducible across implementations).
"""
import pytest
import torch

from glue_factory_b200 import synth
from oracle import sp_open as osp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def weights(tmp_path_factory, sp_weights):
    d = tmp_path_factory.mktemp("w")
    torch.save(sp_weights, d / "sp.pth")
    return str(d / "sp.pth")


def _model(weights, dev, **conf):
    from gluefactory.models import get_model

    return get_model("extractors.superpoint_open")({"weights": weights, **conf}).to(dev).eval()


def _oracle_detect(heat, conf):
    """Per-image keypoints of the oracle on a given heat map (no stacking, counts may differ)."""
    bordered = osp.remove_borders(osp.nms(heat, conf["nms_radius"]), conf.get("remove_borders", 4))
    return [osp.select_keypoints(bordered[i], conf["detection_threshold"], conf["max_num_keypoints"])
            for i in range(heat.shape[0])]


IMAGES = [(60, 160, 192), (61, 160, 192), (62, 160, 192)]


def _batch():
    imgs = [synth.synthetic_image(s, h, w) for s, h, w in IMAGES]
    return torch.stack(imgs)


@pytest.mark.parametrize("k", [1100, 1500])  # the three images hold ~1080 / ~1130 / ~1085 candidates
def test_force_num_keypoints_pads_like_the_reference(dev, weights, k):
    """configs/train+eval/superpoint-open+lightglue_megadepth.yaml:28 sets force_num_keypoints: True."""
    conf = dict(max_num_keypoints=k, detection_threshold=0.01, nms_radius=3, force_num_keypoints=True)
    model = _model(weights, dev, **conf)
    img = _batch()
    torch.manual_seed(0)
    with torch.no_grad():
        pred = model({"image": img.to(dev)})
    B = img.shape[0]
    assert pred["keypoints"].shape == (B, k, 2) and pred["keypoint_scores"].shape == (B, k)
    assert pred["descriptors"].shape == (B, k, 256)
    want = _oracle_detect(pred["heatmap"].cpu(), conf)
    counts = [w[0].shape[0] for w in want]
    assert len(set(counts)) > 1 and min(counts) < k, counts  # the batch really has differing counts
    for b in range(B):
        xy, sc, _ = want[b]
        n = xy.shape[0]
        got_xy, got_sc = pred["keypoints"][b].cpu(), pred["keypoint_scores"][b].cpu()
        assert torch.equal(got_xy[:n], xy + 0.5) and torch.equal(got_sc[:n], sc)
        if n < k:
            assert (got_sc[n:] == 0).all()
            pad = got_xy[n:] - 0.5
            lo, hi = xy.min(0).values, xy.max(0).values
            assert (pad >= lo).all() and (pad <= hi).all()
    norms = pred["descriptors"].norm(dim=-1)
    assert (norms - 1).abs().max().item() < 1e-4


def test_unequal_counts_raise_without_force(dev, weights):
    """torch.stack of per-image keypoint lists fails in the reference (:196)."""
    model = _model(weights, dev, max_num_keypoints=1500, detection_threshold=0.01, nms_radius=3)
    with torch.no_grad(), pytest.raises(RuntimeError, match="equal size"):
        model({"image": _batch().to(dev)})


def test_fewer_candidates_than_k_are_returned_in_row_major_order(dev, weights):
    """B = 1, fewer candidates than max_num_keypoints: no top-k call in the reference (:52-56), row-major order."""
    conf = dict(max_num_keypoints=4000, detection_threshold=0.01, nms_radius=3)
    model = _model(weights, dev, **conf)
    img = _batch()[1:2]
    with torch.no_grad():
        pred = model({"image": img.to(dev)})
    (xy, sc, idx), = _oracle_detect(pred["heatmap"].cpu(), conf)
    assert 0 < xy.shape[0] < 4000
    assert torch.equal(pred["keypoints"][0].cpu(), xy + 0.5) and torch.equal(pred["keypoint_scores"][0].cpu(), sc)
    assert (idx[1:] > idx[:-1]).all()


def test_no_keypoint_cap_and_dense_outputs(dev, weights):
    """max_num_keypoints=None (every candidate, :155-160) and dense_outputs (:213-214)."""
    conf = dict(max_num_keypoints=None, detection_threshold=0.02, nms_radius=4, dense_outputs=True)
    model = _model(weights, dev, **conf)
    img = _batch()[:1]
    with torch.no_grad():
        pred = model({"image": img.to(dev)})
        heat, dense = model.dense_forward(img.to(dev))
    (xy, sc, _), = _oracle_detect(pred["heatmap"].cpu(), conf)
    assert torch.equal(pred["keypoints"][0].cpu(), xy + 0.5) and torch.equal(pred["keypoint_scores"][0].cpu(), sc)
    assert pred["dense_descriptors"].shape == (1, 256, 20, 24)
    assert torch.equal(pred["dense_descriptors"], dense.permute(0, 3, 1, 2))
    assert torch.equal(pred["heatmap"], heat)


def test_exactly_k_maxima_with_negative_threshold_sort_by_score(dev):
    """with detection_threshold < 0 the reference's candidate set is every interior pixel (> k), so its
    top-k returns score-descending order even when exactly k positive maxima exist."""
    from glue_factory_b200 import ops

    g = torch.Generator().manual_seed(5)
    H, W, k = 96, 128, 48
    scores = torch.zeros(1, H, W)
    ys = torch.arange(8, 8 + 6 * 12, 12)
    xs = torch.arange(8, 8 + 8 * 14, 14)
    vals = torch.rand(48, generator=g) * 0.5 + 0.1
    i = 0
    for y in ys:
        for x in xs:
            scores[0, y, x] = vals[i]
            i += 1
    assert i == k
    _, kpts, ksc, idx, counts = ops.sp_detect_fast(scores.to(dev), 3, 4, -1.0, k)
    bordered = osp.remove_borders(osp.nms(scores, 3), 4)
    xy, sc, lin = osp.select_keypoints(bordered[0], -1.0, k)
    assert counts.tolist() == [k]
    assert torch.equal(ksc[0].cpu(), sc) and torch.equal(kpts[0].cpu(), xy + 0.5)
    assert torch.equal(idx[0].cpu().long(), lin)
    assert (sc[1:] <= sc[:-1]).all()
